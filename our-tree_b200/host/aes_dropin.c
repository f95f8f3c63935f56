/*
 * aes_dropin.c — the reference's aes-modes/aes.h entry points on top of the
 * B200 engine.  Key schedules run here on the host in C (they
 * are O(240 bytes) per key and must fill aes_context identically); every
 * cipher round runs on the GPU through bc_*_host.
 */
#define _GNU_SOURCE            /* process_vm_readv */
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/uio.h>
#include <unistd.h>

#include "aes.h"
#include "bulkcipher.h"
#include "../csrc/bc_internal.h"
#include "../csrc/bc_sbox.inc"

static const uint8_t SBOX[256] = { BC_SBOX_INIT };

static uint8_t xtime(uint8_t v) { return (uint8_t)((v << 1) ^ ((v & 0x80) ? 0x1B : 0)); }

static uint32_t sub_word(uint32_t w)
{
    return (uint32_t)SBOX[w & 0xFF] | ((uint32_t)SBOX[(w >> 8) & 0xFF] << 8) |
           ((uint32_t)SBOX[(w >> 16) & 0xFF] << 16) | ((uint32_t)SBOX[w >> 24] << 24);
}

uint32_t bc_aes_sub_word(uint32_t w) { return sub_word(w); }    /* for aesni_dropin.c's KEY_256_ASSIST_2 */

/* InvMixColumns of one little-endian column word: what RT0..3[FSb[.]] yields at
 * aes-modes/aes.c:580-589. */
static uint32_t inv_mix_column(uint32_t w)
{
    uint8_t b[4] = { (uint8_t)w, (uint8_t)(w >> 8), (uint8_t)(w >> 16), (uint8_t)(w >> 24) }, o[4];
    for (int i = 0; i < 4; i++) {
        uint8_t acc = 0;
        static const uint8_t coef[4] = { 0x0E, 0x0B, 0x0D, 0x09 };
        for (int j = 0; j < 4; j++) {
            uint8_t v = b[(i + j) & 3], c = coef[j], p = 0;
            while (c) { if (c & 1) p ^= v; v = xtime(v); c >>= 1; }
            acc ^= p;
        }
        o[i] = acc;
    }
    return (uint32_t)o[0] | ((uint32_t)o[1] << 8) | ((uint32_t)o[2] << 16) | ((uint32_t)o[3] << 24);
}

/* FIPS-197 expansion on little-endian words (aes.c:469-539); returns nr or 0.  Declared in bulkcipher.h. */
int bc_aes_expand_key(const unsigned char *key, unsigned int keybits, uint32_t rk[60])
{
    if (keybits != 128 && keybits != 192 && keybits != 256) return 0;
    const int nk = (int)keybits / 32, nr = nk + 6, total = 4 * (nr + 1);
    for (int i = 0; i < nk; i++)
        rk[i] = (uint32_t)key[4 * i] | ((uint32_t)key[4 * i + 1] << 8) | ((uint32_t)key[4 * i + 2] << 16) |
                ((uint32_t)key[4 * i + 3] << 24);
    uint8_t rcon = 1;
    for (int i = nk; i < total; i++) {
        uint32_t t = rk[i - 1];
        if (i % nk == 0) { t = sub_word((t >> 8) | (t << 24)) ^ rcon; rcon = xtime(rcon); }
        else if (nk == 8 && i % nk == 4) t = sub_word(t);
        rk[i] = rk[i - nk] ^ t;
    }
    return nr;
}

/* the equivalent-inverse schedule of aes.c:563-594 from an encryption schedule */
void bc_aes_invert_key(const uint32_t enc[60], int nr, uint32_t dec[60])
{
    for (int r = 0; r <= nr; r++)
        for (int c = 0; c < 4; c++) {
            uint32_t w = enc[4 * (nr - r) + c];
            dec[4 * r + c] = (r == 0 || r == nr) ? w : inv_mix_column(w);
        }
}

static void ctx_store(aes_context *ctx, const uint32_t rk[60], int nr)
{
    ctx->nr = nr;
    ctx->rk = ctx->buf;                                   /* aes.c:466 */
    for (int i = 0; i < 4 * (nr + 1); i++) ctx->buf[i] = (unsigned long)rk[i];
}

static int ctx_load(const aes_context *ctx, uint32_t rk[60])
{
    if (!ctx || !ctx->rk || (ctx->nr != 10 && ctx->nr != 12 && ctx->nr != 14)) {
        bc_set_error("aes_context holds no key schedule (call aes_setkey_enc/dec first)");
        return BC_ERR_BAD_ARG;
    }
    for (int i = 0; i < 4 * (ctx->nr + 1); i++) rk[i] = (uint32_t)ctx->rk[i];
    return BC_OK;
}

int aes_setkey_enc(aes_context *ctx, const unsigned char *key, unsigned int keysize)
{
    uint32_t rk[60];
    int nr = bc_aes_expand_key(key, keysize, rk);
    if (!nr) return POLARSSL_ERR_AES_INVALID_KEY_LENGTH;  /* aes.c:455-461 */
    ctx_store(ctx, rk, nr);
    return 0;
}

int aes_setkey_dec(aes_context *ctx, const unsigned char *key, unsigned int keysize)
{
    uint32_t enc[60], dec[60];
    int nr = bc_aes_expand_key(key, keysize, enc);
    if (!nr) return POLARSSL_ERR_AES_INVALID_KEY_LENGTH;  /* aes.c:555-561 */
    bc_aes_invert_key(enc, nr, dec);
    ctx_store(ctx, dec, nr);
    memset(enc, 0, sizeof enc);                           /* aes.c:596 */
    return 0;
}

/* ---- aes_crypt_ecb: one block per call, made bearable for loop callers -----------------------
 *
 * The reference's "Plain ECB" driver walks a buffer 16 bytes at a time (aes-modes/test.c:37-41).
 * One GPU round trip per block is ~30 us, i.e. well under 1 MB/s.  When a thread's calls walk
 * consecutive addresses, the library therefore reads AHEAD: it copies the next window of the
 * caller's input (4 KiB, growing to 4 MiB while the windows keep being used up), runs the whole
 * window through the GPU in one call and keeps (input copy, output) per thread.  A later call is
 * served from that window only if its 16 input bytes EQUAL the copy the output was computed from
 * and the key schedule and mode are the same — ECB is a pure function of (key, mode, block), so the
 * answer is exact whatever happened to the caller's memory in between; anything else is a miss and
 * goes to the GPU.  Nothing is ever computed on the host.
 *
 * Reading ahead must not fault on memory the caller never promised: the copy is made with
 * process_vm_readv on our own pid, one iovec element per page, which stops at the first
 * unreadable page and reports how far it got.  Where that call is not permitted the read-ahead
 * switches itself off and every call is a GPU round trip as before.
 * BULKCIPHER_ECB_READAHEAD=0 disables it. */
#define ECB_AHEAD_MIN ((size_t)4096)
#define ECB_AHEAD_MAX ((size_t)4 << 20)

typedef struct {
    uint32_t rk[60];
    int nr, mode;
    unsigned char *in, *out;        /* window: copy of the caller's input, and its ECB image */
    size_t cap;
    size_t n_blocks, next;          /* blocks in the window, index of the one expected next */
    const unsigned char *last_in;   /* input pointer of the previous call (run detector) */
    unsigned run;
    size_t win;
    int off;
    uint64_t hits, misses;          /* calls served from the window / sent to the GPU (bc_internal_ecb_ahead_stats) */
} ecb_ahead;

static pthread_key_t g_ecb_key;
static pthread_once_t g_ecb_once = PTHREAD_ONCE_INIT;

static void ecb_ahead_free(void *p)
{
    ecb_ahead *a = (ecb_ahead *)p;
    if (!a) return;
    if (a->in)  { memset(a->in, 0, a->cap);  free(a->in); }
    if (a->out) { memset(a->out, 0, a->cap); free(a->out); }
    memset(a, 0, sizeof *a);
    free(a);
}

static void ecb_key_make(void) { pthread_key_create(&g_ecb_key, ecb_ahead_free); }

static ecb_ahead *ecb_state(void)
{
    pthread_once(&g_ecb_once, ecb_key_make);
    ecb_ahead *a = (ecb_ahead *)pthread_getspecific(g_ecb_key);
    if (!a) {
        a = (ecb_ahead *)calloc(1, sizeof *a);
        if (!a) return NULL;
        const char *e = getenv("BULKCIPHER_ECB_READAHEAD");
        if (e && *e == '0') a->off = 1;
        pthread_setspecific(g_ecb_key, a);
    }
    return a;
}

/* Copy up to `want` bytes starting at `src` without ever faulting; returns the bytes copied. */
static size_t read_ahead(unsigned char *dst, const unsigned char *src, size_t want)
{
    struct iovec local = { dst, want }, remote[1024];
    const size_t page = 4096;
    size_t n = 0, off = 0;
    while (off < want && n < 1024) {
        size_t len = page - (((uintptr_t)src + off) & (page - 1));
        if (len > want - off) len = want - off;
        remote[n].iov_base = (void *)(src + off);
        remote[n].iov_len = len;
        off += len; n++;
    }
    local.iov_len = off;
    const ssize_t got = process_vm_readv(getpid(), &local, 1, remote, n, 0);
    return got > 0 ? (size_t)got : 0;
}

/* test hook (tests/c/host_san_test.c): the fault-free copy on its own */
size_t bc_internal_read_ahead(unsigned char *dst, const unsigned char *src, size_t want) { return read_ahead(dst, src, want); }

int aes_crypt_ecb(aes_context *ctx, int mode, const unsigned char input[16], unsigned char output[16])
{
    uint32_t rk[60];
    int rc = ctx_load(ctx, rk);
    if (rc) return rc;
    const int nr = ctx->nr, m = mode == AES_DECRYPT ? BC_DECRYPT : BC_ENCRYPT;
    ecb_ahead *a = ecb_state();
    if (!a || a->off) return bc_aes_ecb_host(rk, nr, m, input, output, 16);

    if (a->next < a->n_blocks && a->nr == nr && a->mode == m &&
        memcmp(a->rk, rk, 16 * (size_t)(nr + 1)) == 0 && memcmp(a->in + 16 * a->next, input, 16) == 0) {
        memcpy(output, a->out + 16 * a->next, 16);        /* same key, same mode, same 16 input bytes: same block */
        a->next++;
        a->last_in = input;
        a->hits++;
        return 0;
    }
    /* miss: is this thread walking a buffer? */
    a->misses++;
    const int used_up = a->n_blocks && a->next == a->n_blocks;
    a->run = (a->last_in && input == a->last_in + 16) ? a->run + 1 : 0;
    a->last_in = input;
    a->n_blocks = a->next = 0;
    if (a->run >= 2) {
        a->win = used_up && a->win ? (a->win * 4 < ECB_AHEAD_MAX ? a->win * 4 : ECB_AHEAD_MAX) : ECB_AHEAD_MIN;
        if (a->cap < a->win) {
            unsigned char *ni = (unsigned char *)realloc(a->in, a->win), *no = ni ? (unsigned char *)realloc(a->out, a->win) : NULL;
            if (ni) a->in = ni;
            if (no) a->out = no;
            if (ni && no) a->cap = a->win; else a->win = a->cap;
        }
        size_t got = a->win >= 16 ? read_ahead(a->in, input, a->win) : 0;
        if (got == 0 && a->win >= 16) a->off = 1;         /* process_vm_readv not available here: stop trying */
        got -= got % 16;
        if (got >= 32 && memcmp(a->in, input, 16) == 0) {
            rc = bc_aes_ecb_host(rk, nr, m, a->in, a->out, got);
            if (rc) return rc;
            memcpy(a->rk, rk, sizeof rk);
            a->nr = nr; a->mode = m;
            a->n_blocks = got / 16; a->next = 1;
            memcpy(output, a->out, 16);
            return 0;
        }
    }
    return bc_aes_ecb_host(rk, nr, m, input, output, 16);
}

/* test hook: this thread's aes_crypt_ecb calls served from the read-ahead window / sent to the GPU */
void bc_internal_ecb_ahead_stats(uint64_t out[2])
{
    ecb_ahead *a = ecb_state();
    out[0] = a ? a->hits : 0;
    out[1] = a ? a->misses : 0;
}

int aes_crypt_cbc(aes_context *ctx, int mode, size_t length, unsigned char iv[16],
                  const unsigned char *input, unsigned char *output)
{
    if (length % 16) return POLARSSL_ERR_AES_INVALID_INPUT_LENGTH;   /* aes.c:767-768 */
    if (length == 0) return 0;                                        /* iv unchanged */
    uint32_t rk[60];
    int rc = ctx_load(ctx, rk);
    if (rc) return rc;
    if (mode == AES_DECRYPT) {
        unsigned char last[16];
        memcpy(last, input + length - 16, 16);            /* before an in-place overwrite (aes.c:786,792) */
        rc = bc_aes_cbc_decrypt_host(rk, ctx->nr, iv, input, output, length);
        if (rc == 0) memcpy(iv, last, 16);
        return rc;
    }
    /* one message, one serial chain; iv <- last ciphertext block (aes.c:807) */
    return bc_aes_cbc_encrypt_batch_host(rk, ctx->nr, 1, length, iv, iv, input, output);
}

/* aes.c:822-862.  The bytes that finish a block begun by an earlier call and the
 * bytes of a trailing partial block are plain XORs with a keystream block the GPU
 * produced (E(iv) in place, aes.c:835,849); all whole blocks run as one GPU call. */
int aes_crypt_cfb128(aes_context *ctx, int mode, size_t length, int *iv_off, unsigned char iv[16],
                     const unsigned char *input, unsigned char *output)
{
    if (!iv_off || !iv) { bc_set_error("aes_crypt_cfb128: NULL iv/iv_off"); return BC_ERR_BAD_ARG; }
    uint32_t rk[60];
    int rc = ctx_load(ctx, rk);
    if (rc) return rc;
    int n = *iv_off & 15;
    /* 1. finish the keystream block the caller carried in */
    while (n != 0 && length > 0) {
        if (mode == AES_DECRYPT) { unsigned char c = *input++; *output++ = (unsigned char)(c ^ iv[n]); iv[n] = c; }
        else iv[n] = *output++ = (unsigned char)(iv[n] ^ *input++);
        n = (n + 1) & 15;
        length--;
    }
    /* 2. whole blocks */
    const size_t whole = length & ~(size_t)15;
    if (whole) {
        if (mode == AES_DECRYPT) {
            unsigned char last[16];
            memcpy(last, input + whole - 16, 16);                 /* before an in-place overwrite */
            if ((rc = bc_aes_cfb128_decrypt_host(rk, ctx->nr, iv, input, output, whole))) return rc;
            memcpy(iv, last, 16);
        } else if ((rc = bc_aes_cfb128_encrypt_batch_host(rk, ctx->nr, 1, whole, iv, iv, input, output))) {
            return rc;
        }
        input += whole; output += whole; length -= whole;
    }
    /* 3. a trailing partial block: iv <- E(iv), then byte-wise */
    if (length) {
        if ((rc = bc_aes_ecb_host(rk, ctx->nr, BC_ENCRYPT, iv, iv, 16))) return rc;
        while (length--) {
            if (mode == AES_DECRYPT) { unsigned char c = *input++; *output++ = (unsigned char)(c ^ iv[n]); iv[n] = c; }
            else iv[n] = *output++ = (unsigned char)(iv[n] ^ *input++);
            n = (n + 1) & 15;
        }
    }
    *iv_off = n;
    return 0;
}

int aes_crypt_ctr(aes_context *ctx, int length, int *nc_off, unsigned char nonce_counter[16],
                  unsigned char stream_block[16], const unsigned char *input, unsigned char *output)
{
    if (length < 0 || !nc_off) { bc_set_error("aes_crypt_ctr: bad length/nc_off"); return BC_ERR_BAD_ARG; }
    uint32_t rk[60];
    int rc = ctx_load(ctx, rk);
    if (rc) return rc;
    int n = *nc_off & 15;
    size_t len = (size_t)length;
    /* finish the keystream block the caller carried in (aes.c:891-894 with n != 0):
     * a plain XOR with bytes the previous call already produced */
    while (n != 0 && len > 0) {
        *output++ = (unsigned char)(*input++ ^ stream_block[n]);
        n = (n + 1) & 15;
        len--;
    }
    if (len > 0) {
        rc = bc_aes_ctr_host(rk, ctx->nr, nonce_counter, 128, 0, input, output, len, stream_block);
        if (rc) return rc;
        /* the counter moves once per block STARTED (aes.c:881-890) */
        size_t started = (len + 15) / 16;
        for (int i = 15; i >= 0 && started; i--) {
            size_t sum = (size_t)nonce_counter[i] + (started & 0xFF);
            nonce_counter[i] = (unsigned char)sum;
            started = (started >> 8) + (sum >> 8);
        }
        n = (int)(len & 15);
    }
    *nc_off = n;
    return 0;
}

/* ------------------------------------------------------------- self test */

static int hexeq(const unsigned char *a, const char *hex, size_t n)
{
    for (size_t i = 0; i < n; i++) {
        unsigned v;
        sscanf(hex + 2 * i, "%2x", &v);
        if (a[i] != (unsigned char)v) return 0;
    }
    return 1;
}

static void unhex(unsigned char *o, const char *hex, size_t n)
{
    for (size_t i = 0; i < n; i++) { unsigned v; sscanf(hex + 2 * i, "%2x", &v); o[i] = (unsigned char)v; }
}

/* Same checks as aes_self_test (aes.c:1084-1330): Rijndael Monte-Carlo ECB and
 * CBC (zero key, 10 000 chained blocks; expected blocks aes.c:912-950) and the
 * RFC 3686 CTR vectors (aes.c:1025-1078), every block through the GPU. */
int aes_self_test(int verbose)
{
    static const char *ecb_dec[3] = { "44416ac2d1f53c583303917e6be9ebe0", "48e31e9e256718f29229319c19f15ba4", "058ccffdbbcb382d1f6f56585d8a4ade" };
    static const char *ecb_enc[3] = { "c34c052cc0da8d73451afe5f03be297f", "f3f6752ae8d7831138f041560631b114", "8b79eecc93a0ee5dff30b4ea21636da4" };
    static const char *cbc_dec[3] = { "faca37e0b0c85373df706e73f7c9af86", "5df678dd17ba4e75b61768c6adef7c7b", "4804e1818fe6297519a3e88c57310413" };
    static const char *cbc_enc[3] = { "8a05fc5e095af4848a08d328d3688e3d", "7bd966d53ad8c1bb85d2adfae87bb104", "fe3c53653e2f45b56fcd88b2cc898ff0" };
    static const char *ctr_key[3] = { "ae6852f8121067cc4bf7a5765577f39e", "7e24067817fae0d743d6ce1f32539163", "7691be035e5020a8ac6e618529f9a0dc" };
    static const char *ctr_nc[3]  = { "00000030000000000000000000000001", "006cb6dbc0543b59da48d90b00000001", "00e0017b27777f3f4a1786f000000001" };
    static const char *ctr_pt[3]  = { "53696e676c6520626c6f636b206d7367",
                                      "000102030405060708090a0b0c0d0e0f101112131415161718191a1b1c1d1e1f",
                                      "000102030405060708090a0b0c0d0e0f101112131415161718191a1b1c1d1e1f20212223" };
    static const char *ctr_ct[3]  = { "e4095d4fb7a7b3792d6175a3261311b8",
                                      "5104a106168a72d9790d41ee8edad388eb2e1efc46da57c8fce630df9141be28",
                                      "c1cf48a89f2ffdd9cf4652e9efdb72d74540a42bde6d7836d59a5ceaaef3105325b2072f" };
    static const int ctr_len[3] = { 16, 32, 36 };
    unsigned char key[32], buf[64], iv[16], prv[16], tmp[16];
    aes_context ctx;
    memset(key, 0, sizeof key);

    for (int i = 0; i < 6; i++) {
        const int u = i >> 1, v = i & 1;
        if (verbose) printf("  AES-ECB-%3d (%s): ", 128 + u * 64, v == AES_DECRYPT ? "dec" : "enc");
        memset(buf, 0, 16);
        if (v == AES_DECRYPT) aes_setkey_dec(&ctx, key, 128 + u * 64); else aes_setkey_enc(&ctx, key, 128 + u * 64);
        for (int j = 0; j < 10000; j++) if (aes_crypt_ecb(&ctx, v, buf, buf)) goto fail;
        if (!hexeq(buf, v == AES_DECRYPT ? ecb_dec[u] : ecb_enc[u], 16)) goto fail;
        if (verbose) printf("passed\n");
    }
    for (int i = 0; i < 6; i++) {
        const int u = i >> 1, v = i & 1;
        if (verbose) printf("  AES-CBC-%3d (%s): ", 128 + u * 64, v == AES_DECRYPT ? "dec" : "enc");
        memset(iv, 0, 16); memset(prv, 0, 16); memset(buf, 0, 16);
        if (v == AES_DECRYPT) {
            aes_setkey_dec(&ctx, key, 128 + u * 64);
            for (int j = 0; j < 10000; j++) if (aes_crypt_cbc(&ctx, v, 16, iv, buf, buf)) goto fail;
            if (!hexeq(buf, cbc_dec[u], 16)) goto fail;
        } else {
            aes_setkey_enc(&ctx, key, 128 + u * 64);
            for (int j = 0; j < 10000; j++) {
                if (aes_crypt_cbc(&ctx, v, 16, iv, buf, buf)) goto fail;
                memcpy(tmp, prv, 16); memcpy(prv, buf, 16); memcpy(buf, tmp, 16);
            }
            if (!hexeq(prv, cbc_enc[u], 16)) goto fail;
        }
        if (verbose) printf("passed\n");
    }
    for (int i = 0; i < 6; i++) {
        const int u = i >> 1, v = i & 1;
        unsigned char nc[16], sb[16];
        int off = 0, len = ctr_len[u];
        if (verbose) printf("  AES-CTR-128 (%s): ", v == AES_DECRYPT ? "dec" : "enc");
        unhex(nc, ctr_nc[u], 16);
        unhex(key, ctr_key[u], 16);
        aes_setkey_enc(&ctx, key, 128);
        unhex(buf, v == AES_DECRYPT ? ctr_ct[u] : ctr_pt[u], (size_t)len);
        if (aes_crypt_ctr(&ctx, len, &off, nc, sb, buf, buf)) goto fail;
        if (!hexeq(buf, v == AES_DECRYPT ? ctr_pt[u] : ctr_ct[u], (size_t)len)) goto fail;
        if (verbose) printf("passed\n");
    }
    if (verbose) printf("\n");
    return 0;
fail:
    if (verbose) printf("failed (%s)\n", bc_last_error());
    return 1;
}
