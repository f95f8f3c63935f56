/*
 * oracle.c — plain-C restatement of the reference's AES (ECB/CBC/CTR), AES-NI
 * bulk calls, ARC4 and RC4.  TEST INFRASTRUCTURE ONLY (see oracle.h): it is the
 * checker for the CUDA path and the "port" CPU baseline; the product never
 * calls it.  Parity status: pinned (tests/test_oracle.py).
 *
 * The code is written from the algorithms' definitions (FIPS-197, SP 800-38A,
 * the RC4 description) and from the *behaviour* of the reference, which each
 * function cites as file:line relative to the reference tree.  State words are
 * little-endian 32-bit columns like the reference's GET_ULONG_LE convention
 * (aes-modes/aes.c:42-60) so that round keys compare word for word.
 */
#include "oracle.h"
#include <string.h>

/* ------------------------------------------------------------------ tables */

static uint8_t  SB[256], ISB[256];          /* aes.c FSb / RSb   (:385-404) */
static uint32_t FT[4][256], RT[4][256];     /* aes.c FT0-3 / RT0-3 (:409-434) */
static int      tables_ready;

static uint8_t gf_mul(uint8_t a, uint8_t b)
{
    uint8_t p = 0;
    while (b) {
        if (b & 1) p ^= a;
        a = (uint8_t)((a << 1) ^ ((a & 0x80) ? 0x1B : 0));   /* XTIME, aes.c:355 */
        b >>= 1;
    }
    return p;
}

static uint32_t rotl32(uint32_t v, int n) { return (v << n) | (v >> (32 - n)); }

/* Same tables as aes_gen_tables (aes.c:361-435), derived directly from the
 * definitions: S(x) = affine(x^-1), FT0[x] = {02,01,01,03}·S(x) from the low
 * byte up (:415-418), RT0[x] = {0E,09,0D,0B}·S^-1(x) (:426-429), and table k is
 * table 0 rotated left by 8k bits (:420-422, :431-433). */
static void build_tables(void)
{
    if (tables_ready) return;
    for (int x = 0; x < 256; x++) {
        uint8_t inv = 0;
        if (x) for (int c = 1; c < 256; c++) if (gf_mul((uint8_t)x, (uint8_t)c) == 1) { inv = (uint8_t)c; break; }
        uint8_t s = inv, r = inv;
        for (int i = 0; i < 4; i++) { r = (uint8_t)((r << 1) | (r >> 7)); s ^= r; }
        s ^= 0x63;
        SB[x] = s;
        ISB[s] = (uint8_t)x;
    }
    for (int x = 0; x < 256; x++) {
        uint8_t s = SB[x], q = ISB[x];
        uint32_t f = (uint32_t)gf_mul(s, 2) | ((uint32_t)s << 8) | ((uint32_t)s << 16) |
                     ((uint32_t)gf_mul(s, 3) << 24);
        uint32_t r = (uint32_t)gf_mul(q, 0x0E) | ((uint32_t)gf_mul(q, 0x09) << 8) |
                     ((uint32_t)gf_mul(q, 0x0D) << 16) | ((uint32_t)gf_mul(q, 0x0B) << 24);
        for (int k = 0; k < 4; k++) {
            FT[k][x] = k ? rotl32(f, 8 * k) : f;
            RT[k][x] = k ? rotl32(r, 8 * k) : r;
        }
    }
    tables_ready = 1;
}

static uint32_t load_le32(const uint8_t *p)
{
    return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}

static void store_le32(uint8_t *p, uint32_t v)
{
    p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24);
}

static uint32_t sub_word(uint32_t w)
{
    return (uint32_t)SB[w & 0xFF] | ((uint32_t)SB[(w >> 8) & 0xFF] << 8) |
           ((uint32_t)SB[(w >> 16) & 0xFF] << 16) | ((uint32_t)SB[w >> 24] << 24);
}

/* ------------------------------------------------------------ key schedule */

/* aes_setkey_enc, aes.c:442-542.  One generic FIPS-197 expansion loop covers
 * the three unrolled cases (:478-489, :494-507, :512-533): on LE words RotWord
 * is a right rotate by 8 and the round constant lands in the low byte. */
int orc_aes_setkey_enc(orc_aes_key *k, const uint8_t *key, unsigned keybits)
{
    build_tables();                                   /* lazy init, aes.c:448-452 */
    if (keybits != 128 && keybits != 192 && keybits != 256)
        return ORC_ERR_KEYLEN;                        /* aes.c:455-461 */
    int nk = (int)keybits / 32;
    k->nr = nk + 6;
    int total = 4 * (k->nr + 1);
    for (int i = 0; i < nk; i++) k->rk[i] = load_le32(key + 4 * i);   /* aes.c:469-472 */
    uint8_t rcon = 1;
    for (int i = nk; i < total; i++) {
        uint32_t t = k->rk[i - 1];
        if (i % nk == 0) {
            t = sub_word((t >> 8) | (t << 24)) ^ rcon;
            rcon = gf_mul(rcon, 2);                   /* RCON[], aes.c:378-383 */
        } else if (nk == 8 && i % nk == 4) {
            t = sub_word(t);                          /* extra SubWord, aes.c:523-527 */
        }
        k->rk[i] = k->rk[i - nk] ^ t;
    }
    return 0;
}

/* aes_setkey_dec, aes.c:547-599: equivalent-inverse-cipher schedule.  Round
 * keys in reverse order; the nr-1 middle ones go through InvMixColumns, which
 * the reference gets as RTk[FSb[byte k]] (:580-589). */
int orc_aes_setkey_dec(orc_aes_key *k, const uint8_t *key, unsigned keybits)
{
    orc_aes_key e;
    int rc = orc_aes_setkey_enc(&e, key, keybits);
    if (rc) return rc;                                /* aes.c:555-561, :569-571 */
    k->nr = e.nr;
    for (int r = 0; r <= e.nr; r++) {
        const uint32_t *src = e.rk + 4 * (e.nr - r);
        for (int c = 0; c < 4; c++) {
            uint32_t w = src[c];
            if (r != 0 && r != e.nr)
                w = RT[0][SB[w & 0xFF]] ^ RT[1][SB[(w >> 8) & 0xFF]] ^
                    RT[2][SB[(w >> 16) & 0xFF]] ^ RT[3][SB[w >> 24]];
            k->rk[4 * r + c] = w;
        }
    }
    memset(&e, 0, sizeof e);                          /* aes.c:596 */
    return 0;
}

/* --------------------------------------------------------------- one block */

/* aes_crypt_ecb, aes.c:650-752 with AES_FROUND/AES_RROUND (:601-645).  Column c
 * of the next state takes byte j from column (c+j)%4 when encrypting and from
 * column (c-j)%4 when decrypting — that is ShiftRows / InvShiftRows folded into
 * the table indices.  `mode` only selects the tables; the schedule in k must
 * already match (aes.c:677,711). */
void orc_aes_ecb_block(const orc_aes_key *k, int mode, const uint8_t in[16], uint8_t out[16])
{
    build_tables();
    const uint32_t (*T)[256] = (mode == ORC_DECRYPT) ? RT : FT;
    const uint8_t *S = (mode == ORC_DECRYPT) ? ISB : SB;
    int dir = (mode == ORC_DECRYPT) ? 3 : 1;          /* +1 or -1 (mod 4) column step */
    const uint32_t *rk = k->rk;
    uint32_t s[4], t[4];

    for (int c = 0; c < 4; c++) s[c] = load_le32(in + 4 * c) ^ *rk++;    /* aes.c:672-675 */

    for (int round = 1; round < k->nr; round++) {
        for (int c = 0; c < 4; c++)
            t[c] = *rk++ ^ T[0][s[c] & 0xFF] ^
                           T[1][(s[(c + dir) & 3] >> 8) & 0xFF] ^
                           T[2][(s[(c + 2 * dir) & 3] >> 16) & 0xFF] ^
                           T[3][s[(c + 3 * dir) & 3] >> 24];
        memcpy(s, t, sizeof s);
    }
    for (int c = 0; c < 4; c++) {                     /* last round, aes.c:687-743 */
        uint32_t w = (uint32_t)S[s[c] & 0xFF] |
                     ((uint32_t)S[(s[(c + dir) & 3] >> 8) & 0xFF] << 8) |
                     ((uint32_t)S[(s[(c + 2 * dir) & 3] >> 16) & 0xFF] << 16) |
                     ((uint32_t)S[s[(c + 3 * dir) & 3] >> 24] << 24);
        store_le32(out + 4 * c, w ^ *rk++);           /* aes.c:746-749 */
    }
}

/* The harness loop over aes_crypt_ecb, aes-modes/test.c:37-41. */
void orc_aes_ecb(const orc_aes_key *k, int mode, size_t length, const uint8_t *in, uint8_t *out)
{
    for (size_t off = 0; off + 16 <= length; off += 16)
        orc_aes_ecb_block(k, mode, in + off, out + off);
}

/* ------------------------------------------------------------------- modes */

/* aes_crypt_cbc, aes.c:757-816.  iv is updated to the last ciphertext block in
 * both directions (:792, :807); in == out is legal (:786, :806). */
int orc_aes_cbc(const orc_aes_key *k, int mode, size_t length, uint8_t iv[16],
                const uint8_t *in, uint8_t *out)
{
    if (length % 16) return ORC_ERR_INPUTLEN;         /* aes.c:767-768 */
    uint8_t blk[16];
    for (size_t off = 0; off < length; off += 16) {
        if (mode == ORC_DECRYPT) {
            uint8_t ct[16];
            memcpy(ct, in + off, 16);                 /* keep C_i before in-place overwrite */
            orc_aes_ecb_block(k, mode, ct, blk);
            for (int i = 0; i < 16; i++) out[off + i] = blk[i] ^ iv[i];
            memcpy(iv, ct, 16);
        } else {
            for (int i = 0; i < 16; i++) blk[i] = in[off + i] ^ iv[i];
            orc_aes_ecb_block(k, mode, blk, out + off);
            memcpy(iv, out + off, 16);
        }
    }
    return 0;
}

/* aes_crypt_ctr, aes.c:869-900.  Byte-granular; a keystream block is produced
 * when the offset is 0 and the 16-byte counter is then incremented as one
 * 128-bit big-endian integer (:884-888).  The reference takes `int length`;
 * size_t here only widens the range. */
int orc_aes_ctr(const orc_aes_key *k, size_t length, int *nc_off, uint8_t nonce_counter[16],
                uint8_t stream_block[16], const uint8_t *in, uint8_t *out)
{
    int n = *nc_off;
    for (size_t p = 0; p < length; p++) {
        if (n == 0) {
            orc_aes_ecb_block(k, ORC_ENCRYPT, nonce_counter, stream_block);
            for (int i = 15; i >= 0; i--)
                if (++nonce_counter[i] != 0) break;
        }
        out[p] = in[p] ^ stream_block[n];
        n = (n + 1) & 15;
    }
    *nc_off = n;
    return 0;
}

/* aes_crypt_cfb128, aes.c:822-862 (compiled out of the reference build because
 * POLARSSL_CIPHER_MODE_CFB is undefined, but declared at aes.h:124 and covered by
 * the SP 800-38A vectors at aes.c:958-1015).  Byte-granular; iv doubles as the
 * keystream block, *iv_off is the position inside it; always the ENCRYPT schedule. */
int orc_aes_cfb128(const orc_aes_key *k, int mode, size_t length, int *iv_off, uint8_t iv[16],
                   const uint8_t *in, uint8_t *out)
{
    int n = *iv_off;
    for (size_t p = 0; p < length; p++) {
        if (n == 0) orc_aes_ecb_block(k, ORC_ENCRYPT, iv, iv);
        if (mode == ORC_DECRYPT) {
            uint8_t c = in[p];
            out[p] = c ^ iv[n];
            iv[n] = c;
        } else {
            iv[n] = out[p] = iv[n] ^ in[p];
        }
        n = (n + 1) & 15;
    }
    *iv_off = n;
    return 0;
}

/* ------------------------------------------------------------ aesni.h calls */

/* Packed schedule = 16 bytes per round key, i.e. the LE words back to back
 * (SURVEY finding 2; Appendix A "repacking"). */
void orc_aesni_pack(const orc_aes_key *k, uint8_t *sched)
{
    for (int i = 0; i < 4 * (k->nr + 1); i++) store_le32(sched + 4 * i, k->rk[i]);
}

/* AES_256_Key_Expansion, aesni.c:38-77: the AESKEYGENASSIST sequence computes
 * exactly the FIPS-197 AES-256 encryption schedule, stored packed. */
void orc_aesni_key256(const uint8_t *userkey, uint8_t *sched)
{
    orc_aes_key k;
    orc_aes_setkey_enc(&k, userkey, 256);
    orc_aesni_pack(&k, sched);
}

static void unpack_sched(orc_aes_key *k, const uint8_t *sched, int rounds)
{
    k->nr = rounds;
    for (int i = 0; i < 4 * (rounds + 1); i++) k->rk[i] = load_le32(sched + 4 * i);
}

/* aesni.c:86,107,130 — a ragged tail is rounded UP to a whole block (the
 * reference over-reads/over-writes; callers should only pass multiples of 16). */
static unsigned long blocks_rounded_up(unsigned long length)
{
    return length / 16 + (length % 16 ? 1 : 0);
}

/* AES_ECB_encrypt, aesni.c:79-97: xor key[0], rounds-1 × AESENC, AESENCLAST.
 * AESENC is ShiftRows∘SubBytes∘MixColumns∘AddRoundKey = one table round. */
void orc_aesni_ecb_encrypt(const uint8_t *in, uint8_t *out, unsigned long length,
                           const uint8_t *sched, int rounds)
{
    orc_aes_key k;
    unpack_sched(&k, sched, rounds);
    unsigned long nb = blocks_rounded_up(length);
    for (unsigned long i = 0; i < nb; i++) orc_aes_ecb_block(&k, ORC_ENCRYPT, in + 16 * i, out + 16 * i);
}

/* AES_ECB_decrypt, aesni.c:99-118: AESDEC rounds on the equivalent-inverse
 * schedule (the repacked aes_setkey_dec output, SURVEY finding 2). */
void orc_aesni_ecb_decrypt(const uint8_t *in, uint8_t *out, unsigned long length,
                           const uint8_t *sched, int rounds)
{
    orc_aes_key k;
    unpack_sched(&k, sched, rounds);
    unsigned long nb = blocks_rounded_up(length);
    for (unsigned long i = 0; i < nb; i++) orc_aes_ecb_block(&k, ORC_DECRYPT, in + 16 * i, out + 16 * i);
}

/* AES_CTR_encrypt, aesni.c:120-152.  Counter block i (0-based) is
 * nonce[0..3] ‖ ivec[0..3] ‖ BE64(q + 1 + i) with q = BE32(ivec[4..7]) << 32;
 * the add is a 64-bit wrapping add (_mm_add_epi64, :139,:143) that never
 * carries into bytes 0..7. */
void orc_aesni_ctr_encrypt(const uint8_t *in, uint8_t *out, const uint8_t ivec[8],
                           const uint8_t nonce[4], unsigned long length,
                           const uint8_t *sched, int rounds)
{
    orc_aes_key k;
    unpack_sched(&k, sched, rounds);
    unsigned long nb = blocks_rounded_up(length);
    uint64_t q = ((uint64_t)ivec[4] << 56) | ((uint64_t)ivec[5] << 48) |
                 ((uint64_t)ivec[6] << 40) | ((uint64_t)ivec[7] << 32);
    uint8_t cb[16], ks[16];
    memcpy(cb, nonce, 4);
    memcpy(cb + 4, ivec, 4);
    for (unsigned long i = 0; i < nb; i++) {
        uint64_t c = q + 1 + (uint64_t)i;
        for (int b = 0; b < 8; b++) cb[8 + b] = (uint8_t)(c >> (56 - 8 * b));
        orc_aes_ecb_block(&k, ORC_ENCRYPT, cb, ks);
        for (int b = 0; b < 16; b++) out[16 * i + b] = in[16 * i + b] ^ ks[b];
    }
}

/* -------------------------------------------------------------------- ARC4 */

/* arc4_setup, arc4.c:43-67 (KSA). */
void orc_arc4_setup(orc_arc4_ctx *c, const uint8_t *key, unsigned keylen)
{
    c->x = c->y = 0;
    for (int i = 0; i < 256; i++) c->m[i] = (uint8_t)i;
    unsigned k = 0;
    uint8_t j = 0;
    for (int i = 0; i < 256; i++) {
        if (k >= keylen) k = 0;                       /* arc4.c:60 */
        uint8_t a = c->m[i];
        j = (uint8_t)(j + a + key[k++]);
        c->m[i] = c->m[j];
        c->m[j] = a;
    }
}

/* arc4_prep, arc4.c:72-97 (PRGA into a keystream buffer; x,y persisted). */
int orc_arc4_prep(orc_arc4_ctx *c, size_t length, uint8_t *keystream)
{
    uint8_t x = (uint8_t)c->x, y = (uint8_t)c->y;
    for (size_t i = 0; i < length; i++) {
        x = (uint8_t)(x + 1);
        uint8_t a = c->m[x];
        y = (uint8_t)(y + a);
        uint8_t b = c->m[y];
        c->m[x] = b;
        c->m[y] = a;
        keystream[i] = c->m[(uint8_t)(a + b)];
    }
    c->x = x;
    c->y = y;
    return 0;
}

/* arc4_crypt, arc4.c:101-112: a bare XOR with the prepared keystream. */
int orc_arc4_crypt(size_t length, const uint8_t *in, const uint8_t *keystream, uint8_t *out)
{
    for (size_t i = 0; i < length; i++) out[i] = in[i] ^ keystream[i];
    return 0;
}

/* -------------------------------------------------------------------- RC4 */

/* rc4_init, rc4.c:56-72, with the upstream unsigned-byte index types
 * (rc4.c:36 names the FreeBSD original), i.e. every index reduced mod 256. */
void orc_rc4_init(orc_rc4_state *s, const uint8_t *key, int keylen)
{
    for (int i = 0; i < 256; i++) s->perm[i] = (uint8_t)i;
    s->index1 = s->index2 = 0;
    uint8_t j = 0;
    for (int i = 0; i < 256; i++) {
        j = (uint8_t)(j + s->perm[i] + key[i % keylen]);     /* rc4.c:69 */
        uint8_t t = s->perm[i]; s->perm[i] = s->perm[j]; s->perm[j] = t;
    }
}

/* rc4_crypt, rc4.c:80-99: PRGA fused with the XOR; in == out allowed (:76). */
void orc_rc4_crypt(orc_rc4_state *s, const uint8_t *in, uint8_t *out, int buflen)
{
    uint8_t i1 = (uint8_t)s->index1, i2 = (uint8_t)s->index2;
    for (int i = 0; i < buflen; i++) {
        i1 = (uint8_t)(i1 + 1);
        i2 = (uint8_t)(i2 + s->perm[i1]);
        uint8_t t = s->perm[i1]; s->perm[i1] = s->perm[i2]; s->perm[i2] = t;
        out[i] = in[i] ^ s->perm[(uint8_t)(s->perm[i1] + s->perm[i2])];
    }
    s->index1 = i1;
    s->index2 = i2;
}

/* ------------------------------------------------------------- batch forms */

/* n independent streams, each arc4_setup → arc4_prep → arc4_crypt like
 * test.c:86-90 + :44-58 (BASELINE config 4). */
void orc_arc4_batch(size_t n_streams, const uint8_t *keys, unsigned keylen,
                    size_t length, const uint8_t *in, uint8_t *out)
{
    uint8_t ks[4096];
    for (size_t s = 0; s < n_streams; s++) {
        orc_arc4_ctx c;
        orc_arc4_setup(&c, keys + s * keylen, keylen);
        for (size_t off = 0; off < length; off += sizeof ks) {
            size_t n = length - off < sizeof ks ? length - off : sizeof ks;
            orc_arc4_prep(&c, n, ks);
            orc_arc4_crypt(n, in + s * length + off, ks, out + s * length + off);
        }
    }
}

/* n independent CBC-encrypt messages (BASELINE config 3), each one an
 * aes_crypt_cbc(AES_ENCRYPT) call with its own iv. */
void orc_aes_cbc_encrypt_batch(const orc_aes_key *k, size_t n_msgs, size_t msg_len,
                               uint8_t *ivs, const uint8_t *in, uint8_t *out)
{
    for (size_t m = 0; m < n_msgs; m++)
        orc_aes_cbc(k, ORC_ENCRYPT, msg_len, ivs + 16 * m, in + m * msg_len, out + m * msg_len);
}

/* ------------------------------------------------------- harness test input */

#include <stdlib.h>

/* The reference harness's input generator: srand(seed) once
 * (aes-modes/test.c:354, test.c:131), then msg[i] = rand() % 255
 * (aes-modes/test.c:54-56) followed by key[i] = rand() % 255 (:60-62).
 * glibc rand(); used only to rebuild the 16 MiB BASELINE config-1 input that
 * the golden digests in tests/golden/ were taken on. */
void orc_harness_input(unsigned seed, uint8_t *msg, size_t n_msg, uint8_t *key, size_t n_key)
{
    srand(seed);
    for (size_t i = 0; i < n_msg; i++) msg[i] = (uint8_t)(rand() % 255);
    for (size_t i = 0; i < n_key; i++) key[i] = (uint8_t)(rand() % 255);
}
