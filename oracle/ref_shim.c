/*
 * ref_shim.c — thin conveniences linked into oracle/_ref/libref.so next to the
 * UNMODIFIED reference objects.  TEST INFRASTRUCTURE ONLY.  This file contains
 * no cipher arithmetic: every byte of output comes from the reference's own
 * aes.c / aesni.c / arc4.c (compiled by oracle/Makefile with ref_* renames).
 * The loops below are the reference harness's call patterns
 * (aes-modes/test.c:37-41, test.c:86-90 + :44-58) so that Python can time and
 * compare whole buffers with one foreign call.
 */
#include <stddef.h>
#include <string.h>
#include "aes.h"     /* $(REF)/aes-modes/aes.h   */
#include "aesni.h"   /* $(REF)/aes-modes/aesni.h */
#include "arc4.h"    /* $(REF)/arc4.h            */

size_t ref_sizeof_aes_context(void)  { return sizeof(aes_context); }
size_t ref_sizeof_arc4_context(void) { return sizeof(arc4_context); }

/* Low 32 bits of each rk[] slot, back to back: the layout aesni.c expects
 * (SURVEY finding 2 / Appendix A). */
void ref_pack_schedule(const aes_context *ctx, unsigned char *sched)
{
    for (int i = 0; i < 4 * (ctx->nr + 1); i++) {
        unsigned long w = ctx->rk[i];
        sched[4 * i]     = (unsigned char)(w);
        sched[4 * i + 1] = (unsigned char)(w >> 8);
        sched[4 * i + 2] = (unsigned char)(w >> 16);
        sched[4 * i + 3] = (unsigned char)(w >> 24);
    }
}

/* One aes_crypt_ecb call per 16 bytes, as ecb_test_thread does. */
void ref_aes_ecb_bulk(aes_context *ctx, int mode, size_t length,
                      const unsigned char *in, unsigned char *out)
{
    for (size_t off = 0; off + 16 <= length; off += 16)
        aes_crypt_ecb(ctx, mode, in + off, out + off);
}

/* aes_crypt_ctr takes `int length`; chain calls below 2 GiB, carrying the
 * resume state exactly as the API defines it. */
int ref_aes_ctr_bulk(aes_context *ctx, size_t length, int *nc_off,
                     unsigned char nonce_counter[16], unsigned char stream_block[16],
                     const unsigned char *in, unsigned char *out)
{
    const size_t step = (size_t)1 << 30;
    while (length) {
        size_t n = length < step ? length : step;
        int rc = aes_crypt_ctr(ctx, (int)n, nc_off, nonce_counter, stream_block, in, out);
        if (rc) return rc;
        in += n; out += n; length -= n;
    }
    return 0;
}

/* n independent streams: arc4_setup -> arc4_prep -> arc4_crypt each. */
void ref_arc4_batch(size_t n_streams, const unsigned char *keys, unsigned int keylen,
                    size_t length, const unsigned char *in, unsigned char *out)
{
    unsigned char ks[4096];
    for (size_t s = 0; s < n_streams; s++) {
        arc4_context c;
        arc4_setup(&c, keys + s * keylen, keylen);
        for (size_t off = 0; off < length; off += sizeof ks) {
            size_t n = length - off < sizeof ks ? length - off : sizeof ks;
            arc4_prep(&c, n, ks);
            arc4_crypt(n, in + s * length + off, ks, out + s * length + off);
        }
    }
}

/* n independent CBC-encrypt messages, each with its own iv (updated). */
void ref_aes_cbc_encrypt_batch(aes_context *ctx, size_t n_msgs, size_t msg_len,
                               unsigned char *ivs, const unsigned char *in, unsigned char *out)
{
    for (size_t m = 0; m < n_msgs; m++)
        aes_crypt_cbc(ctx, AES_ENCRYPT, msg_len, ivs + 16 * m, in + m * msg_len, out + m * msg_len);
}
