/*
 * rc4_dropin.c — the reference's arc4.h and rc4.h entry points on top of the
 * B200 engine (single-stream forms; the batched kernel with one stream).
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "arc4.h"
#include "rc4.h"
#include "bulkcipher.h"

void bc_set_error_str(const char *msg);              /* bc_engine.c */
void bc_void_call_failed(int rc, const char *fn);   /* bc_engine.c: stderr + bc_void_status(), abort only on request */

/* The reference's calls are void: a failure is reported (stderr, bc_last_error(), bc_void_status())
 * and the output buffer is left untouched — never garbage, never a silent CPU fallback. */
static void die_on(int rc, const char *fn)
{
    if (rc != BC_OK) bc_void_call_failed(rc, fn);
}

/* arc4.c:43-67 */
void arc4_setup(arc4_context *ctx, const unsigned char *key, unsigned int keylen)
{
    unsigned char k0[1];
    if (keylen == 0) { k0[0] = key[0]; key = k0; keylen = 1; }   /* the reference re-reads key[0] (arc4.c:58-62) */
    if (keylen > 256) keylen = 256;                               /* later bytes are never reached */
    die_on(bc_arc4_batch_host(1, key, keylen, 0, NULL, NULL, NULL, ctx), "arc4_setup");
}

/* arc4.c:72-97 */
int arc4_prep(arc4_context *ctx, size_t length, unsigned char *keystream)
{
    if (length == 0) return 0;
    return bc_arc4_batch_host(1, NULL, 0, length, NULL, keystream, ctx, ctx);
}

/* arc4.c:101-112 */
int arc4_crypt(size_t length, const unsigned char *input, unsigned char *keystream, unsigned char *output)
{
    return bc_xor_host(input, keystream, output, length);
}

/* arc4.c:148-183 with the vectors of arc4.c:124-143 */
int arc4_self_test(int verbose)
{
    static const unsigned char key[3][8] = { { 0x01, 0x23, 0x45, 0x67, 0x89, 0xAB, 0xCD, 0xEF },
                                             { 0x01, 0x23, 0x45, 0x67, 0x89, 0xAB, 0xCD, 0xEF },
                                             { 0, 0, 0, 0, 0, 0, 0, 0 } };
    static const unsigned char pt[3][8] = { { 0x01, 0x23, 0x45, 0x67, 0x89, 0xAB, 0xCD, 0xEF },
                                            { 0, 0, 0, 0, 0, 0, 0, 0 }, { 0, 0, 0, 0, 0, 0, 0, 0 } };
    static const unsigned char ct[3][8] = { { 0x75, 0xB7, 0x87, 0x80, 0x99, 0xE0, 0xC5, 0x96 },
                                            { 0x74, 0x94, 0xC2, 0xE7, 0x10, 0x4B, 0x08, 0x79 },
                                            { 0xDE, 0x18, 0x89, 0x41, 0xA3, 0x37, 0x5D, 0x3A } };
    for (int i = 0; i < 3; i++) {
        unsigned char ibuf[8], obuf[8], ks[8];
        arc4_context ctx;
        if (verbose) printf("  ARC4 test #%d: ", i + 1);
        memcpy(ibuf, pt[i], 8);
        arc4_setup(&ctx, key[i], 8);
        if (arc4_prep(&ctx, 8, ks) || arc4_crypt(8, ibuf, ks, obuf) || memcmp(obuf, ct[i], 8) != 0) {
            if (verbose) printf("failed\n");
            return 1;
        }
        if (verbose) printf("passed\n");
    }
    if (verbose) printf("\n");
    return 0;
}

/* rc4.c:56-72 (intended semantics) — same KSA, state mapped perm<->m */
void rc4_init(struct rc4_state *const state, const char *key, int keylen)
{
    arc4_context c;
    if (keylen < 1) {
        bc_set_error_str("keylen must be >= 1 (the reference divides by it, rc4.c:69)");
        die_on(BC_ERR_BAD_ARG, "rc4_init");
        return;
    }
    /* key[i % keylen] over 256 steps only ever reads the first 256 bytes */
    const int rc = bc_arc4_batch_host(1, key, (unsigned)(keylen > 256 ? 256 : keylen), 0, NULL, NULL, NULL, &c);
    if (rc != BC_OK) { die_on(rc, "rc4_init"); return; }         /* state left untouched */
    memcpy(state->perm, c.m, 256);
    state->index1 = 0;
    state->index2 = 0;
}

/* rc4.c:80-99 (intended semantics) */
void rc4_crypt(struct rc4_state *const state, const char *inbuf, char *outbuf, int buflen)
{
    arc4_context c;
    if (buflen <= 0) return;
    c.x = state->index1 & 0xFF;
    c.y = state->index2 & 0xFF;
    memcpy(c.m, state->perm, 256);
    const int rc = bc_arc4_batch_host(1, NULL, 0, (size_t)buflen, inbuf, outbuf, &c, &c);
    if (rc != BC_OK) { die_on(rc, "rc4_crypt"); return; }        /* state and output left untouched */
    memcpy(state->perm, c.m, 256);
    state->index1 = c.x;
    state->index2 = c.y;
}
