/*
 * arc4.h — drop-in for the reference's arc4.h (PolarSSL ARC4 with the local
 * setup / prep / crypt split, arc4.c:43-112), backed by the B200 engine.
 * A single RC4 stream is a strictly serial chain ("replicas
 * only"), so these per-stream calls exist for parity; throughput comes from
 * bc_arc4_batch_host / bc_arc4_batch_dev (bulkcipher.h) over many streams.
 */
#ifndef BULKCIPHER_DROPIN_ARC4_H
#define BULKCIPHER_DROPIN_ARC4_H

#include <stddef.h>
#include <string.h>

/* ABI of arc4.h:35-41 (264 bytes) */
typedef struct {
    int x;
    int y;
    unsigned char m[256];
} arc4_context;

#ifdef __cplusplus
extern "C" {
#endif

/* arc4.c:43-67 — KSA.  Only the first 256 key bytes matter; keylen 0 behaves
 * like the reference (key[0] repeated).  void: a GPU failure is
 * reported on stderr and through bc_void_status() (bulkcipher.h); ctx is left untouched. */
void arc4_setup(arc4_context *ctx, const unsigned char *key, unsigned int keylen);
/* arc4.c:72-97 — PRGA into `keystream`; ctx->x, y, m persist.  Returns 0. */
int arc4_prep(arc4_context *ctx, size_t length, unsigned char *keystream);
/* arc4.c:101-112 — output = input ^ keystream (no context involved).  Returns 0. */
int arc4_crypt(size_t length, const unsigned char *input, unsigned char *keystream,
               unsigned char *output);
/* arc4.c:148-183 — the three Rescorla vectors through the GPU path.  0 = pass. */
int arc4_self_test(int verbose);

#ifdef __cplusplus
}
#endif
#endif
