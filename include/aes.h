/*
 * aes.h — drop-in for the reference's aes-modes/aes.h, backed by the B200
 * bulk-cipher engine (libbulkcipher.so).  Same symbols, same struct layout,
 * same return codes, same outputs as aes-modes/aes.c, so aes-modes/test.c
 * relinks unchanged; the cipher rounds run on the GPU (there is no CPU
 * fallback: without a CUDA device the calls return BC_ERR_* from bulkcipher.h).
 *
 * Buffers are HOST pointers, as in the reference (device pointers are also
 * accepted and then used in place without copies).  in == out is legal for
 * every mode (aes.c:786,806,891).  For buffers already in HBM, lengths over
 * 2 GiB, CTR shards and batched CBC-encrypt use bulkcipher.h directly.
 */
#ifndef BULKCIPHER_DROPIN_AES_H
#define BULKCIPHER_DROPIN_AES_H

#include <stddef.h>
#include <string.h>

#define AES_ENCRYPT 1     /* aes-modes/aes.h:33 */
#define AES_DECRYPT 0     /* aes-modes/aes.h:34 */

#define POLARSSL_ERR_AES_INVALID_KEY_LENGTH   -0x0020   /* aes-modes/aes.h:35 */
#define POLARSSL_ERR_AES_INVALID_INPUT_LENGTH -0x0022   /* aes-modes/aes.h:36 */

/* ABI of aes-modes/aes.h:41-47 (560 bytes on LP64): each 32-bit round-key word
 * sits in the low half of one `unsigned long` slot; rk points into buf. */
typedef struct {
    int nr;
    unsigned long *rk;
    unsigned long buf[68];
} aes_context;

#ifdef __cplusplus
extern "C" {
#endif

/* aes-modes/aes.c:442-542.  keysize 128/192/256 bits, else -0x20 with ctx untouched.
 * Runs on the host: it is O(240 bytes) per key and must fill ctx->buf identically. */
int aes_setkey_enc(aes_context *ctx, const unsigned char *key, unsigned int keysize);
/* aes-modes/aes.c:547-599: equivalent-inverse-cipher schedule. */
int aes_setkey_dec(aes_context *ctx, const unsigned char *key, unsigned int keysize);

/* aes-modes/aes.c:650-752: one 16-byte block.  `mode` only selects the tables;
 * ctx must hold the matching schedule.  A lone call is one GPU round trip.  A
 * thread whose calls walk consecutive addresses (the reference's loop,
 * aes-modes/test.c:37-41) is served from a per-thread read-ahead window: the
 * library copies the next 4 KiB .. 4 MiB of the caller's input without ever
 * faulting (process_vm_readv), runs the window through the GPU in one call, and
 * answers a later call from it only when that call's 16 input bytes equal the
 * bytes the stored output was computed from (same key schedule, same mode) —
 * exact by construction.  Bulk callers are still better off with
 * AES_ECB_encrypt (aesni.h) or bc_aes_ecb_host. */
int aes_crypt_ecb(aes_context *ctx, int mode, const unsigned char input[16], unsigned char output[16]);

/* aes-modes/aes.c:757-816.  length % 16 != 0 -> -0x22, nothing written.  iv is
 * updated to the last ciphertext block.  Decrypt is block-parallel on the GPU;
 * encrypt is one serial chain (one GPU thread) — batch independent messages
 * with bc_aes_cbc_encrypt_batch_host instead. */
int aes_crypt_cbc(aes_context *ctx, int mode, size_t length, unsigned char iv[16],
                  const unsigned char *input, unsigned char *output);

/* aes-modes/aes.h:124 / aes.c:822-862 (the reference declares it but compiles it out).
 * Byte-granular and resumable through *iv_off; ctx holds the ENCRYPT schedule for
 * both directions.  Decrypt is block-parallel on the GPU, encrypt one serial chain. */
int aes_crypt_cfb128(aes_context *ctx, int mode, size_t length, int *iv_off, unsigned char iv[16],
                     const unsigned char *input, unsigned char *output);

/* aes-modes/aes.c:869-900.  Byte-granular, resumable through *nc_off /
 * nonce_counter / stream_block, 128-bit big-endian counter increment. */
int aes_crypt_ctr(aes_context *ctx, int length, int *nc_off, unsigned char nonce_counter[16],
                  unsigned char stream_block[16], const unsigned char *input, unsigned char *output);

/* aes-modes/aes.c:1084-1330 (the reference compiles it only under
 * POLARSSL_SELF_TEST): Monte-Carlo ECB/CBC for 128/192/256 and the RFC 3686 CTR
 * vectors, here run through the GPU path.  0 = pass, 1 = fail. */
int aes_self_test(int verbose);

#ifdef __cplusplus
}
#endif
#endif
