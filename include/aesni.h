/*
 * aesni.h — drop-in for the reference's aes-modes/aesni.h.  The reference
 * implements these bulk calls with AES-NI intrinsics on the CPU
 * (aes-modes/aesni.c); here the same symbols run on the B200 through
 * libbulkcipher.so.  `key` is the packed schedule: 16 bytes per round key
 * (what AES_256_Key_Expansion writes, or the low 32 bits of each
 * aes_context.rk[] slot back to back — SURVEY finding 2).
 *
 * The functions are `void` in the reference, so a GPU failure cannot be
 * returned: it is reported on stderr, the output buffer is left untouched (no
 * silent garbage, no CPU fallback) and the code waits in bc_void_status()
 * (bulkcipher.h); BULKCIPHER_ABORT_ON_ERROR=1 makes it abort() instead.  Use
 * the bc_* calls of bulkcipher.h where a return code is wanted.
 */
#ifndef BULKCIPHER_DROPIN_AESNI_H
#define BULKCIPHER_DROPIN_AESNI_H

#if !defined(ALIGN16)
# if defined(__GNUC__)
#  define ALIGN16 __attribute__((aligned(16)))
# else
#  define ALIGN16 __declspec(align(16))
# endif
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* aes-modes/aesni.c:7-11 answers "is the accelerated path usable?" from CPUID;
 * here: nonzero iff a CUDA device is usable. */
int CheckAESSupport(void);

/* aes-modes/aesni.c:38-77: AES-256 encryption schedule, 15 x 16 packed bytes.
 * Host computation, no alignment requirement on `key`. */
void AES_256_Key_Expansion(const unsigned char *userkey, unsigned char *key);
/* Companions the reference lacks: 176- and 208-byte schedules. */
void AES_128_Key_Expansion(const unsigned char *userkey, unsigned char *key);
void AES_192_Key_Expansion(const unsigned char *userkey, unsigned char *key);

/* aes-modes/aesni.c:79-97.  length is rounded UP to whole blocks like the
 * reference (:86) — pass multiples of 16. */
void AES_ECB_encrypt(const unsigned char *in, unsigned char *out, unsigned long length,
                     const unsigned char *key, int number_of_rounds);
/* aes-modes/aesni.c:99-118; key = equivalent-inverse schedule (repacked aes_setkey_dec). */
void AES_ECB_decrypt(const unsigned char *in, unsigned char *out, unsigned long length,
                     const char *key, int number_of_rounds);
/* aes-modes/aesni.c:120-152.  Counter block i = nonce[4] | ivec[0..3] |
 * BE64((BE32(ivec[4..7]) << 32) + 1 + i), 64-bit wrapping add. */
void AES_CTR_encrypt(const unsigned char *in, unsigned char *out, const unsigned char ivec[8],
                     const unsigned char nonce[4], unsigned long length,
                     const unsigned char *key, int number_of_rounds);

#ifdef __cplusplus
}
#endif
#endif
