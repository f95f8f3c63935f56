/*
 * aesni_dropin.c — the reference's aes-modes/aesni.h bulk calls on top of the
 * B200 engine.  The reference runs AESENC/AESDEC on the CPU (aesni.c:79-152);
 * these are the same symbols with the same outputs, computed on the GPU.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "aesni.h"
#include "bulkcipher.h"

int  bc_aes_expand_key(const unsigned char *key, unsigned int keybits, uint32_t rk[60]);   /* aes_dropin.c */

/* The reference's calls are void: fail loudly instead of returning garbage. */
static void die_on(int rc, const char *fn)
{
    if (rc == BC_OK) return;
    fprintf(stderr, "libbulkcipher: %s failed (%d): %s — no CPU fallback, aborting\n", fn, rc, bc_last_error());
    abort();
}

int CheckAESSupport(void)
{
    return bc_device_count() > 0 ? 0x2000000 : 0;     /* the bit aesni.c:10 tests, as "path usable" */
}

static void expand_packed(const unsigned char *userkey, unsigned bits, unsigned char *key)
{
    uint32_t rk[60];
    int nr = bc_aes_expand_key(userkey, bits, rk);
    for (int i = 0; i < 4 * (nr + 1); i++) {
        key[4 * i] = (unsigned char)rk[i];
        key[4 * i + 1] = (unsigned char)(rk[i] >> 8);
        key[4 * i + 2] = (unsigned char)(rk[i] >> 16);
        key[4 * i + 3] = (unsigned char)(rk[i] >> 24);
    }
}

void AES_256_Key_Expansion(const unsigned char *userkey, unsigned char *key) { expand_packed(userkey, 256, key); }
void AES_128_Key_Expansion(const unsigned char *userkey, unsigned char *key) { expand_packed(userkey, 128, key); }
void AES_192_Key_Expansion(const unsigned char *userkey, unsigned char *key) { expand_packed(userkey, 192, key); }

static int unpack(const void *key, int rounds, uint32_t rk[60])
{
    if (rounds != 10 && rounds != 12 && rounds != 14) return 0;
    const unsigned char *k = (const unsigned char *)key;
    for (int i = 0; i < 4 * (rounds + 1); i++)
        rk[i] = (uint32_t)k[4 * i] | ((uint32_t)k[4 * i + 1] << 8) | ((uint32_t)k[4 * i + 2] << 16) |
                ((uint32_t)k[4 * i + 3] << 24);
    return 1;
}

static unsigned long round_up_blocks(unsigned long length)   /* aesni.c:86,107,130 */
{
    return (length / 16 + (length % 16 ? 1 : 0)) * 16;
}

void AES_ECB_encrypt(const unsigned char *in, unsigned char *out, unsigned long length,
                     const unsigned char *key, int number_of_rounds)
{
    uint32_t rk[60];
    if (!unpack(key, number_of_rounds, rk)) die_on(BC_ERR_BAD_ARG, "AES_ECB_encrypt (rounds must be 10/12/14)");
    die_on(bc_aes_ecb_host(rk, number_of_rounds, BC_ENCRYPT, in, out, round_up_blocks(length)), "AES_ECB_encrypt");
}

void AES_ECB_decrypt(const unsigned char *in, unsigned char *out, unsigned long length,
                     const char *key, int number_of_rounds)
{
    uint32_t rk[60];
    if (!unpack(key, number_of_rounds, rk)) die_on(BC_ERR_BAD_ARG, "AES_ECB_decrypt (rounds must be 10/12/14)");
    die_on(bc_aes_ecb_host(rk, number_of_rounds, BC_DECRYPT, in, out, round_up_blocks(length)), "AES_ECB_decrypt");
}

void AES_CTR_encrypt(const unsigned char *in, unsigned char *out, const unsigned char ivec[8],
                     const unsigned char nonce[4], unsigned long length,
                     const unsigned char *key, int number_of_rounds)
{
    uint32_t rk[60];
    if (!unpack(key, number_of_rounds, rk)) die_on(BC_ERR_BAD_ARG, "AES_CTR_encrypt (rounds must be 10/12/14)");
    /* first counter block: nonce | ivec[0..3] | BE64((BE32(ivec[4..7]) << 32) + 1)  (aesni.c:132-139) */
    unsigned char cb[16];
    memcpy(cb, nonce, 4);
    memcpy(cb + 4, ivec, 4);
    uint64_t q = ((uint64_t)ivec[4] << 56) | ((uint64_t)ivec[5] << 48) | ((uint64_t)ivec[6] << 40) |
                 ((uint64_t)ivec[7] << 32);
    q += 1;
    for (int b = 0; b < 8; b++) cb[8 + b] = (unsigned char)(q >> (56 - 8 * b));
    die_on(bc_aes_ctr_host(rk, number_of_rounds, cb, 64, 0, in, out, round_up_blocks(length), NULL), "AES_CTR_encrypt");
}
