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

void bc_set_error_str(const char *msg);              /* bc_engine.c */
void bc_void_call_failed(int rc, const char *fn);   /* bc_engine.c: stderr + bc_void_status(), abort only on request */

/* The reference's calls are void: a failure is reported (stderr, bc_last_error(), bc_void_status())
 * and the output buffer is left untouched — never garbage, never a silent CPU fallback. */
static void die_on(int rc, const char *fn)
{
    if (rc != BC_OK) bc_void_call_failed(rc, fn);
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

/* KEY_256_ASSIST_1 / KEY_256_ASSIST_2 (aesni.c:13-36): the two helper steps of the reference's
 * AES-256 expansion.  Not declared in aesni.h but exported by aesni.o, so they are part of the
 * symbol set a relinked program may reference.  Operands are 16-byte values as four
 * little-endian words (what an __m128i is in memory); plain C, no intrinsics:
 *   _1: temp2 <- word 3 of temp2 in all four lanes; temp1 <- prefix-XOR of its words, ^ temp2
 *   _2: t <- SubWord(word 3 of temp1) in all four lanes (AESKEYGENASSIST rcon 0, lane 2);
 *       temp3 <- prefix-XOR of its words, ^ t */
uint32_t bc_aes_sub_word(uint32_t w);                                                       /* aes_dropin.c */

static void ld4(const void *p, uint32_t w[4])
{
    const unsigned char *b = (const unsigned char *)p;
    for (int i = 0; i < 4; i++)
        w[i] = (uint32_t)b[4 * i] | ((uint32_t)b[4 * i + 1] << 8) | ((uint32_t)b[4 * i + 2] << 16) | ((uint32_t)b[4 * i + 3] << 24);
}

static void st4(void *p, const uint32_t w[4])
{
    unsigned char *b = (unsigned char *)p;
    for (int i = 0; i < 4; i++) {
        b[4 * i] = (unsigned char)w[i]; b[4 * i + 1] = (unsigned char)(w[i] >> 8);
        b[4 * i + 2] = (unsigned char)(w[i] >> 16); b[4 * i + 3] = (unsigned char)(w[i] >> 24);
    }
}

static void prefix_xor_then(uint32_t w[4], uint32_t t)
{
    w[1] ^= w[0]; w[2] ^= w[1]; w[3] ^= w[2];
    for (int i = 0; i < 4; i++) w[i] ^= t;
}

void KEY_256_ASSIST_1(void *temp1, void *temp2)
{
    uint32_t a[4], b[4];
    ld4(temp1, a); ld4(temp2, b);
    b[0] = b[1] = b[2] = b[3];                       /* _mm_shuffle_epi32(*temp2, 0xff), aesni.c:15 */
    prefix_xor_then(a, b[3]);
    st4(temp1, a); st4(temp2, b);
}

void KEY_256_ASSIST_2(void *temp1, void *temp3)
{
    uint32_t a[4], c[4];
    ld4(temp1, a); ld4(temp3, c);
    prefix_xor_then(c, bc_aes_sub_word(a[3]));       /* aesni.c:27-28: keygenassist lane 2 = SubWord(X3) */
    st4(temp3, c);
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
    if (!unpack(key, number_of_rounds, rk)) { bc_set_error_str("number_of_rounds must be 10, 12 or 14"); die_on(BC_ERR_BAD_ARG, "AES_ECB_encrypt"); return; }
    die_on(bc_aes_ecb_host(rk, number_of_rounds, BC_ENCRYPT, in, out, round_up_blocks(length)), "AES_ECB_encrypt");
}

void AES_ECB_decrypt(const unsigned char *in, unsigned char *out, unsigned long length,
                     const char *key, int number_of_rounds)
{
    uint32_t rk[60];
    if (!unpack(key, number_of_rounds, rk)) { bc_set_error_str("number_of_rounds must be 10, 12 or 14"); die_on(BC_ERR_BAD_ARG, "AES_ECB_decrypt"); return; }
    die_on(bc_aes_ecb_host(rk, number_of_rounds, BC_DECRYPT, in, out, round_up_blocks(length)), "AES_ECB_decrypt");
}

void AES_CTR_encrypt(const unsigned char *in, unsigned char *out, const unsigned char ivec[8],
                     const unsigned char nonce[4], unsigned long length,
                     const unsigned char *key, int number_of_rounds)
{
    uint32_t rk[60];
    if (!unpack(key, number_of_rounds, rk)) { bc_set_error_str("number_of_rounds must be 10, 12 or 14"); die_on(BC_ERR_BAD_ARG, "AES_CTR_encrypt"); return; }
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
