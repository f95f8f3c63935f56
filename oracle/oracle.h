/*
 * oracle.h — CPU restatement of the reference's bulk-cipher algorithms.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs may load it, and only as the checker.  The product
 * library (libbulkcipher.so) never links, loads or calls it.
 *
 * Parity status: PINNED.  The restatement is checked (tests/test_oracle.py)
 * against (1) the reference's own known-answer vectors — aes.c:912-950
 * (Rijndael Monte-Carlo, ECB+CBC, 128/192/256, enc+dec), aes.c:1025-1078
 * (RFC 3686 CTR #1-#3), arc4.c:124-143 (3 RC4 vectors); (2) FIPS-197 App. C and
 * SP 800-38A F.1/F.2/F.5 vectors; (3) the reference objects themselves
 * (oracle/_ref/libref.so, compiled from /root/reference by oracle/Makefile)
 * on seeded random inputs; (4) the SHA-256 golden digests in tests/golden/.
 *
 * All file:line citations are relative to the reference tree.
 */
#ifndef ORACLE_H
#define ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_ENCRYPT 1   /* aes-modes/aes.h:33 AES_ENCRYPT */
#define ORC_DECRYPT 0   /* aes-modes/aes.h:34 AES_DECRYPT */
#define ORC_ERR_KEYLEN   (-0x0020)  /* aes-modes/aes.h:35 */
#define ORC_ERR_INPUTLEN (-0x0022)  /* aes-modes/aes.h:36 */

/* Round-key schedule as 32-bit little-endian column words (the low halves of
 * aes_context.rk[], aes-modes/aes.h:41-47). */
typedef struct {
    int      nr;        /* 10 / 12 / 14 */
    uint32_t rk[60];    /* 4*(nr+1) words used */
} orc_aes_key;

int  orc_aes_setkey_enc(orc_aes_key *k, const uint8_t *key, unsigned keybits);
int  orc_aes_setkey_dec(orc_aes_key *k, const uint8_t *key, unsigned keybits);
void orc_aes_ecb_block(const orc_aes_key *k, int mode, const uint8_t in[16], uint8_t out[16]);
/* bulk ECB over whole blocks (what the harness loop aes-modes/test.c:37-41 does) */
void orc_aes_ecb(const orc_aes_key *k, int mode, size_t length, const uint8_t *in, uint8_t *out);
int  orc_aes_cbc(const orc_aes_key *k, int mode, size_t length, uint8_t iv[16],
                 const uint8_t *in, uint8_t *out);
int  orc_aes_ctr(const orc_aes_key *k, size_t length, int *nc_off, uint8_t nonce_counter[16],
                 uint8_t stream_block[16], const uint8_t *in, uint8_t *out);

int  orc_aes_cfb128(const orc_aes_key *k, int mode, size_t length, int *iv_off, uint8_t iv[16],
                    const uint8_t *in, uint8_t *out);

/* aesni.h-style calls on a packed schedule (16 bytes per round key). */
void orc_aesni_key256(const uint8_t *userkey, uint8_t *sched /*240 B*/);
void orc_aesni_pack(const orc_aes_key *k, uint8_t *sched /*16*(nr+1) B*/);
void orc_aesni_ecb_encrypt(const uint8_t *in, uint8_t *out, unsigned long length,
                           const uint8_t *sched, int rounds);
void orc_aesni_ecb_decrypt(const uint8_t *in, uint8_t *out, unsigned long length,
                           const uint8_t *sched, int rounds);
void orc_aesni_ctr_encrypt(const uint8_t *in, uint8_t *out, const uint8_t ivec[8],
                           const uint8_t nonce[4], unsigned long length,
                           const uint8_t *sched, int rounds);

/* arc4.h */
typedef struct { int x, y; uint8_t m[256]; } orc_arc4_ctx;     /* arc4.h:35-41 */
void orc_arc4_setup(orc_arc4_ctx *c, const uint8_t *key, unsigned keylen);
int  orc_arc4_prep (orc_arc4_ctx *c, size_t length, uint8_t *keystream);
int  orc_arc4_crypt(size_t length, const uint8_t *in, const uint8_t *keystream, uint8_t *out);

/* rc4.h — intended semantics (upstream FreeBSD u_char types, rc4.c:36); the
 * vendored rc4.c is undefined behaviour as shipped (SURVEY finding 4). */
typedef struct { uint8_t perm[256]; int index1, index2; } orc_rc4_state;  /* rc4.h:43-47 */
void orc_rc4_init (orc_rc4_state *s, const uint8_t *key, int keylen);
void orc_rc4_crypt(orc_rc4_state *s, const uint8_t *in, uint8_t *out, int buflen);

/* batched conveniences used by the parity tests and the CPU baseline */
void orc_arc4_batch(size_t n_streams, const uint8_t *keys, unsigned keylen,
                    size_t length, const uint8_t *in, uint8_t *out);
void orc_aes_cbc_encrypt_batch(const orc_aes_key *k, size_t n_msgs, size_t msg_len,
                               uint8_t *ivs /* [n][16], updated */, const uint8_t *in, uint8_t *out);

/* reference harness input: srand(seed); msg[i]=rand()%255; key[i]=rand()%255 */
void orc_harness_input(unsigned seed, uint8_t *msg, size_t n_msg, uint8_t *key, size_t n_key);

#ifdef __cplusplus
}
#endif
#endif
