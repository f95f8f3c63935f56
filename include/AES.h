/*
 * AES.h — drop-in for the reference's aes-gpu/Source/AES.h (:84-147): the C++
 * facade `class AES : public BlockCipher`, here a thin header-only adapter over
 * the B200 engine (libbulkcipher.so) so that main_ecb_e.cu / main_ecb_d.cu
 * compile as plain C++ and produce correct results.  NOT a port of AES.cu: the
 * reference kernels keep their state in block-shared scalars and ask for the
 * whole buffer as dynamic shared memory (AES.cu:250,290; SURVEY finding 7).
 *
 * Same public interface and word convention: a block is uint[4], each word
 * big-endian (GETWORD / PUTWORD, AES.cu:42-43); makeKey accepts the key size in
 * bytes or bits and throws a C string on anything else (AES.cu:205-218);
 * encrypt / decrypt take HOST word arrays of n blocks.  Unlike the reference,
 * GPU errors are not ignored: they throw the engine's message.
 */
#ifndef BULKCIPHER_DROPIN_AES_GPU_H
#define BULKCIPHER_DROPIN_AES_GPU_H

#include <string.h>

#include "BlockCipher.h"
#include "bulkcipher.h"

#define AES_BLOCKBITS   128
#define AES_BLOCKSIZE   16
#define AES_MAXKEYBITS  256
#define AES_MAXKEYSIZE  32
#define MAXKC           (AES_MAXKEYBITS / 32)
#define MAXKB           (AES_MAXKEYBITS / 8)
#define MAXNR           14

class AES : public BlockCipher {
public:
    AES() : Nr(0) { memset(e_sched, 0, sizeof e_sched); memset(d_sched, 0, sizeof d_sched); }
    virtual ~AES() { memset(e_sched, 0, sizeof e_sched); memset(d_sched, 0, sizeof d_sched); }

    inline uint blockBits() const { return AES_BLOCKBITS; }
    inline uint blockSize() const { return AES_BLOCKSIZE; }
    inline uint keyBits() const { return (Nr - 6) << 5; }
    inline uint keySize() const { return (Nr - 6) << 2; }

    /* AES.cu:190-203 */
    void byte2int(const byte *b, uint *i) {
        for (int k = 0; k < 4; k++)
            i[k] = ((uint)b[4 * k] << 24) | ((uint)b[4 * k + 1] << 16) | ((uint)b[4 * k + 2] << 8) | (uint)b[4 * k + 3];
    }
    void int2byte(const uint *i, byte *b) {
        for (int k = 0; k < 4; k++) {
            b[4 * k] = (byte)(i[k] >> 24); b[4 * k + 1] = (byte)(i[k] >> 16);
            b[4 * k + 2] = (byte)(i[k] >> 8); b[4 * k + 3] = (byte)i[k];
        }
    }

    /* AES.cu:205-228 */
    void makeKey(const byte *cipherKey, uint keySize, uint dir) {
        switch (keySize) {
        case 16: case 24: case 32: keySize <<= 3; break;
        case 128: case 192: case 256: break;
        default: throw "Invalid AES key size";
        }
        if (dir > DIR_BOTH) throw "Invalid direction";
        if (dir == DIR_NONE) return;
        Nr = (uint)bc_aes_expand_key(cipherKey, keySize, e_sched);
        if (dir & DIR_DECRYPT) bc_aes_invert_key(e_sched, (int)Nr, d_sched);
    }

    /* AES.cu:230-282 — n blocks of 4 big-endian words each, host memory */
    void encrypt(const uint *pt, uint *ct, uint n = 1) {
        if (bc_aes_ecb_be32_host(e_sched, (int)Nr, BC_ENCRYPT, pt, ct, (size_t)n * 16)) throw bc_last_error();
    }
    void decrypt(const uint *ct, uint *pt, uint n = 1) {
        if (bc_aes_ecb_be32_host(d_sched, (int)Nr, BC_DECRYPT, ct, pt, (size_t)n * 16)) throw bc_last_error();
    }

private:
    uint Nr;
    uint32_t e_sched[60];
    uint32_t d_sched[60];
};

#endif
