/*
 * BlockCipher.h — drop-in for the reference's aes-gpu/Source/BlockCipher.h
 * (:48-107): the abstract block-cipher interface its C++ class AES implements.
 * Blocks travel as arrays of 32-bit words (big-endian within a word, see
 * byte2int / int2byte).
 *
 * One deliberate fix: the reference typedefs `uint` as `unsigned long`
 * "assuming sizeof(uint) == 4" (AES.h:31), which is 8 bytes on LP64 Linux and
 * breaks its own size arithmetic (SURVEY finding 7).  Here `uint` IS 32 bits.
 */
#ifndef BULKCIPHER_DROPIN_BLOCKCIPHER_H
#define BULKCIPHER_DROPIN_BLOCKCIPHER_H

#include <stdint.h>

#ifndef USUAL_TYPES
#define USUAL_TYPES
typedef unsigned char byte;
typedef uint32_t      uint;
#endif

#define DIR_NONE    0
#define DIR_ENCRYPT 1
#define DIR_DECRYPT 2
#define DIR_BOTH    (DIR_ENCRYPT | DIR_DECRYPT)

class BlockCipher {
public:
    virtual ~BlockCipher() {}
    virtual uint blockBits() const = 0;
    virtual uint blockSize() const = 0;
    virtual uint keyBits() const = 0;
    virtual uint keySize() const = 0;
    /* one block, byte[16] <-> uint[4] */
    virtual void byte2int(const byte *b, uint *i) = 0;
    virtual void int2byte(const uint *i, byte *b) = 0;
    /* key schedule for encryption, decryption or both */
    virtual void makeKey(const byte *cipherKey, uint keyBits, uint dir) = 0;
    /* n blocks of 4 words each */
    virtual void encrypt(const uint *pt, uint *ct, uint n) = 0;
    virtual void decrypt(const uint *ct, uint *pt, uint n) = 0;
};

#endif
