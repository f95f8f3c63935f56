/*
 * aes_gpu_facade_test.cpp — the call sequences of the reference's
 * aes-gpu/Source/main_ecb_e.cu (:12-52) and main_ecb_d.cu (:13-40) against the
 * drop-in `class AES : BlockCipher` (include/AES.h), with the verification the
 * reference never does: FIPS-197 vectors through byte2int / encrypt / int2byte,
 * decrypt(encrypt(x)) == x on a random buffer, and word interface == byte interface.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>

#include "AES.h"

static int check_fips(uint keySize, const char *expect_hex)
{
    byte key[32], pt[16], ct[16], back[16];
    uint w[4], o[4];
    for (int i = 0; i < 32; i++) key[i] = (byte)i;
    for (int i = 0; i < 16; i++) pt[i] = (byte)(i * 0x11);
    AES aes;
    aes.makeKey(key, keySize, DIR_BOTH);                 /* bytes or bits, AES.cu:206-214 */
    aes.byte2int(pt, w);
    aes.encrypt(w, o, 1);
    aes.int2byte(o, ct);
    char hex[33];
    for (int i = 0; i < 16; i++) sprintf(hex + 2 * i, "%02x", ct[i]);
    aes.decrypt(o, w, 1);
    aes.int2byte(w, back);
    int ok = !strcmp(hex, expect_hex) && !memcmp(back, pt, 16) && aes.keyBits() == (keySize <= 32 ? keySize * 8 : keySize);
    printf("  FIPS-197 AES-%u via class AES: %s\n", aes.keyBits(), ok ? "OK" : "MISMATCH");
    return ok;
}

int main()
{
    int ok = 1;
    try {
        ok &= check_fips(16, "69c4e0d86a7b0430d8cdb78070b4c55a");
        ok &= check_fips(192, "dda97ca4864cdfe06eaf70a0ec0d7191");
        ok &= check_fips(32, "8ea2b7ca516745bfeafc49904b496089");

        /* main_ecb_e.cu's timed region: makeKey + encrypt of a rand() buffer */
        const uint ptSize = 10485760 / sizeof(uint);     /* words; the reference's second size */
        byte key[32];
        for (int i = 0; i < 32; i++) key[i] = (byte)rand();
        uint *pt = (uint *)malloc(ptSize * sizeof(uint)), *ct = (uint *)malloc(ptSize * sizeof(uint)), *back = (uint *)malloc(ptSize * sizeof(uint));
        for (uint i = 0; i < ptSize; i++) pt[i] = (uint)rand();
        AES *aes = new AES();
        struct timeval start, end;
        gettimeofday(&start, NULL);
        aes->makeKey(key, 32 << 3, DIR_ENCRYPT);
        aes->encrypt(pt, ct, ptSize >> 2);
        gettimeofday(&end, NULL);
        long long t = (long long)(end.tv_sec - start.tv_sec) * 1000000 + (end.tv_usec - start.tv_usec);
        printf("AES ECB test, %u: %lld,  (microseconds, host buffers)\n", (unsigned)(ptSize * sizeof(uint)), t);
        aes->makeKey(key, 256, DIR_BOTH);
        aes->decrypt(ct, back, ptSize >> 2);
        int rt = !memcmp(back, pt, ptSize * sizeof(uint));
        printf("  decrypt(encrypt(x)) == x on %u words: %s\n", ptSize, rt ? "OK" : "MISMATCH");
        ok &= rt;
        /* word interface == byte-stream interface on the first block */
        byte b[16], e[16];
        uint32_t rk[60];
        aes->int2byte(pt, b);
        int nr = bc_aes_expand_key(key, 256, rk);
        if (bc_aes_ecb_host(rk, nr, BC_ENCRYPT, b, e, 16)) throw bc_last_error();
        uint w[4];
        aes->byte2int(e, w);
        int same = !memcmp(w, ct, 16);
        printf("  uint[] interface == byte interface: %s\n", same ? "OK" : "MISMATCH");
        ok &= same;
        try { aes->makeKey(key, 17, DIR_ENCRYPT); ok = 0; } catch (const char *msg) { printf("  bad key size throws: \"%s\"\n", msg); }
        delete aes;
    } catch (const char *msg) {
        fprintf(stderr, "error: %s\n", msg);
        return 2;
    }
    printf("# %s\n", ok ? "all outputs verified" : "VERIFICATION FAILED");
    return ok ? 0 : 3;
}
