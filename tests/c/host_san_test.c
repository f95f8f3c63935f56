/*
 * host_san_test.c — the C host layer under AddressSanitizer + UBSan, no GPU needed.
 * Links the host C sources with stubs for the CUDA launchers and checks what runs on the host:
 * key schedules against FIPS-197 (appendix A.1/A.3 round keys), argument checks and error
 * codes, the "no device => fail, never emulate" rule, and the parallel staging copy.
 * Built and run by tests/test_abi.py::test_host_layer_under_sanitizers.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>

#include "aes.h"
#include "aesni.h"
#include "arc4.h"
#include "bulkcipher.h"

/* the device half of the C-ABI is not linked here: every launcher reports a CUDA failure */
#define STUB(name, ...) int name(__VA_ARGS__) { return BC_ERR_CUDA; }
STUB(bc_aes_ctr_dev, const uint32_t *a, int b, const unsigned char *c, int d, uint64_t e, const void *f, void *g, size_t h, void *i, void *j)
STUB(bc_aes_ecb_dev, const uint32_t *a, int b, int c, const void *d, void *e, size_t f, void *g)
STUB(bc_aes_ecb_be32_dev, const uint32_t *a, int b, int c, const void *d, void *e, size_t f, void *g)
STUB(bc_aes_cbc_decrypt_dev, const uint32_t *a, int b, const unsigned char *c, const void *d, void *e, size_t f, void *g)
STUB(bc_aes_cfb128_decrypt_dev, const uint32_t *a, int b, const unsigned char *c, const void *d, void *e, size_t f, void *g)
STUB(bc_aes_cbc_encrypt_batch_dev, const uint32_t *a, int b, size_t c, size_t d, const void *e, void *f, const void *g, void *h, void *i)
STUB(bc_aes_cfb128_encrypt_batch_dev, const uint32_t *a, int b, size_t c, size_t d, const void *e, void *f, const void *g, void *h, void *i)
STUB(bc_arc4_batch_dev, size_t a, const void *b, unsigned c, size_t d, const void *e, void *f, const void *g, void *h, void *i)
STUB(bc_xor_dev, const void *a, const void *b, void *c, size_t d, void *e)

void bc_internal_par_memcpy(void *dst, const void *src, size_t n);   /* host/bc_engine.c */
size_t bc_internal_next_chunk(size_t nbytes, size_t off, size_t idx, size_t chunk, int ramp);   /* host/bc_engine.c */
size_t bc_internal_read_ahead(unsigned char *dst, const unsigned char *src, size_t want);       /* host/aes_dropin.c */

static int fails;
#define CHECK(c) do { if (!(c)) { printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #c); fails++; } } while (0)

int main(void)
{
    /* FIPS-197 A.1: last round key of the 128-bit example; A.3: last of the 256-bit example */
    static const unsigned char k128[16] = { 0x2b, 0x7e, 0x15, 0x16, 0x28, 0xae, 0xd2, 0xa6, 0xab, 0xf7, 0x15, 0x88, 0x09, 0xcf, 0x4f, 0x3c };
    static const unsigned char last128[16] = { 0xd0, 0x14, 0xf9, 0xa8, 0xc9, 0xee, 0x25, 0x89, 0xe1, 0x3f, 0x0c, 0xc8, 0xb6, 0x63, 0x0c, 0xa6 };
    static const unsigned char k256[32] = { 0x60, 0x3d, 0xeb, 0x10, 0x15, 0xca, 0x71, 0xbe, 0x2b, 0x73, 0xae, 0xf0, 0x85, 0x7d, 0x77, 0x81,
                                            0x1f, 0x35, 0x2c, 0x07, 0x3b, 0x61, 0x08, 0xd7, 0x2d, 0x98, 0x10, 0xa3, 0x09, 0x14, 0xdf, 0xf4 };
    static const unsigned char last256[16] = { 0xfe, 0x48, 0x90, 0xd1, 0xe6, 0x18, 0x8d, 0x0b, 0x04, 0x6d, 0xf3, 0x44, 0x70, 0x6c, 0x63, 0x1e };
    aes_context ctx, dctx;
    unsigned char sched[240], buf[64], iv[16] = { 0 };
    memset(&ctx, 0, sizeof ctx);
    CHECK(aes_setkey_enc(&ctx, k128, 128) == 0 && ctx.nr == 10 && ctx.rk == ctx.buf);
    AES_128_Key_Expansion(k128, sched);
    CHECK(memcmp(sched + 160, last128, 16) == 0);
    for (int i = 0; i < 44; i++) CHECK((uint32_t)ctx.rk[i] == ((uint32_t)sched[4 * i] | (uint32_t)sched[4 * i + 1] << 8 | (uint32_t)sched[4 * i + 2] << 16 | (uint32_t)sched[4 * i + 3] << 24));
    AES_256_Key_Expansion(k256, sched);
    CHECK(memcmp(sched + 224, last256, 16) == 0);
    /* the equivalent-inverse schedule starts with the last and ends with the first encryption round key */
    CHECK(aes_setkey_dec(&dctx, k128, 128) == 0);
    for (int i = 0; i < 4; i++) { CHECK(dctx.rk[i] == ctx.rk[40 + i]); CHECK(dctx.rk[40 + i] == ctx.rk[i]); }
    /* error codes, ctx untouched on a bad key size */
    aes_context z;
    memset(&z, 0x77, sizeof z);
    CHECK(aes_setkey_enc(&z, k128, 100) == POLARSSL_ERR_AES_INVALID_KEY_LENGTH && z.nr == 0x77777777);
    CHECK(aes_setkey_dec(&z, k128, 0) == POLARSSL_ERR_AES_INVALID_KEY_LENGTH);
    CHECK(aes_crypt_cbc(&ctx, AES_ENCRYPT, 17, iv, buf, buf) == POLARSSL_ERR_AES_INVALID_INPUT_LENGTH);
    CHECK(aes_crypt_cbc(&ctx, AES_ENCRYPT, 0, iv, buf, buf) == 0);
    /* no device: every compute path fails and leaves the output alone */
    memset(buf, 0xAB, sizeof buf);
    if (bc_device_count() == 0) {
        int off = 0;
        unsigned char nc[16] = { 0 }, sb[16] = { 0 };
        CHECK(aes_crypt_ecb(&ctx, AES_ENCRYPT, buf, buf + 16) != 0);
        CHECK(aes_crypt_ctr(&ctx, 32, &off, nc, sb, buf, buf + 32) != 0);
        CHECK(bc_arc4_batch_host(1, k128, 16, 16, buf, buf + 16, NULL, NULL) != 0);
        CHECK(CheckAESSupport() == 0);
        for (int i = 0; i < 64; i++) CHECK(buf[i] == 0xAB);
        CHECK(strlen(bc_last_error()) > 0);
    }
    /* parallel staging copy: odd sizes around the share boundaries, overlapping nothing */
    for (size_t n = 1; n < ((size_t)9 << 20); n = n * 3 + 4097) {
        unsigned char *src = malloc(n), *dst = malloc(n + 2);
        for (size_t i = 0; i < n; i++) src[i] = (unsigned char)(i * 131 + 7);
        dst[0] = dst[n + 1] = 0x5A;
        bc_internal_par_memcpy(dst + 1, src, n);
        CHECK(memcmp(dst + 1, src, n) == 0 && dst[0] == 0x5A && dst[n + 1] == 0x5A);
        free(src);
        free(dst);
    }
    /* the pipeline's chunk schedule: covers the transfer exactly, pages stay whole, no chunk above the granule */
    {
        const size_t M = (size_t)1 << 20, sizes[] = { 8 * 64 * M, 4096 * M, 1000 * M + 7, 8 * 64 * M + 4095, 8 * 64 * M + 4096 + 5, 513 * M, 128 * M, 16 };
        for (unsigned k = 0; k < sizeof sizes / sizeof sizes[0]; k++)
            for (int ramp = 0; ramp < 2; ramp++) {
                const size_t chunk = k == 6 ? 16 * M : 64 * M;
                size_t off = 0, idx = 0, first = 0, last = 0;
                while (off < sizes[k]) {
                    const size_t n = bc_internal_next_chunk(sizes[k], off, idx, chunk, ramp && sizes[k] >= 8 * chunk);
                    CHECK(n > 0 && n <= chunk && (off + n == sizes[k] || n % 4096 == 0));
                    if (n == 0) break;
                    if (idx == 0) first = n;
                    last = n;
                    off += n; idx++;
                }
                CHECK(off == sizes[k]);
                if (ramp && sizes[k] >= 8 * chunk) CHECK(first == chunk / 16 && last < chunk / 16 + 4096 && last >= chunk / 16);
            }
    }
    /* the read-ahead copy of aes_crypt_ecb never faults: it stops at the first unreadable page */
    {
        unsigned char *m = mmap(NULL, 3 * 4096, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
        CHECK(m != MAP_FAILED);
        for (int i = 0; i < 2 * 4096; i++) m[i] = (unsigned char)(i * 7 + 1);
        CHECK(munmap(m + 2 * 4096, 4096) == 0);
        unsigned char *dst = malloc(4 * 4096);
        const size_t got = bc_internal_read_ahead(dst, m + 100, 4 * 4096);
        CHECK(got == 0 || got == 2 * 4096 - 100);                 /* 0: process_vm_readv not permitted in this sandbox */
        if (got) CHECK(memcmp(dst, m + 100, got) == 0);
        CHECK(bc_internal_read_ahead(dst, m + 4096 - 8, 8) == 8 || got == 0);
        free(dst);
        munmap(m, 2 * 4096);
        /* a block walk without a device: every call fails, nothing is written, nothing is cached */
        if (bc_device_count() == 0) {
            unsigned char in[16 * 8] = { 1 }, out[16 * 8];
            memset(out, 0xCD, sizeof out);
            for (int b = 0; b < 8; b++) CHECK(aes_crypt_ecb(&ctx, AES_ENCRYPT, in + 16 * b, out + 16 * b) != 0);
            for (unsigned i = 0; i < sizeof out; i++) CHECK(out[i] == 0xCD);
            CHECK(bc_aes_ecb_host_multi(1, (const uint32_t *)sched, 10, BC_ENCRYPT, in, out, sizeof in) != 0);
            CHECK(bc_void_status() == 0);
            AES_ECB_encrypt(in, out, sizeof in, sched, 14);
            CHECK(bc_void_status() != 0 && out[0] == 0xCD);
        }
    }
    printf(fails ? "host_san_test: %d failure(s)\n" : "host_san_test: ok\n", fails);
    return fails ? 1 : 0;
}
