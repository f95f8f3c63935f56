/*
 * call_latency.c — what one small host-buffer call costs, and where: AES-128-CTR through
 * bc_aes_ctr_host on pageable (malloc) and pinned (bc_host_alloc) buffers, called from the main
 * thread and — the way the reference harness calls the API — from a freshly created pthread per
 * call (aes-modes/test.c:316-327 creates and joins its threads inside every timed iteration).
 * Prints the median of 20 calls in microseconds for a few sizes.
 */
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "aes.h"
#include "bulkcipher.h"

static uint32_t rk[60];
static int nr;
static unsigned char counter[16];

typedef struct { const unsigned char *in; unsigned char *out; size_t n; int rc; } job;

static void *one_call(void *p)
{
    job *j = (job *)p;
    j->rc = bc_aes_ctr_host(rk, nr, counter, 128, 0, j->in, j->out, j->n, NULL);
    return NULL;
}

static double now_us(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3;
}

typedef struct { aes_context *ctx; const unsigned char *in; unsigned char *out; size_t n; double us; } ecb_job;

static void *ecb_plain(void *p)
{
    ecb_job *j = (ecb_job *)p;
    const double t0 = now_us();
    for (size_t off = 0; off < j->n; off += 16) aes_crypt_ecb(j->ctx, AES_ENCRYPT, j->in + off, j->out + off);
    j->us = now_us() - t0;
    return NULL;
}

static int cmp(const void *a, const void *b) { double x = *(const double *)a, y = *(const double *)b; return x < y ? -1 : x > y; }

static double median_us(job *j, int fresh_thread)
{
    double t[20];
    for (int i = 0; i < 22; i++) {
        const double t0 = now_us();
        if (fresh_thread) {
            pthread_t th;
            pthread_create(&th, NULL, one_call, j);
            pthread_join(th, NULL);
        } else {
            one_call(j);
        }
        if (j->rc) { fprintf(stderr, "call failed: %s\n", bc_last_error()); exit(1); }
        if (i >= 2) t[i - 2] = now_us() - t0;
    }
    qsort(t, 20, sizeof t[0], cmp);
    return t[10];
}

int main(void)
{
    unsigned char key[16];
    for (int i = 0; i < 16; i++) key[i] = (unsigned char)i;
    nr = bc_aes_expand_key(key, 128, rk);
    const size_t sizes[] = { 4096, 65536, 1 << 20, 10 << 20, 100 << 20 };
    printf("bc_aes_ctr_host, median of 20 calls, microseconds\n");
    printf("| bytes | pageable, main thread | pageable, new thread per call | pinned, main thread | pinned, new thread per call |\n|---|---|---|---|---|\n");
    for (unsigned s = 0; s < sizeof sizes / sizeof sizes[0]; s++) {
        const size_t n = sizes[s];
        unsigned char *a = malloc(n), *b = malloc(n);
        unsigned char *pa = bc_host_alloc(n), *pb = bc_host_alloc(n);
        if (!a || !b || !pa || !pb) { fprintf(stderr, "allocation failed: %s\n", bc_last_error()); return 1; }
        memset(a, 3, n); memset(b, 0, n); memset(pa, 3, n); memset(pb, 0, n);
        job jp = { a, b, n, 0 }, jq = { pa, pb, n, 0 };
        const double m1 = median_us(&jp, 0), m2 = median_us(&jp, 1), m3 = median_us(&jq, 0), m4 = median_us(&jq, 1);
        if (memcmp(b, pb, n)) { fprintf(stderr, "pageable and pinned outputs differ\n"); return 1; }
        printf("| %zu | %.0f | %.0f | %.0f | %.0f |\n", n, m1, m2, m3, m4);
        free(a); free(b); bc_host_free(pa); bc_host_free(pb);
    }

    /* device-resident buffers from C (no Python, no CUDA events): one call + sync, and 200 calls back to back */
    {
        const size_t dsizes[] = { 4096, 65536, 1 << 20, 4 << 20, 16 << 20, 64 << 20 };
        void *d_in = bc_dev_alloc(64 << 20), *d_out = bc_dev_alloc(64 << 20);
        if (!d_in || !d_out) { fprintf(stderr, "bc_dev_alloc failed: %s\n", bc_last_error()); return 1; }
        printf("\nbc_aes_ctr_dev on device-resident buffers, called from C\n| bytes | one call + sync, us | per call, 200 back to back, us | GB/s back to back |\n|---|---|---|---|\n");
        for (unsigned s = 0; s < sizeof dsizes / sizeof dsizes[0]; s++) {
            const size_t n = dsizes[s];
            double t[20];
            for (int i = 0; i < 22; i++) {
                const double t0 = now_us();
                if (bc_aes_ctr_dev(rk, nr, counter, 128, 0, d_in, d_out, n, NULL, NULL) || bc_device_sync()) { fprintf(stderr, "%s\n", bc_last_error()); return 1; }
                if (i >= 2) t[i - 2] = now_us() - t0;
            }
            qsort(t, 20, sizeof t[0], cmp);
            const double t0 = now_us();
            for (int i = 0; i < 200; i++) bc_aes_ctr_dev(rk, nr, counter, 128, 0, d_in, d_out, n, NULL, NULL);
            bc_device_sync();
            const double per = (now_us() - t0) / 200;
            printf("| %zu | %.1f | %.2f | %.1f |\n", n, t[10], per, n / per / 1e3);
        }
        bc_dev_free(d_in); bc_dev_free(d_out);
    }

    /* the reference's "Plain ECB" loop, one aes_crypt_ecb call per 16 bytes (aes-modes/test.c:37-41) */
    {
        const size_t n = 16 << 20;
        unsigned char *a = malloc(n), *b = malloc(n), *c = malloc(n);
        aes_context ctx;
        for (size_t i = 0; i < n; i++) a[i] = (unsigned char)(i * 131 + (i >> 11));
        aes_setkey_enc(&ctx, key, 128);
        const double t0 = now_us();
        for (size_t off = 0; off < n; off += 16)
            if (aes_crypt_ecb(&ctx, AES_ENCRYPT, a + off, b + off)) { fprintf(stderr, "aes_crypt_ecb: %s\n", bc_last_error()); return 1; }
        const double dt = now_us() - t0;
        if (bc_aes_ecb_host(rk, nr, BC_ENCRYPT, a, c, n) || memcmp(b, c, n)) { fprintf(stderr, "block loop differs from the bulk call\n"); return 1; }
        printf("\naes_crypt_ecb block loop over %zu MiB (read-ahead window): %.1f ms = %.0f MB/s, equal to the bulk call\n",
               n >> 20, dt / 1e3, n / dt);
        /* the same loop with the window off (the switch is read when a thread makes its first call) */
        setenv("BULKCIPHER_ECB_READAHEAD", "0", 1);
        ecb_job ej = { &ctx, a, b, 4000 * 16, 0 };
        pthread_t th;
        pthread_create(&th, NULL, ecb_plain, &ej);
        pthread_join(th, NULL);
        printf("aes_crypt_ecb block loop, one GPU round trip per block (BULKCIPHER_ECB_READAHEAD=0): %.1f us per block = %.2f MB/s\n",
               ej.us / 4000, 4000 * 16 / ej.us);
        free(a); free(b); free(c);
    }
    return 0;
}
