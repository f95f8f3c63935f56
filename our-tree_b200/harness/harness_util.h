/* harness_util.h — timing, CSV printing and a compact SHA-256 for the two
 * harnesses.  The harnesses are the counterparts of the reference's test.c and
 * aes-modes/test.c: same input generator (srand(1337); rand()%255), same
 * "<Mode>, <bytes>, <threads>, t1, t2, ..." microsecond CSV lines
 * (aes-modes/test.c:47,95; test.c:61), plus what the reference lacks: GB/s,
 * and verification of every output. */
#ifndef HARNESS_UTIL_H
#define HARNESS_UTIL_H

#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

static inline double now_us(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3;
}

static int cmp_double(const void *a, const void *b)
{
    double x = *(const double *)a, y = *(const double *)b;
    return (x > y) - (x < y);
}

/* prints the reference-format line and returns the median in microseconds */
static inline double print_csv_line(const char *mode, size_t bytes, int threads, double *us, int n)
{
    printf("%s, %zu, %d, ", mode, bytes, threads);
    for (int i = 0; i < n; i++) printf("%lld, ", (long long)us[i]);
    qsort(us, (size_t)n, sizeof us[0], cmp_double);
    double med = n ? us[n / 2] : 0.0;
    printf("  # median %.0f us = %.3f GB/s\n", med, med > 0 ? bytes / med / 1e3 : 0.0);
    fflush(stdout);
    return med;
}

/* ---- SHA-256 (FIPS 180-4), used to compare outputs with the committed digests ---- */
typedef struct { uint32_t h[8]; uint64_t len; unsigned char buf[64]; size_t fill; } sha256_t;

static const uint32_t SHA_K[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, 0xd807aa98, 0x12835b01,
    0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174, 0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc,
    0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da, 0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147,
    0x06ca6351, 0x14292967, 0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
    0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070, 0x19a4c116, 0x1e376c08,
    0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3, 0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208,
    0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2 };

static inline uint32_t ror32(uint32_t v, int n) { return (v >> n) | (v << (32 - n)); }

static void sha256_block(sha256_t *s, const unsigned char *p)
{
    uint32_t w[64], a[8];
    for (int i = 0; i < 16; i++) w[i] = ((uint32_t)p[4 * i] << 24) | ((uint32_t)p[4 * i + 1] << 16) | ((uint32_t)p[4 * i + 2] << 8) | p[4 * i + 3];
    for (int i = 16; i < 64; i++)
        w[i] = w[i - 16] + (ror32(w[i - 15], 7) ^ ror32(w[i - 15], 18) ^ (w[i - 15] >> 3)) + w[i - 7] +
               (ror32(w[i - 2], 17) ^ ror32(w[i - 2], 19) ^ (w[i - 2] >> 10));
    memcpy(a, s->h, sizeof a);
    for (int i = 0; i < 64; i++) {
        uint32_t t1 = a[7] + (ror32(a[4], 6) ^ ror32(a[4], 11) ^ ror32(a[4], 25)) + ((a[4] & a[5]) ^ (~a[4] & a[6])) + SHA_K[i] + w[i];
        uint32_t t2 = (ror32(a[0], 2) ^ ror32(a[0], 13) ^ ror32(a[0], 22)) + ((a[0] & a[1]) ^ (a[0] & a[2]) ^ (a[1] & a[2]));
        memmove(a + 1, a, 7 * sizeof a[0]);
        a[4] += t1;
        a[0] = t1 + t2;
    }
    for (int i = 0; i < 8; i++) s->h[i] += a[i];
}

static void sha256_hex(const unsigned char *data, size_t n, char out[65])
{
    sha256_t s = { { 0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19 }, 0, { 0 }, 0 };
    size_t off = 0;
    for (; off + 64 <= n; off += 64) sha256_block(&s, data + off);
    unsigned char tail[128] = { 0 };
    size_t r = n - off;
    memcpy(tail, data + off, r);
    tail[r] = 0x80;
    size_t tl = (r + 9 <= 64) ? 64 : 128;
    uint64_t bits = (uint64_t)n * 8;
    for (int i = 0; i < 8; i++) tail[tl - 1 - i] = (unsigned char)(bits >> (8 * i));
    sha256_block(&s, tail);
    if (tl == 128) sha256_block(&s, tail + 64);
    for (int i = 0; i < 8; i++) sprintf(out + 8 * i, "%08x", s.h[i]);
}

static int check_digest(const char *what, const unsigned char *data, size_t n, const char *want)
{
    char got[65];
    sha256_hex(data, n, got);
    int ok = strcmp(got, want) == 0;
    printf("  verify %-34s %s\n", what, ok ? "OK" : "MISMATCH");
    if (!ok) printf("    got  %s\n    want %s\n", got, want);
    return ok;
}

#endif
