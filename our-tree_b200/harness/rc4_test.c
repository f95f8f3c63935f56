/*
 * rc4_test.c — counterpart of the reference's test.c (RC4 benchmark) for the
 * B200 engine.  The reference generates ONE keystream serially (arc4_setup +
 * arc4_prep, timed as "Generated a new key in", test.c:84-91) and threads only
 * the XOR (test.c:44-58); a single RC4 stream cannot be parallelised, so the
 * GPU figure that matters is the batched one: many independent keys, KSA +
 * PRGA + XOR fused (BASELINE config 4).  Both are here, with verification.
 *
 *   rc4_test [--iters K] [--streams N] [--length L] [--single-bytes B]
 */
#include <pthread.h>

#include "arc4.h"
#include "rc4.h"
#include "bulkcipher.h"
#include "harness_util.h"

#define KEY_LENGTH_BYTES 16                 /* test.c:12 */
#define MAX_ITERS 64

int main(int argc, char *argv[])
{
    int iters = 10, ok = 1;
    size_t n_streams = 4096, length = 4096, single = 1 << 20;
    for (int i = 1; i < argc; i++) {
        if (!strcmp(argv[i], "--iters") && i + 1 < argc) iters = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--streams") && i + 1 < argc) n_streams = strtoull(argv[++i], NULL, 10);
        else if (!strcmp(argv[i], "--length") && i + 1 < argc) length = strtoull(argv[++i], NULL, 10);
        else if (!strcmp(argv[i], "--single-bytes") && i + 1 < argc) single = strtoull(argv[++i], NULL, 10);
        else { fprintf(stderr, "usage: %s [--iters K] [--streams N] [--length L] [--single-bytes B]\n", argv[0]); return 1; }
    }
    if (iters < 1 || iters > MAX_ITERS) return 1;
    if (bc_device_count() < 1) { fprintf(stderr, "no usable CUDA device: %s\n", bc_last_error()); return 2; }
    srand(1337);                            /* test.c:131 */
    double us[MAX_ITERS];

    printf("# B200 bulk-cipher RC4 harness, %d iterations\n", iters);
    printf("# ARC4 self test (arc4.c:148-183 on the GPU): %s\n", arc4_self_test(0) == 0 ? "passed" : "FAILED");

    /* ---- the reference's own shape: one key, one stream (test.c:60-126) ---- */
    unsigned char *msg = bc_host_alloc(single), *ks = bc_host_alloc(single), *out = bc_host_alloc(single);
    unsigned char key[KEY_LENGTH_BYTES];
    if (!msg || !ks || !out) { fprintf(stderr, "host alloc failed: %s\n", bc_last_error()); return 2; }
    for (size_t i = 0; i < single; i++) msg[i] = (unsigned char)(rand() % 255);
    for (int i = 0; i < KEY_LENGTH_BYTES; i++) key[i] = (unsigned char)(rand() % 255);
    arc4_context ctx;
    double t0 = now_us();
    arc4_setup(&ctx, key, KEY_LENGTH_BYTES);
    arc4_prep(&ctx, single, ks);
    printf("\nGenerated a new key in %lld, \n", (long long)(now_us() - t0));          /* test.c:84-91 */
    for (int it = 0; it < iters; it++) {
        t0 = now_us();
        arc4_crypt(single, msg, ks, out);
        us[it] = now_us() - t0;
    }
    print_csv_line("RC4", single, 1, us, iters);
    struct rc4_state st;                                                              /* rc4.h face, same bytes */
    unsigned char *out2 = bc_host_alloc(single);
    rc4_init(&st, (const char *)key, KEY_LENGTH_BYTES);
    rc4_crypt(&st, (const char *)msg, (char *)out2, (int)single);
    ok &= !memcmp(out, out2, single);
    printf("  verify %-34s %s\n", "arc4 setup/prep/crypt == rc4_crypt", ok ? "OK" : "MISMATCH");

    /* ---- batched: n_streams independent keys x length bytes (BASELINE config 4 shape) ---- */
    size_t total = n_streams * length;
    unsigned char *bm = bc_host_alloc(total), *bo = bc_host_alloc(total), *keys = bc_host_alloc(n_streams * KEY_LENGTH_BYTES);
    if (!bm || !bo || !keys) { fprintf(stderr, "host alloc failed: %s\n", bc_last_error()); return 2; }
    srand(1337);
    for (size_t i = 0; i < total; i++) bm[i] = (unsigned char)(rand() % 255);
    memcpy(keys, bm, n_streams * KEY_LENGTH_BYTES);            /* stream s key = msg[16s..16s+15] */
    for (int it = 0; it < iters; it++) {
        t0 = now_us();
        if (bc_arc4_batch_host(n_streams, keys, KEY_LENGTH_BYTES, length, bm, bo, NULL, NULL)) { fprintf(stderr, "%s\n", bc_last_error()); return 2; }
        us[it] = now_us() - t0;
    }
    print_csv_line("RC4 batched (host buffers)", total, (int)n_streams, us, iters);
    if (n_streams == 4096 && length == 4096)
        ok &= check_digest("4096 streams x 4 KiB", bo, total, "109f647611066c1a8cab28e08773d3feb2a02a151e4ecf5972825473596410b5");
    /* stream 0 through the single-stream entry points must give the same bytes */
    arc4_setup(&ctx, keys, KEY_LENGTH_BYTES);
    arc4_prep(&ctx, length, ks);
    arc4_crypt(length, bm, ks, out);
    int same = !memcmp(out, bo, length);
    printf("  verify %-34s %s\n", "batched stream 0 == single stream", same ? "OK" : "MISMATCH");
    ok &= same;

    void *d_in = bc_dev_alloc(total), *d_out = bc_dev_alloc(total), *d_keys = bc_dev_alloc(n_streams * KEY_LENGTH_BYTES);
    bc_memcpy_h2d(d_in, bm, total);
    bc_memcpy_h2d(d_keys, keys, n_streams * KEY_LENGTH_BYTES);
    for (int it = 0; it < iters + 1; it++) {
        t0 = now_us();
        if (bc_arc4_batch_dev(n_streams, d_keys, KEY_LENGTH_BYTES, length, d_in, d_out, NULL, NULL, NULL) || bc_device_sync()) { fprintf(stderr, "%s\n", bc_last_error()); return 2; }
        if (it) us[it - 1] = now_us() - t0;
    }
    print_csv_line("RC4 batched (device resident)", total, (int)n_streams, us, iters);
    bc_memcpy_d2h(bm, d_out, total);
    same = !memcmp(bm, bo, total);
    printf("  verify %-34s %s\n", "device path == host path", same ? "OK" : "MISMATCH");
    ok &= same;
    printf("# %s\n", ok ? "all outputs verified" : "VERIFICATION FAILED");
    return ok ? 0 : 3;
}
