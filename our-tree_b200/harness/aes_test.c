/*
 * aes_test.c — counterpart of the reference's aes-modes/test.c for the B200
 * engine.  Same shape (random message from srand(1337)/rand()%255, fresh key
 * per iteration, N pthreads on equal contiguous slices, microsecond CSV lines)
 * but it runs the GPU path through the reference's own entry points, it checks
 * every output, and it adds the device-resident measurements.
 *
 *   aes_test [--bits 128|256] [--iters K] [--sizes a,b,...] [--threads a,b,...]
 *            [--gpus N] [--config1]
 *
 * Differences from aes-modes/test.c, on purpose (SURVEY findings 5-6):
 *  - "Plain CTR" really is CTR (the reference times ECB there, test.c:162);
 *  - CTR slices continue ONE counter stream (slice t starts at its first block
 *    index) instead of restarting the counter on every slice (test.c:273-285);
 *  - outputs are verified: decrypt round trip, slices == unsliced, and with
 *    --config1 the SHA-256 digests of the 16 MiB BASELINE config-1 case.
 */
#include <pthread.h>

#include "aes.h"
#include "aesni.h"
#include "bulkcipher.h"
#include "harness_util.h"

#define MAX_ITERS 64
static int ITERATIONS = 10;                 /* aes-modes/test.c:10 */
static int KEY_BITS = 128;

static unsigned char *msg, *out, *chk;
static size_t MESSAGE_LENGTH;
static unsigned char key[32];
static uint32_t rk_enc[60], rk_dec[60];
static int nr;

typedef struct { int thread_id, total_threads, mode; } AESInfo;   /* aes-modes/test.c:19-22 */
enum { M_ECB, M_CTR, M_CBC_DEC };
static const unsigned char NONCE_COUNTER[16] = { 'H', 'l', 'o', 'E', 'e', 'l', 'A', 'S', '3', '1', '5', 'A', 0xff, 0xff, 0xff, 0xf0 };

static void fill_message(size_t n)
{
    for (size_t i = 0; i < n; i++) msg[i] = (unsigned char)(rand() % 255);   /* aes-modes/test.c:54-56 */
}

static void new_key(void)
{
    aes_context e, d;
    for (int i = 0; i < KEY_BITS / 8; i++) key[i] = (unsigned char)(rand() % 255);   /* :60-62 */
    aes_setkey_enc(&e, key, (unsigned)KEY_BITS);
    aes_setkey_dec(&d, key, (unsigned)KEY_BITS);
    nr = e.nr;
    for (int i = 0; i < 4 * (nr + 1); i++) { rk_enc[i] = (uint32_t)e.rk[i]; rk_dec[i] = (uint32_t)d.rk[i]; }
}

/* one slice per thread, like ecb_test_thread / aes_ctr_test_thread (test.c:28-44, :273-285) */
static void *slice_thread(void *a)
{
    AESInfo *info = (AESInfo *)a;
    size_t blocks = MESSAGE_LENGTH / 16, per = blocks / (size_t)info->total_threads;
    size_t b0 = per * (size_t)info->thread_id;
    size_t nb = (info->thread_id == info->total_threads - 1) ? blocks - b0 : per;   /* last slice takes the remainder */
    const unsigned char *in = msg + 16 * b0;
    unsigned char *o = out + 16 * b0;
    int rc = 0;
    if (info->mode == M_ECB) rc = bc_aes_ecb_host(rk_enc, nr, BC_ENCRYPT, in, o, 16 * nb);
    else if (info->mode == M_CTR) rc = bc_aes_ctr_host(rk_enc, nr, NONCE_COUNTER, 128, b0, in, o, 16 * nb, NULL);
    else {
        const unsigned char *iv = b0 ? msg + 16 * b0 - 16 : NONCE_COUNTER;       /* previous ciphertext block */
        rc = bc_aes_cbc_decrypt_host(rk_dec, nr, iv, in, o, 16 * nb);
    }
    if (rc) { fprintf(stderr, "GPU call failed: %s\n", bc_last_error()); exit(2); }
    return NULL;
}

static int host_test(const char *name, int mode, size_t len, int num_thread)
{
    double us[MAX_ITERS];
    int ok = 1;
    MESSAGE_LENGTH = len;
    fill_message(len);
    for (int iter = 0; iter < ITERATIONS; iter++) {
        new_key();
        pthread_t th[64];
        AESInfo infos[64];
        double t0 = now_us();
        for (int t = 0; t < num_thread; t++) {
            infos[t].thread_id = t; infos[t].total_threads = num_thread; infos[t].mode = mode;
            pthread_create(&th[t], NULL, slice_thread, &infos[t]);
        }
        for (int t = 0; t < num_thread; t++) pthread_join(th[t], NULL);
        us[iter] = now_us() - t0;
        if (iter == 0) {                        /* verify once per (size, threads): inverse / unsliced equality */
            int rc;
            if (mode == M_ECB) rc = bc_aes_ecb_host(rk_dec, nr, BC_DECRYPT, out, chk, len), ok &= !rc && !memcmp(chk, msg, len);
            else if (mode == M_CTR) rc = bc_aes_ctr_host(rk_enc, nr, NONCE_COUNTER, 128, 0, out, chk, len, NULL), ok &= !rc && !memcmp(chk, msg, len);
            else rc = bc_aes_cbc_decrypt_host(rk_dec, nr, NONCE_COUNTER, msg, chk, len), ok &= !rc && !memcmp(chk, out, len);
        }
    }
    print_csv_line(name, len, num_thread, us, ITERATIONS);
    if (!ok) printf("  VERIFY FAILED for %s %zu/%d\n", name, len, num_thread);
    return ok;
}

/* device-resident: what the kernels do when the data already lives in HBM */
static int device_test(size_t len)
{
    double us[MAX_ITERS];
    int ok = 1;
    void *d_in = bc_dev_alloc(len), *d_out = bc_dev_alloc(len);
    if (!d_in || !d_out) { fprintf(stderr, "device alloc failed: %s\n", bc_last_error()); return 0; }
    size_t head = len < ((size_t)1 << 20) ? len : ((size_t)1 << 20);
    for (size_t off = 0; off < len; off += MESSAGE_LENGTH) {       /* tile the host message over the device buffer */
        size_t n = len - off < MESSAGE_LENGTH ? len - off : MESSAGE_LENGTH;
        bc_memcpy_h2d((char *)d_in + off, msg, n);
    }
    new_key();
    static const char *names[3] = { "Device ECB", "Device CTR", "Device CBC-dec" };
    for (int mode = 0; mode < 3; mode++) {
        for (int iter = 0; iter < ITERATIONS + 1; iter++) {
            double t0 = now_us();
            int rc = mode == M_ECB ? bc_aes_ecb_dev(rk_enc, nr, BC_ENCRYPT, d_in, d_out, len, NULL)
                   : mode == M_CTR ? bc_aes_ctr_dev(rk_enc, nr, NONCE_COUNTER, 128, 0, d_in, d_out, len, NULL, NULL)
                                   : bc_aes_cbc_decrypt_dev(rk_dec, nr, NONCE_COUNTER, d_in, d_out, len, NULL);
            if (!rc) rc = bc_device_sync();
            if (rc) { fprintf(stderr, "GPU call failed: %s\n", bc_last_error()); exit(2); }
            if (iter) us[iter - 1] = now_us() - t0;               /* first pass is the warm-up */
        }
        print_csv_line(names[mode], len, 1, us, ITERATIONS);
        /* head of the device result == the host-pointer path on the same bytes */
        bc_memcpy_d2h(chk, d_out, head);
        int rc = mode == M_ECB ? bc_aes_ecb_host(rk_enc, nr, BC_ENCRYPT, msg, out, head)
               : mode == M_CTR ? bc_aes_ctr_host(rk_enc, nr, NONCE_COUNTER, 128, 0, msg, out, head, NULL)
                               : bc_aes_cbc_decrypt_host(rk_dec, nr, NONCE_COUNTER, msg, out, head);
        ok &= !rc && !memcmp(chk, out, head);
    }
    bc_dev_free(d_in);
    bc_dev_free(d_out);
    if (!ok) printf("  VERIFY FAILED (device vs host path)\n");
    return ok;
}

/* BASELINE config 1: 16 MiB, reference generator exactly, AES-128 ECB/CBC/CTR, zero IV;
 * digests are those of the reference's aes.c (tests/golden/digests.json). */
static int config1(void)
{
    const size_t n = (size_t)16 << 20;
    unsigned char k32[32], iv[16], sb[16], nc[16];
    aes_context e, d;
    int ok = 1, off = 0;
    srand(1337);
    for (size_t i = 0; i < n; i++) msg[i] = (unsigned char)(rand() % 255);
    for (int i = 0; i < 32; i++) k32[i] = (unsigned char)(rand() % 255);
    aes_setkey_enc(&e, k32, 128);
    aes_setkey_dec(&d, k32, 128);
    printf("config 1: 16 MiB, srand(1337) generator, AES-128, through aes.h / aesni.h entry points\n");
    ok &= check_digest("input", msg, n, "4abae3aadfa673d106a41a04daae422d55e43962fbadb7466dcc4e0a0c02e600");
    unsigned char sched[176];
    AES_128_Key_Expansion(k32, sched);
    AES_ECB_encrypt(msg, out, n, sched, 10);
    ok &= check_digest("AES_ECB_encrypt (ECB-128 enc)", out, n, "838f1c7b577b525200e11a59d2db56de240f1416f48807175c6600d8ba9addad");
    memset(iv, 0, 16);
    aes_crypt_cbc(&d, AES_DECRYPT, n, iv, msg, out);
    ok &= check_digest("aes_crypt_cbc (CBC-128 dec)", out, n, "a9b785e15062073f25465153b099d663f3e2e2f61a14794874062516fc2913a2");
    memset(iv, 0, 16);
    aes_crypt_cbc(&e, AES_ENCRYPT, n, iv, msg, out);
    ok &= check_digest("aes_crypt_cbc (CBC-128 enc)", out, n, "3f8df46f87bf67a25ac6fecf85d0bcf1a2a706fd949f4a0abd19994b75191b34");
    memset(nc, 0, 16);
    aes_crypt_ctr(&e, (int)n, &off, nc, sb, msg, out);
    ok &= check_digest("aes_crypt_ctr (CTR-128)", out, n, "467d1fe6f3378218dd12ce3628a689d99f4172cb606a7c564f8d6ec59b60e462");
    return ok;
}

static int parse_list(const char *s, size_t *v, int max)
{
    int n = 0;
    while (*s && n < max) { v[n++] = (size_t)strtoull(s, (char **)&s, 10); if (*s == ',') s++; }
    return n;
}

int main(int argc, char *argv[])
{
    size_t sizes[16] = { 1048576, 10485760, 104857600, 1048576000 };      /* aes-modes/test.c:358-440 */
    size_t threads[16] = { 1, 2, 4, 8 };
    int n_sizes = 4, n_threads = 4, gpus = 1, do_config1 = 0, ok = 1;
    for (int i = 1; i < argc; i++) {
        if (!strcmp(argv[i], "--bits") && i + 1 < argc) KEY_BITS = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--iters") && i + 1 < argc) ITERATIONS = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--sizes") && i + 1 < argc) n_sizes = parse_list(argv[++i], sizes, 16);
        else if (!strcmp(argv[i], "--threads") && i + 1 < argc) n_threads = parse_list(argv[++i], threads, 16);
        else if (!strcmp(argv[i], "--gpus") && i + 1 < argc) gpus = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--config1")) do_config1 = 1;
        else { fprintf(stderr, "usage: %s [--bits 128|256] [--iters K] [--sizes a,b] [--threads a,b] [--gpus N] [--config1]\n", argv[0]); return 1; }
    }
    if (ITERATIONS < 1 || ITERATIONS > MAX_ITERS || (KEY_BITS != 128 && KEY_BITS != 192 && KEY_BITS != 256)) return 1;
    if (!CheckAESSupport()) { fprintf(stderr, "no usable CUDA device: %s\n", bc_last_error()); return 2; }   /* aes-modes/test.c:400 */
    srand(1337);                                                                                            /* :354 */
    size_t maxlen = (size_t)16 << 20;
    for (int i = 0; i < n_sizes; i++) if (sizes[i] > maxlen) maxlen = sizes[i];
    msg = bc_host_alloc(maxlen); out = bc_host_alloc(maxlen); chk = bc_host_alloc(maxlen);   /* pinned: DMA'd directly */
    if (!msg || !out || !chk) { fprintf(stderr, "host alloc failed: %s\n", bc_last_error()); return 2; }
    printf("# B200 bulk-cipher AES harness: AES-%d, %d iterations, SMs=%d, GPUs=%d\n", KEY_BITS, ITERATIONS, bc_sm_count(), bc_device_count());
    printf("# AES self test (aes.c:1084-1330 on the GPU): %s\n", aes_self_test(0) == 0 ? "passed" : "FAILED");

    if (do_config1) ok &= config1();
    for (int s = 0; s < n_sizes; s++) {
        size_t len = sizes[s] / 16 * 16;
        for (int t = 0; t < n_threads; t++) {
            ok &= host_test("GPU ECB", M_ECB, len, (int)threads[t]);
            ok &= host_test("GPU CTR", M_CTR, len, (int)threads[t]);
            ok &= host_test("GPU CBC-dec", M_CBC_DEC, len, (int)threads[t]);
        }
        ok &= device_test(len);
        if (gpus > 1) {                        /* one host thread per GPU, sharded by block offset, no collective */
            double us[MAX_ITERS];
            for (int iter = 0; iter < ITERATIONS; iter++) {
                double t0 = now_us();
                if (bc_aes_ctr_host_multi(gpus, rk_enc, nr, NONCE_COUNTER, 128, msg, out, len)) { fprintf(stderr, "%s\n", bc_last_error()); return 2; }
                us[iter] = now_us() - t0;
            }
            print_csv_line("Multi-GPU CTR", len, gpus, us, ITERATIONS);
            bc_aes_ctr_host(rk_enc, nr, NONCE_COUNTER, 128, 0, msg, chk, len, NULL);
            ok &= !memcmp(chk, out, len);
        }
    }
    printf("# %s\n", ok ? "all outputs verified" : "VERIFICATION FAILED");
    bc_host_free(msg); bc_host_free(out); bc_host_free(chk);
    return ok ? 0 : 3;
}
