/*
 * bulkcipher.h — the B200 bulk-cipher engine's own C-ABI (extension symbols).
 *
 * The drop-in symbols of the reference live in aes.h / aesni.h / arc4.h /
 * rc4.h next to this file and are implemented on top of the calls below.
 * These bc_* entry points are what the BASELINE configs force beyond the
 * reference API (SURVEY §8b "Extensions the configs force"): device-resident
 * buffers, size_t lengths, an explicit starting block for CTR shards, batched
 * CBC-encrypt and batched RC4.  Plain pointers and sizes only; `stream` is a
 * cudaStream_t passed as void* (NULL = the default stream).
 *
 * Every function returns 0 on success or a negative BC_ERR_* code;
 * bc_last_error() gives the calling thread's last message.  There is no CPU
 * fallback anywhere: without a usable CUDA device every call fails with
 * BC_ERR_CUDA / BC_ERR_NO_DEVICE.
 *
 * Round keys (`rk`) are 4*(nr+1) little-endian 32-bit words = the low halves
 * of aes_context.rk[] (aes-modes/aes.h:41-47), i.e. exactly the packed
 * schedule aesni.c consumes (aes-modes/aesni.c:79-97).  `rk` is a HOST pointer;
 * it travels to the GPU as a kernel argument (constant bank), so calls are
 * re-entrant and no global key state exists.
 */
#ifndef BULKCIPHER_H
#define BULKCIPHER_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BC_OK             0
#define BC_ERR_CUDA       (-0x7001)  /* a CUDA runtime call or launch failed        */
#define BC_ERR_BAD_ARG    (-0x7002)  /* NULL pointer, bad nr / mode / length         */
#define BC_ERR_NO_DEVICE  (-0x7003)  /* no CUDA device visible                       */
#define BC_ERR_OVERLAP    (-0x7004)  /* in/out partially overlap (must be == or disjoint) */
#define BC_ERR_TOO_LARGE  (-0x7005)  /* request exceeds a documented fixed limit     */

#define BC_ENCRYPT 1                 /* == AES_ENCRYPT, aes-modes/aes.h:33 */
#define BC_DECRYPT 0                 /* == AES_DECRYPT, aes-modes/aes.h:34 */

const char *bc_last_error(void);
/* The reference's void entry points (aesni.h bulk calls, arc4_setup, rc4_init, rc4_crypt) cannot
 * return a failure.  When one fails it prints the message to stderr, leaves its output untouched
 * and records the BC_ERR_* code for the calling thread; bc_void_status() returns that code (0 if
 * none) and clears it.  With BULKCIPHER_ABORT_ON_ERROR=1 in the environment a failure aborts the
 * process instead.  (Extension: the reference has no error reporting at all) */
int  bc_void_status(void);
int  bc_device_count(void);
/* Device used by the calling thread's *_host calls and allocations (default 0). */
int  bc_set_device(int device);
int  bc_get_device(void);
int  bc_sm_count(void);
/* Frees every buffer, stream and event the library owns on all devices (optional; the next
 * call allocates again).  No other call may be in flight. */
int  bc_shutdown(void);

/* ---- memory helpers for C callers (harness, multi-GPU driver) ------------ */
/* bc_memcpy_* are synchronous copies on the legacy default stream: follow them with calls on
 * stream NULL, or synchronise before using a non-blocking stream of your own. */
void *bc_host_alloc(size_t bytes);            /* pinned host memory (cudaHostAlloc)   */
void *bc_host_alloc_wc(size_t bytes);         /* pinned + write-combined: CPU writes, GPU reads (an H2D source); free with bc_host_free */
void  bc_host_free(void *p);
void *bc_dev_alloc(size_t bytes);
void  bc_dev_free(void *p);
int   bc_memcpy_h2d(void *d_dst, const void *h_src, size_t bytes);
int   bc_memcpy_d2h(void *h_dst, const void *d_src, size_t bytes);
int   bc_device_sync(void);

/* ---- AES on device-resident buffers -------------------------------------- */

/* CTR keystream XOR.  Replaces the loop of aes_crypt_ctr (aes-modes/aes.c:869-900)
 * and AES_CTR_encrypt (aes-modes/aesni.c:120-152) for a buffer already in HBM.
 * Block i of the buffer uses counter = `counter` (16 bytes, big-endian integer)
 * + first_block + i.  counter_bits = 128: full 128-bit carry (aes.c:884-888);
 * counter_bits = 64: the add wraps inside bytes 8..15 and never carries into
 * bytes 0..7 (_mm_add_epi64, aesni.c:139,143).  nbytes need not be a multiple of
 * 16 (the last block is used partially, like aes.c:879-895).  If
 * d_last_keystream != NULL it receives the 16-byte keystream block of the last
 * block touched (the reference's stream_block on return); it must be 16-byte
 * aligned like the data buffers.  in == out allowed. */
int bc_aes_ctr_dev(const uint32_t *rk, int nr, const unsigned char counter[16], int counter_bits,
                   uint64_t first_block, const void *d_in, void *d_out, size_t nbytes,
                   void *d_last_keystream, void *stream);

/* ECB over nbytes/16 whole blocks.  Replaces the per-block loop over
 * aes_crypt_ecb (aes-modes/aes.c:650-752, harness aes-modes/test.c:37-41) and
 * AES_ECB_encrypt / AES_ECB_decrypt (aes-modes/aesni.c:79-118).  `mode` selects
 * the tables only; rk must be the matching (enc or equivalent-inverse dec)
 * schedule, exactly as in the reference (aes.c:677,711).  in == out allowed. */
int bc_aes_ecb_dev(const uint32_t *rk, int nr, int mode, const void *d_in, void *d_out,
                   size_t nbytes, void *stream);

/* ECB where each block is four BIG-endian 32-bit words in memory: the uint[] block
 * interface of the reference's C++ class AES (aes-gpu/Source/AES.h:84-147, GETWORD /
 * PUTWORD at AES.cu:42-43, byte2int/int2byte at AES.cu:190-203).  include/AES.h wraps it. */
int bc_aes_ecb_be32_dev(const uint32_t *rk, int nr, int mode, const void *d_in, void *d_out,
                        size_t nbytes, void *stream);

/* CBC decrypt, parallel over blocks.  Replaces aes_crypt_cbc(AES_DECRYPT)
 * (aes-modes/aes.c:784-797).  iv (host, 16 bytes) is read only; the caller gets
 * the new iv as the last ciphertext block (aes.c:792).  in == out is allowed
 * and then keeps the ciphertext block in front of every 4 KiB row-group in
 * scratch (16 bytes per 4 KiB of input) that is allocated and freed in stream
 * order from a library-owned memory pool: any number of streams and host
 * threads may run such calls at once, and nothing synchronises the device. */
int bc_aes_cbc_decrypt_dev(const uint32_t *rk, int nr, const unsigned char iv[16],
                           const void *d_in, void *d_out, size_t nbytes, void *stream);

/* No counterpart in the reference.  In-place CBC decrypt like
 * bc_aes_cbc_decrypt_dev(d_in == d_out), for callers that must not trigger a
 * device allocation: services that provision all their HBM up front and run
 * against a fixed budget (a late cudaMalloc can fail or stall them), code that
 * owns a memory pool and has to account for every byte on the device, and
 * latency-sensitive paths where the device-wide synchronisation an allocation
 * or free can imply would stall unrelated streams.  bc_aes_cbc_decrypt_dev
 * takes its scratch from a memory pool that grows on demand; this entry point instead
 * parks the per-4-KiB boundary blocks in a fixed 1 MiB array that is part of
 * the library image (each GPU that loads the library holds it until exit) and
 * calls no cudaMalloc / cudaFree or any other allocating routine.
 *   - Limit: nbytes <= 256 MiB (65536 row-groups) per call, multiple of 16;
 *     larger requests fail with BC_ERR_TOO_LARGE and touch nothing.  Longer
 *     buffers: split on a multiple of 4096 bytes and pass the last ciphertext
 *     block of one piece (read it before the call) as the iv of the next.
 *   - The first call on each device looks up the array's device address
 *     (cudaGetSymbolAddress, with that device current); make that first call
 *     before starting any stream capture.  That first call (per device and
 *     key size) is also where the CUDA runtime loads the kernel image and the
 *     library sets the kernel's shared-memory attribute: treat it as a warm-up
 *     and make it before the budgeted or latency-critical phase begins.
 *   - All calls on a device share the one array, whatever the stream: two
 *     calls whose kernels can overlap on the GPU race on it.  Order them (same
 *     stream, or an event) — calls that follow one another need no clearing,
 *     every entry a call reads it has just rewritten itself. */
int bc_aes_cbc_decrypt_inplace_noalloc_dev(const uint32_t *rk, int nr, const unsigned char iv[16],
                                           void *d_buf, size_t nbytes, void *stream);

/* CBC encrypt of n_msgs independent messages of msg_len bytes each (multiple
 * of 16), message-major contiguous.  Each message is one aes_crypt_cbc(
 * AES_ENCRYPT) call of the reference (aes-modes/aes.c:801-812): serial inside a
 * message, parallel across messages.  d_ivs is [n_msgs][16]; if d_ivs_out !=
 * NULL it receives each message's last ciphertext block (aes.c:807). */
int bc_aes_cbc_encrypt_batch_dev(const uint32_t *rk, int nr, size_t n_msgs, size_t msg_len,
                                 const void *d_ivs, void *d_ivs_out,
                                 const void *d_in, void *d_out, void *stream);

/* CFB128, the mode aes-modes/aes.h:124 declares and aes.c:822-862 implements behind
 * POLARSSL_CIPHER_MODE_CFB (off in the reference build).  Whole blocks; rk is the
 * ENCRYPTION schedule in both directions (aes.c:835,849).  Decrypt is parallel:
 * P_i = E(C_{i-1}) ^ C_i, same chaining/scratch rules as bc_aes_cbc_decrypt_dev.
 * Encrypt is serial per message, C_i = E(C_{i-1}) ^ P_i, batched across messages
 * like bc_aes_cbc_encrypt_batch_dev. */
int bc_aes_cfb128_decrypt_dev(const uint32_t *rk, int nr, const unsigned char iv[16],
                              const void *d_in, void *d_out, size_t nbytes, void *stream);
int bc_aes_cfb128_encrypt_batch_dev(const uint32_t *rk, int nr, size_t n_msgs, size_t msg_len,
                                    const void *d_ivs, void *d_ivs_out,
                                    const void *d_in, void *d_out, void *stream);

/* ---- RC4 / ARC4 on device-resident buffers ------------------------------- */

/* n_streams independent RC4 streams, stream-major contiguous: for stream s,
 * key = d_keys[s*keylen .. +keylen), data = d_in[s*length .. +length).  Each
 * stream is arc4_setup -> arc4_prep(length) -> arc4_crypt (arc4.c:43-112) =
 * rc4_init -> rc4_crypt (rc4.c:56-99, intended semantics), KSA+PRGA+XOR fused.
 * If d_ctx_out != NULL it receives n_streams arc4_context structs (264 bytes:
 * int x, int y, m[256]; arc4.h:35-41) holding the state after `length` bytes.
 * If d_ctx_in != NULL the KSA is skipped and streams resume from those
 * contexts (keys ignored).  1 <= keylen <= 256.  in == out allowed. */
int bc_arc4_batch_dev(size_t n_streams, const void *d_keys, unsigned keylen, size_t length,
                      const void *d_in, void *d_out, const void *d_ctx_in, void *d_ctx_out,
                      void *stream);

/* out = a ^ b: the whole of arc4_crypt once a keystream exists (arc4.c:101-112).
 * The three buffers may alias elementwise. */
int bc_xor_dev(const void *d_a, const void *d_b, void *d_out, size_t nbytes, void *stream);

/* ---- the same operations on HOST buffers --------------------------------- */
/* Chunked, multi-stream H2D -> kernel -> D2H pipeline through library-owned
 * device buffers.  Buffers from bc_host_alloc / cudaHostRegister are DMA'd
 * directly; pageable memory is staged through internal pinned buffers.  These
 * are what the drop-in symbols call; they are PCIe-bound. */
int bc_aes_ctr_host(const uint32_t *rk, int nr, const unsigned char counter[16], int counter_bits,
                    uint64_t first_block, const void *in, void *out, size_t nbytes,
                    unsigned char last_keystream[16]);
int bc_aes_ecb_host(const uint32_t *rk, int nr, int mode, const void *in, void *out, size_t nbytes);
int bc_aes_ecb_be32_host(const uint32_t *rk, int nr, int mode, const void *in, void *out, size_t nbytes);
int bc_aes_cbc_decrypt_host(const uint32_t *rk, int nr, const unsigned char iv[16],
                            const void *in, void *out, size_t nbytes);
int bc_aes_cbc_encrypt_batch_host(const uint32_t *rk, int nr, size_t n_msgs, size_t msg_len,
                                  const void *ivs, void *ivs_out, const void *in, void *out);
int bc_aes_cfb128_decrypt_host(const uint32_t *rk, int nr, const unsigned char iv[16],
                               const void *in, void *out, size_t nbytes);
int bc_aes_cfb128_encrypt_batch_host(const uint32_t *rk, int nr, size_t n_msgs, size_t msg_len,
                                     const void *ivs, void *ivs_out, const void *in, void *out);
/* in == NULL (here and in bc_arc4_batch_dev): write the raw keystream, which is
 * what arc4_prep produces (arc4.c:72-97).  length == 0 with ctx_out: key setup only. */
int bc_arc4_batch_host(size_t n_streams, const void *keys, unsigned keylen, size_t length,
                       const void *in, void *out, const void *ctx_in, void *ctx_out);
int bc_xor_host(const void *a, const void *b, void *out, size_t nbytes);

/* Registration cache for pageable host buffers (off by default).  The reference's harnesses
 * malloc() a buffer once and pass it to the API ten times (aes-modes/test.c:287-350); pageable
 * memory has to be staged through pinned buffers by host memcpy, pinned memory is DMA'd directly.
 * With the cache on, a range of at least 64 MiB that a bc_*_host call (or a drop-in built on one)
 * sees for the second time is pinned in place (cudaHostRegister) and later calls take the pinned
 * path.  Every use re-checks, from /proc/self/pagemap, that each page of the range is still the
 * physical page that was pinned, so freed-and-reused addresses are noticed and fall back to
 * staging; that needs CAP_SYS_ADMIN (frame numbers are hidden otherwise) and the call returns 0
 * where it cannot be had.  It stays opt-in because of what the library cannot see: between the
 * caller's free() and the library's next look at those addresses CUDA still holds them as
 * registered, and an unrelated cudaHostAlloc / cudaHostRegister that lands there fails with
 * cudaErrorAlreadyMapped.  Call bc_register_pageable(0) — it releases every idle registration —
 * before freeing big buffers in a process that also pins memory elsewhere.  Returns 1 if the cache
 * is on after the call, else 0.  BULKCIPHER_REGISTER_PAGEABLE=1 in the environment switches it on
 * for unmodified callers (e.g. the relinked reference harness). */
int bc_register_pageable(int enable);

/* ---- multi-GPU (one host thread per GPU, no collective) ------------------ */
/* The reference's slice drivers with a device per slice instead of a core per slice
 * (aes-modes/test.c:28-44 ECB, :194-206 AES-NI ECB, :273-285 CTR; test.c:44-58 RC4 XOR):
 * HOST buffers are cut into equal contiguous ranges, GPU g of n_gpus (devices 0..n_gpus-1) takes
 * range g, the last GPU takes the remainder (the rule of aes-modes/test.c:33-35).  Each shard is
 * an ordinary bc_*_host call on its own device from its own host thread, bound to the CPUs of
 * the GPU's NUMA node; nothing is exchanged between GPUs.  The result is identical
 * to the single bc_*_host call with the same arguments.
 *
 * CTR: by 16-byte block ranges; shard g's counter starts at its first block index (the
 * reference harness restarts the counter per slice, aes-modes/test.c:282 — a bug not kept). */
int bc_aes_ctr_host_multi(int n_gpus, const uint32_t *rk, int nr, const unsigned char counter[16],
                          int counter_bits, const void *in, void *out, size_t nbytes);
/* ECB: by block ranges (whole blocks only). */
int bc_aes_ecb_host_multi(int n_gpus, const uint32_t *rk, int nr, int mode,
                          const void *in, void *out, size_t nbytes);
/* CBC / CFB128 decrypt: by block ranges; shard g > 0 takes its chaining block from the last
 * ciphertext block of shard g-1 (aes.c:784-797), read from `in` before anything is written, so
 * in == out is fine. */
int bc_aes_cbc_decrypt_host_multi(int n_gpus, const uint32_t *rk, int nr, const unsigned char iv[16],
                                  const void *in, void *out, size_t nbytes);
int bc_aes_cfb128_decrypt_host_multi(int n_gpus, const uint32_t *rk, int nr, const unsigned char iv[16],
                                     const void *in, void *out, size_t nbytes);
/* Batched CBC / CFB128 encrypt: by message index (a single message is a serial chain and does
 * not shard: replicas only). */
int bc_aes_cbc_encrypt_batch_host_multi(int n_gpus, const uint32_t *rk, int nr, size_t n_msgs, size_t msg_len,
                                        const void *ivs, void *ivs_out, const void *in, void *out);
int bc_aes_cfb128_encrypt_batch_host_multi(int n_gpus, const uint32_t *rk, int nr, size_t n_msgs, size_t msg_len,
                                           const void *ivs, void *ivs_out, const void *in, void *out);
/* Batched RC4: by stream index (keys, data and contexts of stream s stay together). */
int bc_arc4_batch_host_multi(int n_gpus, size_t n_streams, const void *keys, unsigned keylen, size_t length,
                             const void *in, void *out, const void *ctx_in, void *ctx_out);

/* Host-side locality helpers used by the calls above (and by one-process-per-GPU callers such as
 * bench.py before they allocate pinned memory).  bc_device_numa_node: the NUMA node the GPU's PCIe
 * slot reports in sysfs, or -1.  bc_bind_host_thread_to_device: restrict the calling thread to
 * that node's CPUs (within its current affinity mask) and prefer the node for its next page
 * allocations; a no-op on single-node boxes.  The reference pins nothing (pthread_create with
 * default attributes, aes-modes/test.c:80). */
int bc_device_numa_node(int device);
int bc_bind_host_thread_to_device(int device);

/* ---- synthetic input for the harness -------------------------------------- */
/* Device counterpart of the reference harness's generator (srand(1337); msg[i] = rand() % 255,
 * aes-modes/test.c:54-56, :354) for buffers too large to build on the host (BASELINE config 5,
 * 256 GiB): little-endian 64-bit word k of d_out = mix64(seed + first_word + k), where
 *   mix64(x): z = x + 0x9e3779b97f4a7c15; z = (z ^ z>>30) * 0xbf58476d1ce4e5b9;
 *             z = (z ^ z>>27) * 0x94d049bb133111eb; return z ^ z>>31      (splitmix64 output)
 * so the host can replay any window of the logical buffer.  d_out 16-byte aligned. */
int bc_fill_splitmix64_dev(void *d_out, size_t nbytes, uint64_t seed, uint64_t first_word, void *stream);

/* ---- key schedules as plain word arrays (what aes_setkey_enc/dec compute) ---- */
/* FIPS-197 expansion on little-endian words (aes-modes/aes.c:469-539); returns nr (10/12/14) or 0. */
int  bc_aes_expand_key(const unsigned char *key, unsigned int keybits, uint32_t rk[60]);
/* Equivalent-inverse-cipher schedule from an encryption schedule (aes-modes/aes.c:563-594). */
void bc_aes_invert_key(const uint32_t enc[60], int nr, uint32_t dec[60]);

#ifdef __cplusplus
}
#endif
#endif /* BULKCIPHER_H */
