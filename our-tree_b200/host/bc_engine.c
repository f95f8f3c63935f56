/*
 * bc_engine.c — host side of libbulkcipher.so, in C: per-thread error slot,
 * device bookkeeping, library-owned scratch, and the HOST-buffer entry points
 * (bc_*_host): a chunked, multi-stream H2D -> kernel -> D2H pipeline around the
 * device entry points of csrc/.  This is the layer the reference's drop-in
 * symbols (aes_dropin.c, aesni_dropin.c, arc4_dropin.c, rc4_dropin.c) call.
 *
 * The reference has no counterpart for most of this: its only device boundary
 * is cudaMalloc + cudaMemcpy + launch + cudaMemcpy + cudaFree per call, all
 * unchecked and serial (aes-gpu/Source/AES.cu:230-255).  Here every CUDA call
 * is checked, buffers persist per device, and copies overlap compute.
 *
 * There is no CPU fallback: if CUDA is unusable every entry point fails.
 */
#define _GNU_SOURCE
#include <pthread.h>
#include <sched.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/syscall.h>
#include <unistd.h>

#include <cuda_runtime_api.h>

#include "bulkcipher.h"
#include "../csrc/bc_internal.h"

#define BC_MAX_DEVICES 16
#define BC_MAX_SLOTS   8
/* Pipeline shape: `g_slots` chunks in flight, each `g_chunk` bytes (a multiple of 4 KiB).
 * Defaults (64 MiB x 3) were picked on a B200 / PCIe Gen5 box (tools/e2e_probe.py);
 * BULKCIPHER_CHUNK_MIB and BULKCIPHER_SLOTS override them before the first call. */
static size_t g_chunk = (size_t)64 << 20;
static int    g_slots = 3;
static size_t g_pageable_chunk = (size_t)16 << 20;   /* cap for pageable (host-in-the-loop) transfers */
#define BC_SLOTS       g_slots
#define BC_CHUNK_BYTES g_chunk
#define BC_MAX_SCRATCH 8

/* ------------------------------------------------------------ error slot */

static __thread char tls_error[512];
static __thread int  tls_device = 0;

void bc_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(tls_error, sizeof tls_error, fmt, ap);
    va_end(ap);
}

void bc_set_error_str(const char *msg) { bc_set_error("%s", msg); }

int bc_cuda_fail(int cuda_err, const char *what)
{
    bc_set_error("%s: %s (%s)", what, cudaGetErrorString((cudaError_t)cuda_err),
                 cudaGetErrorName((cudaError_t)cuda_err));
    return BC_ERR_CUDA;
}

const char *bc_last_error(void) { return tls_error; }

/* The reference's void entry points (aesni.h bulk calls, arc4_setup, rc4_*) have no way to
 * return a failure.  They report it here instead: the message goes to stderr, the output buffer
 * is left untouched, the code is kept per thread until read.  BULKCIPHER_ABORT_ON_ERROR=1 turns a
 * failure into abort() for hosts that prefer a crash to a missed check. */
static __thread int tls_void_status = 0;

void bc_void_call_failed(int rc, const char *fn)
{
    tls_void_status = rc;
    fprintf(stderr, "libbulkcipher: %s failed (%d): %s - output left untouched (no CPU fallback)\n", fn, rc, tls_error);
    const char *e = getenv("BULKCIPHER_ABORT_ON_ERROR");
    if (e && *e && *e != '0') abort();
}

int bc_void_status(void)
{
    const int rc = tls_void_status;
    tls_void_status = 0;
    return rc;
}

/* ---------------------------------------------------------- device state */

typedef struct {
    cudaStream_t stream;
    cudaEvent_t  done;
    unsigned char *d_buf;       /* BC_CHUNK_BYTES of HBM, processed in place */
    unsigned char *pin_in;      /* staging for pageable callers (lazy) */
    unsigned char *pin_out;
    unsigned char *pin_b;       /* staging for a pageable second operand (bc_xor_host) */
    /* pending copy-out of a staged chunk */
    unsigned char *pend_dst;
    size_t         pend_len;
} bc_slot;

typedef struct {
    void *key;                  /* stream the buffer was last handed out for */
    void *ptr;
    size_t cap;
} bc_scratch;

typedef struct {
    pthread_mutex_t lock;       /* one host-pipeline call at a time per device */
    pthread_mutex_t scratch_lock;
    int        ready;
    int        sm_count;
    bc_slot    slot[BC_MAX_SLOTS];
    bc_scratch scratch[BC_MAX_SCRATCH];
    cudaMemPool_t pool;         /* stream-ordered scratch (bc_stream_scratch_alloc); NULL until first use */
    unsigned char *d_aux;       /* small device buffer: ivs, keys, contexts, last keystream */
    size_t     aux_cap;
    /* enqueue-ahead pipeline for pinned callers: one stream per stage, events chain the slots */
    cudaStream_t s_h2d, s_k, s_d2h;
    cudaEvent_t  e_h2d[BC_MAX_SLOTS], e_k[BC_MAX_SLOTS], e_d2h[BC_MAX_SLOTS];
} bc_dev;

static bc_dev g_dev[BC_MAX_DEVICES];
static pthread_once_t g_once = PTHREAD_ONCE_INIT;

static void init_locks(void)
{
    const char *e = getenv("BULKCIPHER_CHUNK_MIB");
    if (e && atoi(e) >= 1 && atoi(e) <= 1024) g_chunk = (size_t)atoi(e) << 20;
    e = getenv("BULKCIPHER_PAGEABLE_CHUNK_MIB");
    if (e && atoi(e) >= 1 && atoi(e) <= 1024) g_pageable_chunk = (size_t)atoi(e) << 20;
    e = getenv("BULKCIPHER_SLOTS");
    if (e && atoi(e) >= 2 && atoi(e) <= BC_MAX_SLOTS) g_slots = atoi(e);
    for (int i = 0; i < BC_MAX_DEVICES; i++) {
        pthread_mutex_init(&g_dev[i].lock, NULL);
        pthread_mutex_init(&g_dev[i].scratch_lock, NULL);
    }
}

int bc_device_count(void)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) { bc_cuda_fail((int)e, "cudaGetDeviceCount"); return 0; }
    return n;
}

int bc_set_device(int device)
{
    int n = bc_device_count();
    if (n <= 0) { if (!tls_error[0]) bc_set_error("no CUDA device"); return BC_ERR_NO_DEVICE; }
    if (device < 0 || device >= n || device >= BC_MAX_DEVICES) {
        bc_set_error("device %d out of range (have %d)", device, n);
        return BC_ERR_BAD_ARG;
    }
    BC_CUDA(cudaSetDevice(device), "cudaSetDevice");
    tls_device = device;
    return BC_OK;
}

int bc_get_device(void) { return tls_device; }

static bc_dev *cur_dev_state(int *dev_out)
{
    int dev = 0;
    pthread_once(&g_once, init_locks);
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= BC_MAX_DEVICES) return NULL;
    if (dev_out) *dev_out = dev;
    return &g_dev[dev];
}

int bc_cached_sm_count(void)
{
    int dev = 0;
    bc_dev *d = cur_dev_state(&dev);
    if (!d) { bc_set_error("no current CUDA device"); return -1; }
    if (d->sm_count <= 0) {
        int n = 0;
        cudaError_t e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) { bc_cuda_fail((int)e, "cudaDeviceGetAttribute(SM count)"); return -1; }
        d->sm_count = n;
    }
    return d->sm_count;
}

int bc_sm_count(void)
{
    if (bc_set_device(tls_device)) return -1;
    return bc_cached_sm_count();
}

/* Scratch in HBM, grow-only, one buffer per key.  Only the library's own host-pipeline calls use
 * this table, with a fixed key each and under the device lock, so an entry is never shared by two
 * calls in flight.  Per-stream scratch of the device entry points comes from
 * bc_stream_scratch_alloc below (stream-ordered, any number of streams). */
int bc_scratch_get(void *key, size_t bytes, void **d_ptr)
{
    bc_dev *d = cur_dev_state(NULL);
    if (!d) { bc_set_error("no current CUDA device"); return BC_ERR_NO_DEVICE; }
    pthread_mutex_lock(&d->scratch_lock);
    bc_scratch *s = NULL;
    for (int i = 0; i < BC_MAX_SCRATCH && !s; i++)
        if (d->scratch[i].ptr && d->scratch[i].key == key) s = &d->scratch[i];
    for (int i = 0; i < BC_MAX_SCRATCH && !s; i++)
        if (!d->scratch[i].ptr) s = &d->scratch[i];
    if (!s) {
        pthread_mutex_unlock(&d->scratch_lock);
        bc_set_error("internal scratch table full (%d keys)", BC_MAX_SCRATCH);
        return BC_ERR_TOO_LARGE;
    }
    int rc = BC_OK;
    if (s->cap < bytes) {
        if (s->ptr) { cudaDeviceSynchronize(); cudaFree(s->ptr); s->ptr = NULL; s->cap = 0; }
        size_t cap = bytes < ((size_t)1 << 20) ? ((size_t)1 << 20) : bytes;
        cudaError_t e = cudaMalloc(&s->ptr, cap);
        if (e != cudaSuccess) { rc = bc_cuda_fail((int)e, "cudaMalloc(scratch)"); s->ptr = NULL; }
        else s->cap = cap;
    }
    s->key = key;
    *d_ptr = s->ptr;
    pthread_mutex_unlock(&d->scratch_lock);
    return rc;
}

/* Stream-ordered scratch for the device entry points (in-place CBC/CFB decrypt keeps one
 * ciphertext block per 4 KiB row-group): allocated and freed IN STREAM ORDER from a library-owned
 * memory pool, so the buffer belongs to exactly the kernels enqueued between the two calls —
 * whatever the number of streams or host threads — and no call ever synchronises the device.
 * The pool keeps freed memory (release threshold = max), so steady-state calls cost microseconds
 * and touch the driver's allocator only when a larger buffer than ever before is needed. */
int bc_stream_scratch_alloc(void *stream, size_t bytes, void **d_ptr)
{
    int dev = 0;
    bc_dev *d = cur_dev_state(&dev);
    if (!d) { bc_set_error("no current CUDA device"); return BC_ERR_NO_DEVICE; }
    pthread_mutex_lock(&d->scratch_lock);
    if (!d->pool) {
        struct cudaMemPoolProps props;
        memset(&props, 0, sizeof props);
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = dev;
        cudaError_t e = cudaMemPoolCreate(&d->pool, &props);
        if (e == cudaSuccess) {
            unsigned long long keep = ~0ull;
            e = cudaMemPoolSetAttribute(d->pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
        if (e != cudaSuccess) {
            if (d->pool) { cudaMemPoolDestroy(d->pool); d->pool = NULL; }
            pthread_mutex_unlock(&d->scratch_lock);
            return bc_cuda_fail((int)e, "cudaMemPoolCreate(scratch pool)");
        }
    }
    cudaMemPool_t pool = d->pool;
    pthread_mutex_unlock(&d->scratch_lock);
    BC_CUDA(cudaMallocFromPoolAsync(d_ptr, bytes ? bytes : 16, pool, (cudaStream_t)stream), "cudaMallocFromPoolAsync(scratch)");
    return BC_OK;
}

int bc_stream_scratch_free(void *stream, void *d_ptr)
{
    if (!d_ptr) return BC_OK;
    BC_CUDA(cudaFreeAsync(d_ptr, (cudaStream_t)stream), "cudaFreeAsync(scratch)");
    return BC_OK;
}

/* ----------------------------------------------------- memory conveniences */

void *bc_host_alloc(size_t bytes)
{
    void *p = NULL;
    if (bc_set_device(tls_device)) return NULL;
    cudaError_t e = cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable);
    if (e != cudaSuccess) { bc_cuda_fail((int)e, "cudaHostAlloc"); return NULL; }
    return p;
}

/* Write-combined pinned memory: for buffers the CPU only WRITES and the GPU only reads (an H2D
 * source).  The device's reads need no cache snoop; CPU reads from it are very slow. */
void *bc_host_alloc_wc(size_t bytes)
{
    void *p = NULL;
    if (bc_set_device(tls_device)) return NULL;
    cudaError_t e = cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable | cudaHostAllocWriteCombined);
    if (e != cudaSuccess) { bc_cuda_fail((int)e, "cudaHostAlloc(write-combined)"); return NULL; }
    return p;
}

void bc_host_free(void *p) { if (p) cudaFreeHost(p); }

void *bc_dev_alloc(size_t bytes)
{
    void *p = NULL;
    if (bc_set_device(tls_device)) return NULL;
    cudaError_t e = cudaMalloc(&p, bytes ? bytes : 1);
    if (e != cudaSuccess) { bc_cuda_fail((int)e, "cudaMalloc"); return NULL; }
    return p;
}

void bc_dev_free(void *p) { if (p) cudaFree(p); }

int bc_memcpy_h2d(void *d_dst, const void *h_src, size_t bytes)
{
    BC_CUDA(cudaMemcpy(d_dst, h_src, bytes, cudaMemcpyHostToDevice), "cudaMemcpy H2D");
    return BC_OK;
}

int bc_memcpy_d2h(void *h_dst, const void *d_src, size_t bytes)
{
    BC_CUDA(cudaMemcpy(h_dst, d_src, bytes, cudaMemcpyDeviceToHost), "cudaMemcpy D2H");
    return BC_OK;
}

int bc_device_sync(void)
{
    BC_CUDA(cudaDeviceSynchronize(), "cudaDeviceSynchronize");
    return BC_OK;
}

/* ------------------------------------------------------ parallel staging copy */

/* Pageable caller memory has to pass through pinned staging buffers, and one core's memcpy
 * (~10 GB/s) is far below the PCIe rate.  A small pool of helper threads splits every large
 * staging copy; the calling thread takes a share too.  Pure data movement: no cipher work is
 * ever done on the host.  BULKCIPHER_COPY_THREADS overrides the helper count (0 = none). */
#define BC_MAX_COPY_THREADS 16
#define BC_PAR_COPY_MIN ((size_t)4 << 20)

typedef struct { unsigned char *dst; const unsigned char *src; size_t n; } bc_copy_job;

static pthread_mutex_t g_copy_user = PTHREAD_MUTEX_INITIALIZER;   /* one parallel copy at a time */
static pthread_mutex_t g_copy_mu   = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t  g_copy_work = PTHREAD_COND_INITIALIZER;
static pthread_cond_t  g_copy_done = PTHREAD_COND_INITIALIZER;
static bc_copy_job     g_copy_jobs[BC_MAX_COPY_THREADS];
static unsigned        g_copy_gen;
static int             g_copy_pending;
static int             g_copy_threads = -1;                       /* -1: pool not started yet */

/* Staging copies move data that the CPU will not touch again soon (the copy engine reads the
 * staging buffer, the caller reads its output much later), and each helper's share is far below the
 * size at which glibc's memcpy switches to non-temporal stores by itself.  Ordinary stores make the
 * core read every destination line before overwriting it (read-for-ownership): 3 bytes of DRAM
 * traffic per byte copied instead of 2.  On x86-64 large shares are therefore copied with streaming
 * stores (SSE2, baseline ISA).  BULKCIPHER_NT_COPY=0 falls back to memcpy (A/B). */
#if defined(__x86_64__)
#include <emmintrin.h>
static int g_nt_copy = -1;

static void stream_copy(void *dst, const void *src, size_t n)
{
    if (g_nt_copy < 0) { const char *e = getenv("BULKCIPHER_NT_COPY"); g_nt_copy = !(e && *e == '0'); }
    if (!g_nt_copy || n < ((size_t)256 << 10)) { memcpy(dst, src, n); return; }
    unsigned char *d = (unsigned char *)dst;
    const unsigned char *s = (const unsigned char *)src;
    const size_t head = (16 - ((uintptr_t)d & 15)) & 15;             /* streaming stores need a 16-byte aligned target */
    if (head) { memcpy(d, s, head); d += head; s += head; n -= head; }
    size_t blocks = n / 64;
    while (blocks--) {
        const __m128i a = _mm_loadu_si128((const __m128i *)s), b = _mm_loadu_si128((const __m128i *)(s + 16));
        const __m128i c = _mm_loadu_si128((const __m128i *)(s + 32)), e = _mm_loadu_si128((const __m128i *)(s + 48));
        _mm_stream_si128((__m128i *)d, a); _mm_stream_si128((__m128i *)(d + 16), b);
        _mm_stream_si128((__m128i *)(d + 32), c); _mm_stream_si128((__m128i *)(d + 48), e);
        s += 64; d += 64;
    }
    _mm_sfence();                                                     /* streaming stores are weakly ordered: publish before anyone is told */
    if (n % 64) memcpy(d, s, n % 64);
}
#else
static void stream_copy(void *dst, const void *src, size_t n) { memcpy(dst, src, n); }
#endif

static void *copy_worker(void *arg)
{
    const int id = (int)(intptr_t)arg;
    unsigned seen = 0;
    for (;;) {
        pthread_mutex_lock(&g_copy_mu);
        while (seen == g_copy_gen) pthread_cond_wait(&g_copy_work, &g_copy_mu);
        seen = g_copy_gen;
        const bc_copy_job job = g_copy_jobs[id];
        pthread_mutex_unlock(&g_copy_mu);
        if (job.n) stream_copy(job.dst, job.src, job.n);
        pthread_mutex_lock(&g_copy_mu);
        if (--g_copy_pending == 0) pthread_cond_signal(&g_copy_done);
        pthread_mutex_unlock(&g_copy_mu);
    }
    return NULL;
}

/* called with g_copy_user held */
static void copy_pool_start(void)
{
    long want = sysconf(_SC_NPROCESSORS_ONLN) * 3 / 4;            /* measured: 8 helpers 17 GB/s, 12 helpers 20 GB/s */
    const char *e = getenv("BULKCIPHER_COPY_THREADS");
    if (e) want = atoi(e);
    if (want > 12 && !e) want = 12;
    if (want > BC_MAX_COPY_THREADS) want = BC_MAX_COPY_THREADS;
    if (want < 0) want = 0;
    int started = 0;
    for (long i = 0; i < want; i++) {
        pthread_t th;
        pthread_attr_t at;
        pthread_attr_init(&at);
        pthread_attr_setdetachstate(&at, PTHREAD_CREATE_DETACHED);
        if (pthread_create(&th, &at, copy_worker, (void *)(intptr_t)started) != 0) { pthread_attr_destroy(&at); break; }
        pthread_attr_destroy(&at);
        started++;
    }
    g_copy_threads = started;
}

static void par_memcpy(void *dst, const void *src, size_t n)
{
    if (n < BC_PAR_COPY_MIN) { stream_copy(dst, src, n); return; }
    pthread_mutex_lock(&g_copy_user);
    if (g_copy_threads < 0) copy_pool_start();
    const int w = g_copy_threads;
    if (w == 0) { pthread_mutex_unlock(&g_copy_user); memcpy(dst, src, n); return; }
    /* w + 1 shares of whole 4 KiB pages; the caller takes the last (and the remainder) */
    const size_t share = (n / (size_t)(w + 1)) & ~(size_t)4095;
    pthread_mutex_lock(&g_copy_mu);
    for (int i = 0; i < w; i++) {
        g_copy_jobs[i].dst = (unsigned char *)dst + share * (size_t)i;
        g_copy_jobs[i].src = (const unsigned char *)src + share * (size_t)i;
        g_copy_jobs[i].n = share;
    }
    g_copy_pending = w;
    g_copy_gen++;
    pthread_cond_broadcast(&g_copy_work);
    pthread_mutex_unlock(&g_copy_mu);
    stream_copy((unsigned char *)dst + share * (size_t)w, (const unsigned char *)src + share * (size_t)w, n - share * (size_t)w);
    pthread_mutex_lock(&g_copy_mu);
    while (g_copy_pending) pthread_cond_wait(&g_copy_done, &g_copy_mu);
    pthread_mutex_unlock(&g_copy_mu);
    pthread_mutex_unlock(&g_copy_user);
}

/* test hook (tests/c/host_san_test.c): the staging copy without a device */
void bc_internal_par_memcpy(void *dst, const void *src, size_t n) { par_memcpy(dst, src, n); }

/* ------------------------------------------------------- pipeline plumbing */

static int dev_prepare(bc_dev **out)
{
    int rc = bc_set_device(tls_device);
    if (rc) return rc;
    bc_dev *d = cur_dev_state(NULL);
    if (!d) { bc_set_error("no current CUDA device"); return BC_ERR_NO_DEVICE; }
    *out = d;
    return BC_OK;
}

/* called with d->lock held */
static int dev_ensure_slots(bc_dev *d)
{
    if (d->ready) return BC_OK;
    for (int i = 0; i < BC_SLOTS; i++) {
        bc_slot *s = &d->slot[i];
        BC_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking), "cudaStreamCreate");
        BC_CUDA(cudaEventCreateWithFlags(&s->done, cudaEventDisableTiming), "cudaEventCreate");
        BC_CUDA(cudaMalloc((void **)&s->d_buf, BC_CHUNK_BYTES), "cudaMalloc(pipeline buffer)");
        BC_CUDA(cudaEventCreateWithFlags(&d->e_h2d[i], cudaEventDisableTiming), "cudaEventCreate");
        BC_CUDA(cudaEventCreateWithFlags(&d->e_k[i], cudaEventDisableTiming), "cudaEventCreate");
        BC_CUDA(cudaEventCreateWithFlags(&d->e_d2h[i], cudaEventDisableTiming), "cudaEventCreate");
    }
    BC_CUDA(cudaStreamCreateWithFlags(&d->s_h2d, cudaStreamNonBlocking), "cudaStreamCreate");
    BC_CUDA(cudaStreamCreateWithFlags(&d->s_k, cudaStreamNonBlocking), "cudaStreamCreate");
    BC_CUDA(cudaStreamCreateWithFlags(&d->s_d2h, cudaStreamNonBlocking), "cudaStreamCreate");
    d->ready = 1;
    return BC_OK;
}

static int dev_ensure_aux(bc_dev *d, size_t bytes)
{
    if (d->aux_cap >= bytes) return BC_OK;
    if (d->d_aux) { cudaDeviceSynchronize(); cudaFree(d->d_aux); d->d_aux = NULL; d->aux_cap = 0; }
    size_t cap = bytes < 4096 ? 4096 : bytes;
    BC_CUDA(cudaMalloc((void **)&d->d_aux, cap), "cudaMalloc(aux)");
    d->aux_cap = cap;
    return BC_OK;
}

static int slot_ensure_pinned(bc_slot *s)
{
    if (!s->pin_in)  BC_CUDA(cudaHostAlloc((void **)&s->pin_in,  BC_CHUNK_BYTES, cudaHostAllocDefault), "cudaHostAlloc(staging in)");
    if (!s->pin_out) BC_CUDA(cudaHostAlloc((void **)&s->pin_out, BC_CHUNK_BYTES, cudaHostAllocDefault), "cudaHostAlloc(staging out)");
    return BC_OK;
}

/* 0 = pageable host, 1 = pinned/registered host, 2 = device or managed */
static int classify(const void *p)
{
    struct cudaPointerAttributes at;
    if (!p) return 0;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return 0; }
    if (at.type == cudaMemoryTypeHost) return 1;
    if (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged) return 2;
    return 0;
}

/* ------------------------------------------- registration cache for pageable callers */

/* The reference's harnesses malloc() their buffers once per test and hand the same ranges to the
 * API ten times (aes-modes/test.c:287-350, test.c:60-125).  Pageable memory has to be staged through
 * pinned buffers by host memcpy (about 20 GB/s); pinned memory is DMA'd directly (about 47 GB/s).
 * So a large range that is seen a SECOND time is pinned in place (cudaHostRegister) and kept in a
 * small LRU table; later calls on it take the pinned path.
 *
 * Keeping a registration has two hazards, both from a caller that frees a buffer the cache still
 * holds (the library is not told about free()):
 *  1. a later malloc can hand the same addresses back with different physical pages, and the stale
 *     registration would still point at the OLD pages.  Nothing is assumed about that: the physical
 *     frame number of every page of the range is recorded at registration (/proc/self/pagemap) and
 *     compared again for every use — pinned pages cannot move, so any difference means the range was
 *     unmapped or remapped; the entry is dropped and the call is staged.  The check runs while the
 *     transfer is already in flight (the copies are enqueued, then the host would only wait): if it
 *     fails, what the copy engines read from or wrote to the orphaned old pages is discarded and the
 *     call is redone through staging.  The kernel shows frame numbers only to privileged processes
 *     (CAP_SYS_ADMIN); where it does not, the cache cannot be switched on.
 *  2. until the library next sees those addresses, CUDA still considers them registered, so another
 *     cudaHostAlloc / cudaHostRegister in the process that happens to land there fails with
 *     cudaErrorAlreadyMapped.  This the library cannot detect.
 * Because of 2 the cache is OPT-IN: BULKCIPHER_REGISTER_PAGEABLE=1 in the environment, or
 * bc_register_pageable(1); a host that frees big buffers and also pins memory elsewhere should call
 * bc_register_pageable(0) (drops every idle registration) before it frees them.
 * BULKCIPHER_REGISTER_MIN_MIB (default 64) sets the smallest range worth pinning,
 * BULKCIPHER_REGISTER_MAX_GIB (default 16) caps the pinned total. */
#include <fcntl.h>

#define BC_REG_SLOTS 40
typedef struct {
    uintptr_t lo, hi;          /* byte range [lo, hi) exactly as the caller passed it */
    int state;                 /* 0 free, 1 seen once, 2 registered, 3 registration refused (do not retry) */
    int users;                 /* calls in flight that rely on the registration */
    uint64_t tick;
    uint64_t *pfn;             /* frame number per page of [page(lo), page(hi-1)] at registration */
    size_t npages;
} bc_reg;

static bc_reg g_reg[BC_REG_SLOTS];
static pthread_mutex_t g_reg_mu = PTHREAD_MUTEX_INITIALIZER;
static uint64_t g_reg_tick;
static int g_reg_mode = -1;        /* -1 unknown, 0 off, 1 on (frame numbers readable) */
static int g_reg_fd = -1;
static size_t g_reg_min = (size_t)64 << 20, g_reg_cap = (size_t)16 << 30, g_reg_total;
static long g_page = 4096;
static uint64_t g_reg_stats[4];    /* registrations, hits, stale drops, evictions (bc_internal_reg_stats) */

#define PFN_MASK ((1ull << 55) - 1)

static int read_pfns(uintptr_t lo, size_t npages, uint64_t *out)
{
    const off_t off = (off_t)(lo / (uintptr_t)g_page) * 8;
    size_t got = 0;
    while (got < npages * 8) {
        ssize_t r = pread(g_reg_fd, (char *)out + got, npages * 8 - got, off + (off_t)got);
        if (r <= 0) return -1;
        got += (size_t)r;
    }
    for (size_t i = 0; i < npages; i++) {
        if (!(out[i] >> 63)) return -1;                  /* page not present */
        out[i] &= PFN_MASK;
        if (!out[i]) return -1;                          /* frame number hidden */
    }
    return 0;
}

static int g_reg_capable = -1;     /* frame numbers readable? (probed once) */

static void reg_probe(void)
{
    g_reg_capable = 0;
    const char *e = getenv("BULKCIPHER_REGISTER_MIN_MIB");
    if (e && atol(e) >= 1) g_reg_min = (size_t)atol(e) << 20;
    e = getenv("BULKCIPHER_REGISTER_MAX_GIB");
    if (e && atol(e) >= 1) g_reg_cap = (size_t)atol(e) << 30;
    g_page = sysconf(_SC_PAGESIZE);
    if (g_page <= 0) return;
    g_reg_fd = open("/proc/self/pagemap", O_RDONLY | O_CLOEXEC);
    if (g_reg_fd < 0) return;
    volatile char probe = 1;                             /* a touched page of our own stack */
    uint64_t v;
    if (read_pfns((uintptr_t)&probe, 1, &v) != 0) { close(g_reg_fd); g_reg_fd = -1; return; }
    (void)probe;
    g_reg_capable = 1;
}

static void reg_init(void)
{
    if (g_reg_capable < 0) reg_probe();
    const char *e = getenv("BULKCIPHER_REGISTER_PAGEABLE");
    g_reg_mode = (e && *e == '1' && g_reg_capable) ? 1 : 0;       /* opt-in, see above */
}

/* called with g_reg_mu held; the entry must have no users */
static void reg_drop(bc_reg *r)
{
    if (r->state == 2) {
        cudaError_t e = cudaHostUnregister((void *)r->lo);
        if (e != cudaSuccess) cudaGetLastError();        /* the range may be gone already: nothing to undo */
        g_reg_total -= r->hi - r->lo;
    }
    free(r->pfn);
    memset(r, 0, sizeof *r);
}

/* Returns 1 when [p, p+n) may be DMA'd directly (a reference is then held until reg_release). */
static int reg_acquire(const void *p, size_t n, int pageable)
{
    if (!p || n == 0) return 0;
    pthread_mutex_lock(&g_reg_mu);
    if (g_reg_mode < 0) reg_init();
    if (g_reg_total == 0 && g_reg_mode == 0) { pthread_mutex_unlock(&g_reg_mu); return 0; }   /* off and nothing of ours alive */
    const uintptr_t lo = (uintptr_t)p, hi = lo + n;
    /* A range that straddles one of our registrations (overlaps without lying inside it) would look
     * pinned at one end and pageable at the other: retire that registration first. */
    for (int i = 0; i < BC_REG_SLOTS; i++) {
        bc_reg *r = &g_reg[i];
        if (r->state == 2 && r->users == 0 && r->lo < hi && lo < r->hi && !(r->lo <= lo && hi <= r->hi)) reg_drop(r);
    }
    const uintptr_t plo = lo & ~((uintptr_t)g_page - 1);
    const size_t npages = (size_t)((hi - 1 - plo) / (uintptr_t)g_page) + 1;
    bc_reg *hit = NULL, *same = NULL, *freeslot = NULL, *lru = NULL;
    for (int i = 0; i < BC_REG_SLOTS; i++) {
        bc_reg *r = &g_reg[i];
        if (r->state == 0) { if (!freeslot) freeslot = r; continue; }
        if (r->state == 2 && r->lo <= lo && hi <= r->hi) hit = r;
        if (r->lo == lo && r->hi == hi) same = r;
        if (r->users == 0 && (!lru || r->tick < lru->tick)) lru = r;
    }
    int ok = 0;
    if (hit) {
        hit->users++; hit->tick = ++g_reg_tick;
        pthread_mutex_unlock(&g_reg_mu);
        return 2;            /* ours: held; reg_validate() decides whether it still covers the same pages */
    }
    /* switched off (only registrations still in use by other calls are alive), memory we did not pin
     * ourselves (the caller's own cudaHostAlloc / cudaHostRegister), or too small to be worth it */
    if (g_reg_mode == 0 || !pageable || n < g_reg_min) { pthread_mutex_unlock(&g_reg_mu); return 0; }
    uint64_t *now = NULL;
    if (same && same->state == 1 && !(now = (uint64_t *)malloc(npages * 8))) { pthread_mutex_unlock(&g_reg_mu); return 0; }
    if (same && same->state == 1) {
        /* second sighting: pin it.  Ranges that overlap it without containing it must go first. */
        int blocked = 0;
        for (int i = 0; i < BC_REG_SLOTS; i++) {
            bc_reg *r = &g_reg[i];
            if (r->state == 2 && r->lo < hi && lo < r->hi) { if (r->users) blocked = 1; else reg_drop(r); }
        }
        while (!blocked && g_reg_total + n > g_reg_cap) {
            bc_reg *v = NULL;
            for (int i = 0; i < BC_REG_SLOTS; i++)
                if (g_reg[i].state == 2 && g_reg[i].users == 0 && (!v || g_reg[i].tick < v->tick)) v = &g_reg[i];
            if (!v) { blocked = 1; break; }
            g_reg_stats[3]++;
            reg_drop(v);
        }
        if (!blocked) {
            cudaError_t e = cudaHostRegister((void *)lo, n, cudaHostRegisterPortable);
            if (e == cudaSuccess && read_pfns(plo, npages, now) == 0) {
                same->state = 2; same->pfn = now; same->npages = npages; now = NULL;
                same->users = 1; same->tick = ++g_reg_tick; g_reg_total += n; g_reg_stats[0]++; ok = 1;
            } else {
                if (e == cudaSuccess) cudaHostUnregister((void *)lo); else cudaGetLastError();
                same->state = 3;
            }
        }
    } else if (!same) {
        bc_reg *r = freeslot;
        if (!r && lru) { if (lru->state == 2) g_reg_stats[3]++; reg_drop(lru); r = lru; }
        if (r) { r->lo = lo; r->hi = hi; r->state = 1; r->tick = ++g_reg_tick; }
    } else {
        same->tick = ++g_reg_tick;
    }
    free(now);
    pthread_mutex_unlock(&g_reg_mu);
    return ok;
}

/* Is the held range [p, p+n) still backed by the pages that were pinned?  (Reads the page tables:
 * about 1 ms per GiB — callers overlap it with the transfer.) */
static int reg_validate(const void *p, size_t n)
{
    const uintptr_t lo = (uintptr_t)p, hi = lo + n;
    const uintptr_t plo = lo & ~((uintptr_t)g_page - 1);
    const size_t npages = (size_t)((hi - 1 - plo) / (uintptr_t)g_page) + 1;
    uint64_t *now = (uint64_t *)malloc(npages * 8);
    if (!now) return 0;
    int ok = read_pfns(plo, npages, now) == 0;
    pthread_mutex_lock(&g_reg_mu);
    const bc_reg *hit = NULL;
    for (int i = 0; i < BC_REG_SLOTS; i++)
        if (g_reg[i].state == 2 && g_reg[i].lo <= lo && hi <= g_reg[i].hi) hit = &g_reg[i];
    if (ok && hit) {
        const size_t first = (size_t)((plo - (hit->lo & ~((uintptr_t)g_page - 1))) / (uintptr_t)g_page);
        ok = memcmp(now, hit->pfn + first, npages * 8) == 0;
    } else {
        ok = 0;
    }
    g_reg_stats[ok ? 1 : 2]++;
    pthread_mutex_unlock(&g_reg_mu);
    free(now);
    return ok;
}

/* Give up the hold on a range found stale and retire the registration (once nobody else holds it). */
static void reg_release_stale(const void *p, size_t n)
{
    const uintptr_t lo = (uintptr_t)p, hi = lo + n;
    pthread_mutex_lock(&g_reg_mu);
    for (int i = 0; i < BC_REG_SLOTS; i++) {
        bc_reg *r = &g_reg[i];
        if (r->state == 2 && r->users > 0 && r->lo <= lo && hi <= r->hi) {
            if (--r->users == 0) reg_drop(r);
            break;
        }
    }
    pthread_mutex_unlock(&g_reg_mu);
}

static void reg_release(const void *p, size_t n)
{
    const uintptr_t lo = (uintptr_t)p, hi = lo + n;
    pthread_mutex_lock(&g_reg_mu);
    for (int i = 0; i < BC_REG_SLOTS; i++) {
        bc_reg *r = &g_reg[i];
        if (r->state == 2 && r->users > 0 && r->lo <= lo && hi <= r->hi) {
            if (--r->users == 0 && g_reg_mode == 0) reg_drop(r);       /* the cache was switched off meanwhile */
            break;
        }
    }
    pthread_mutex_unlock(&g_reg_mu);
}

static void reg_shutdown(void)
{
    pthread_mutex_lock(&g_reg_mu);
    for (int i = 0; i < BC_REG_SLOTS; i++) if (g_reg[i].state && g_reg[i].users == 0) reg_drop(&g_reg[i]);
    pthread_mutex_unlock(&g_reg_mu);
}

int bc_register_pageable(int enable)
{
    pthread_mutex_lock(&g_reg_mu);
    if (g_reg_capable < 0) reg_probe();
    g_reg_mode = (enable && g_reg_capable) ? 1 : 0;
    if (!g_reg_mode)
        for (int i = 0; i < BC_REG_SLOTS; i++) if (g_reg[i].state && g_reg[i].users == 0) reg_drop(&g_reg[i]);
    const int on = g_reg_mode;
    pthread_mutex_unlock(&g_reg_mu);
    return on;
}

/* test hook: registrations, validated uses, stale drops, evictions so far; returns 1 if the cache is active */
int bc_internal_reg_stats(uint64_t out[4])
{
    pthread_mutex_lock(&g_reg_mu);
    if (g_reg_mode < 0) reg_init();
    memcpy(out, g_reg_stats, sizeof g_reg_stats);
    const int on = g_reg_mode;
    pthread_mutex_unlock(&g_reg_mu);
    return on;
}

/* The pageable operands of one host call: pin the ones seen before, remember which to release. */
#define BC_HOLD_MAX 6
typedef struct { const void *p[BC_HOLD_MAX]; size_t n[BC_HOLD_MAX]; int held[BC_HOLD_MAX]; int count; } reg_hold;

static void hold_add(reg_hold *h, const void *p, size_t n)
{
    if (!p || !n || h->count >= BC_HOLD_MAX) return;
    for (int k = 0; k < h->count; k++) if (h->p[k] == p && h->n[k] == n) return;      /* in == out */
    const int i = h->count++;
    h->p[i] = p; h->n[i] = n; h->held[i] = 0;
    /* every host range goes through the cache: one of OUR registrations is re-validated before each
     * use (it would classify as pinned even after the caller freed and re-mapped the addresses) */
    const int c = classify(p);
    if (c != 2) h->held[i] = reg_acquire(p, n, c == 0);
}

static __thread int tls_ramp;            /* this thread's current host call accepts chunks of any size (CTR, ECB, XOR) */
static __thread reg_hold *tls_hold;      /* the operands of the host call this thread is inside */

/* Check the held ranges that still await it; returns the number found stale.  Stale ones are
 * released and retired, so that classify() sees plain pageable memory afterwards. */
static int hold_validate(reg_hold *h)
{
    int stale = 0;
    for (int i = 0; h && i < h->count; i++) {
        if (h->held[i] != 2) continue;
        if (reg_validate(h->p[i], h->n[i])) { h->held[i] = 1; continue; }
        reg_release_stale(h->p[i], h->n[i]);
        h->held[i] = 0;
        stale++;
    }
    return stale;
}

static int hold_pending(const reg_hold *h)
{
    for (int i = 0; h && i < h->count; i++) if (h->held[i] == 2) return 1;
    return 0;
}

static void hold_begin(reg_hold *h, const void *a, size_t na, const void *b, size_t nb, const void *c, size_t nc)
{
    h->count = 0;
    tls_hold = h;
    hold_add(h, a, na);
    hold_add(h, b, nb);
    hold_add(h, c, nc);
}

static void hold_end(reg_hold *h)
{
    for (int i = 0; i < h->count; i++) if (h->held[i]) reg_release(h->p[i], h->n[i]);
    tls_hold = NULL;
}

/* wait for a slot's previous chunk and finish its staged copy-out */
static int slot_retire(bc_slot *s)
{
    BC_CUDA(cudaEventSynchronize(s->done), "cudaEventSynchronize");
    if (s->pend_dst) {
        par_memcpy(s->pend_dst, s->pin_out, s->pend_len);
        s->pend_dst = NULL;
    }
    return BC_OK;
}

/* Size of pipeline chunk number `idx`, which starts at byte `off` of an nbytes transfer cut into
 * granules of `chunk` bytes.  Without `ramp`: equal granules.  With it: 1/16, 1/8, 1/4, 1/2 granule,
 * full granules (the last of them shortened to whole 4 KiB pages so that the tail fits exactly), then
 * 1/2, 1/4, 1/8 and 1/16 granule plus any odd bytes.  Every chunk but the last is a multiple of
 * 4096 bytes and none exceeds the granule.  Pure arithmetic (tests/c/host_san_test.c walks it). */
size_t bc_internal_next_chunk(size_t nbytes, size_t off, size_t idx, size_t chunk, int ramp)
{
    size_t n = chunk;
    const size_t left = nbytes - off;
    if (ramp) {
        const size_t q = chunk >> 4, tail_total = chunk - q;                 /* 1/2 + 1/4 + 1/8 + 1/16 */
        if (idx < 4) {
            n = chunk >> (4 - idx);
        } else if (left > tail_total && ((left - tail_total) & ~(size_t)4095) > 0) {
            n = (left - tail_total) & ~(size_t)4095;
            if (n > chunk) n = chunk;
        } else {
            n = chunk >> 1;
            while (n > q && left < 2 * n - q) n >>= 1;
            if (n == q) n = left;
        }
    }
    return n > left ? left : n;
}

/* One chunk's kernel: `d` holds n bytes that start at byte `off` of the caller's buffer. */
typedef int (*chunk_fn)(void *arg, size_t chunk_idx, size_t off, size_t n, void *d, cudaStream_t st);

/* Run `fn` over [in, in+nbytes) in chunks of `chunk` bytes (<= BC_CHUNK_BYTES),
 * each chunk H2D -> fn (in place on the device) -> D2H on its own stream. */
/* Pinned (or absent) input and pinned output: nothing needs the host between chunks, so the
 * whole transfer is enqueued up front on three stage streams — all H2D copies back to back on
 * one, kernels on the second, D2H copies on the third — with events handing each slot from
 * stage to stage and back.  Both copy engines then run without waiting for the host. */
static int run_pipeline_enqueued(bc_dev *d, const unsigned char *in, unsigned char *out, size_t nbytes,
                                 size_t chunk, chunk_fn fn, void *arg, int *stale)
{
    int rc = BC_OK;
    size_t idx = 0;
    /* Ramp (position-independent chunks only, tls_ramp): while the first chunk travels in, nothing
     * travels out, and the other way round at the end — one link direction idles for a whole chunk at
     * each end.  Starting and finishing with chunks of 1/16, 1/8, 1/4, 1/2 of the granule shrinks
     * those two single-direction phases 16-fold. */
    const int ramp = tls_ramp && nbytes >= 3 * chunk && (chunk >> 4) >= ((size_t)1 << 20) && !getenv("BULKCIPHER_NO_RAMP");
    size_t n = 0;
    for (size_t off = 0; off < nbytes && rc == BC_OK; off += n, idx++) {
        n = bc_internal_next_chunk(nbytes, off, idx, chunk, ramp);
        const int sl = (int)(idx % BC_SLOTS);
        bc_slot *s = &d->slot[sl];
        if (in) {
            if (idx >= (size_t)BC_SLOTS) BC_CUDA(cudaStreamWaitEvent(d->s_h2d, d->e_d2h[sl], 0), "cudaStreamWaitEvent");
            BC_CUDA(cudaMemcpyAsync(s->d_buf, in + off, n, cudaMemcpyHostToDevice, d->s_h2d), "cudaMemcpyAsync H2D");
            BC_CUDA(cudaEventRecord(d->e_h2d[sl], d->s_h2d), "cudaEventRecord");
            BC_CUDA(cudaStreamWaitEvent(d->s_k, d->e_h2d[sl], 0), "cudaStreamWaitEvent");
        } else if (idx >= (size_t)BC_SLOTS) {
            BC_CUDA(cudaStreamWaitEvent(d->s_k, d->e_d2h[sl], 0), "cudaStreamWaitEvent");
        }
        if ((rc = fn(arg, idx, off, n, s->d_buf, d->s_k))) break;
        BC_CUDA(cudaEventRecord(d->e_k[sl], d->s_k), "cudaEventRecord");
        BC_CUDA(cudaStreamWaitEvent(d->s_d2h, d->e_k[sl], 0), "cudaStreamWaitEvent");
        BC_CUDA(cudaMemcpyAsync(out + off, s->d_buf, n, cudaMemcpyDeviceToHost, d->s_d2h), "cudaMemcpyAsync D2H");
        BC_CUDA(cudaEventRecord(d->e_d2h[sl], d->s_d2h), "cudaEventRecord");
    }
    /* everything is enqueued and the host would only wait now: the time to check that ranges the
     * registration cache vouched for still cover the pages that were pinned */
    if (stale) *stale = hold_validate(tls_hold);
    cudaError_t e1 = cudaStreamSynchronize(d->s_d2h), e2 = cudaStreamSynchronize(d->s_k), e3 = cudaStreamSynchronize(d->s_h2d);
    if (rc == BC_OK && e1 != cudaSuccess) rc = bc_cuda_fail((int)e1, "cudaStreamSynchronize(D2H)");
    if (rc == BC_OK && e2 != cudaSuccess) rc = bc_cuda_fail((int)e2, "cudaStreamSynchronize(kernels)");
    if (rc == BC_OK && e3 != cudaSuccess) rc = bc_cuda_fail((int)e3, "cudaStreamSynchronize(H2D)");
    return rc;
}

static int run_pipeline(bc_dev *d, const unsigned char *in, unsigned char *out, size_t nbytes,
                        size_t chunk, chunk_fn fn, void *arg)
{
    int in_pinned = in && classify(in) == 1, out_pinned = classify(out) == 1;
    if ((in_pinned || !in) && out_pinned && !getenv("BULKCIPHER_HOST_IN_LOOP")) {
        int stale = 0;
        int rc0 = run_pipeline_enqueued(d, in, out, nbytes, chunk, fn, arg, &stale);
        if (!stale) return rc0;
        /* a registration turned out to be stale (freed and re-mapped addresses): the copy engines
         * worked on orphaned pages, nothing of the caller's was read or written.  The entries are
         * retired; do the whole transfer again through staging (same chunking: batched calls cut
         * their chunks on item boundaries). */
    } else if (hold_pending(tls_hold)) {
        hold_validate(tls_hold);                      /* host-in-the-loop path: check before anything moves */
    }
    in_pinned = in && classify(in) == 1;
    out_pinned = classify(out) == 1;
    int rc = BC_OK;
    size_t idx = 0;
    for (size_t off = 0; off < nbytes && rc == BC_OK; off += chunk, idx++) {
        const size_t n = nbytes - off < chunk ? nbytes - off : chunk;
        bc_slot *s = &d->slot[idx % BC_SLOTS];
        if (idx >= BC_SLOTS && (rc = slot_retire(s))) break;
        cudaError_t e;
        if ((in && !in_pinned) || !out_pinned) { if ((rc = slot_ensure_pinned(s))) break; }
        if (in) {                                  /* in == NULL: the kernel only produces output */
            const unsigned char *src = in + off;
            if (!in_pinned) { par_memcpy(s->pin_in, src, n); src = s->pin_in; }
            e = cudaMemcpyAsync(s->d_buf, src, n, cudaMemcpyHostToDevice, s->stream);
            if (e != cudaSuccess) { rc = bc_cuda_fail((int)e, "cudaMemcpyAsync H2D"); break; }
        }
        if ((rc = fn(arg, idx, off, n, s->d_buf, s->stream))) break;
        unsigned char *dst = out_pinned ? out + off : s->pin_out;
        e = cudaMemcpyAsync(dst, s->d_buf, n, cudaMemcpyDeviceToHost, s->stream);
        if (e != cudaSuccess) { rc = bc_cuda_fail((int)e, "cudaMemcpyAsync D2H"); break; }
        if (!out_pinned) { s->pend_dst = out + off; s->pend_len = n; }
        e = cudaEventRecord(s->done, s->stream);
        if (e != cudaSuccess) { rc = bc_cuda_fail((int)e, "cudaEventRecord"); break; }
    }
    /* drain in issue order so staged copy-outs land even after an error */
    const size_t issued = idx < BC_SLOTS ? idx : BC_SLOTS;
    for (size_t k = 0; k < issued; k++) {
        bc_slot *s = &d->slot[(idx - issued + k) % BC_SLOTS];
        int r2 = slot_retire(s);
        if (rc == BC_OK) rc = r2;
    }
    if (rc != BC_OK) { for (int k = 0; k < BC_SLOTS; k++) { cudaStreamSynchronize(d->slot[k].stream); d->slot[k].pend_dst = NULL; } }
    return rc;
}

/* Chunk size for a transfer of nbytes: the full granule for big buffers, about an eighth of
 * the buffer (whole MiB, at least 4 MiB) for smaller ones so that copies still overlap.
 * Pageable buffers keep the host in the loop (staging copies between the waits), where smaller
 * chunks interleave better: capped at g_pageable_chunk (tools/e2e_probe.py ... pageable). */
static size_t pick_chunk(size_t nbytes, const void *in, const void *out)
{
    size_t cap = BC_CHUNK_BYTES;
    if ((in && classify(in) != 1) || classify(out) != 1) cap = g_pageable_chunk < cap ? g_pageable_chunk : cap;
    if (nbytes >= 8 * cap) return cap;
    size_t c = ((nbytes / 8) + ((size_t)1 << 20) - 1) & ~(((size_t)1 << 20) - 1);
    if (c < ((size_t)4 << 20)) c = (size_t)4 << 20;
    return c < cap ? c : cap;
}

/* The *_host calls take both buffers in host memory (staged pipeline) or both in device memory
 * (launched in place, no staging).  One of each would have the pipeline dereference a device
 * pointer on the CPU: refuse it instead of crashing. */
static int check_same_side(const void *in, const void *out)
{
    const int ci = in ? classify(in) : -1, co = out ? classify(out) : -1;
    if (ci >= 0 && co >= 0 && (ci == 2) != (co == 2)) {
        bc_set_error("mixed buffers: `in` is %s memory but `out` is %s memory; pass both on the same side",
                     ci == 2 ? "device" : "host", co == 2 ? "device" : "host");
        return BC_ERR_BAD_ARG;
    }
    return BC_OK;
}

static int check_host_args(const uint32_t *rk, int nr, const void *in, const void *out, size_t nbytes)
{
    if (!rk || (nr != 10 && nr != 12 && nr != 14)) {
        bc_set_error("bad key schedule (rk=%p, nr=%d; nr must be 10, 12 or 14)", (const void *)rk, nr);
        return BC_ERR_BAD_ARG;
    }
    if (nbytes && (!in || !out)) { bc_set_error("NULL data pointer with nbytes=%zu", nbytes); return BC_ERR_BAD_ARG; }
    return nbytes ? check_same_side(in, out) : BC_OK;
}

/* ------------------------------------------------------------- AES-CTR */

typedef struct {
    const uint32_t *rk; int nr; const unsigned char *counter; int bits; uint64_t first_block;
    size_t total; unsigned char *d_last;
} ctr_op;

static int ctr_chunk(void *arg, size_t idx, size_t off, size_t n, void *d, cudaStream_t st)
{
    ctr_op *o = (ctr_op *)arg;
    (void)idx;
    void *last = (o->d_last && off + n == o->total) ? o->d_last : NULL;
    return bc_aes_ctr_dev(o->rk, o->nr, o->counter, o->bits, o->first_block + off / 16, d, d, n, last, st);
}

int bc_aes_ctr_host(const uint32_t *rk, int nr, const unsigned char counter[16], int counter_bits,
                    uint64_t first_block, const void *in, void *out, size_t nbytes,
                    unsigned char last_keystream[16])
{
    int rc = check_host_args(rk, nr, in, out, nbytes);
    if (rc) return rc;
    if (nbytes == 0) return BC_OK;
    bc_dev *d;
    if ((rc = dev_prepare(&d))) return rc;
    if (classify(in) == 2 && classify(out) == 2) {           /* device-resident: no staging at all */
        pthread_mutex_lock(&d->lock);
        void *d_last = NULL;
        if (last_keystream && (rc = dev_ensure_aux(d, 16)) == BC_OK) d_last = d->d_aux;
        if (rc == BC_OK) rc = bc_aes_ctr_dev(rk, nr, counter, counter_bits, first_block, in, out, nbytes, d_last, NULL);
        if (rc == BC_OK && d_last) rc = bc_memcpy_d2h(last_keystream, d_last, 16);
        if (rc == BC_OK) rc = bc_device_sync();
        pthread_mutex_unlock(&d->lock);
        return rc;
    }
    reg_hold h;
    hold_begin(&h, in, nbytes, out, nbytes, NULL, 0);
    pthread_mutex_lock(&d->lock);
    if ((rc = dev_ensure_slots(d)) == BC_OK && (rc = dev_ensure_aux(d, 16)) == BC_OK) {
        ctr_op o = { rk, nr, counter, counter_bits, first_block, nbytes, last_keystream ? d->d_aux : NULL };
        tls_ramp = 1;
        rc = run_pipeline(d, (const unsigned char *)in, (unsigned char *)out, nbytes, pick_chunk(nbytes, in, out), ctr_chunk, &o);
        tls_ramp = 0;
        if (rc == BC_OK && last_keystream) rc = bc_memcpy_d2h(last_keystream, d->d_aux, 16);
    }
    pthread_mutex_unlock(&d->lock);
    hold_end(&h);
    return rc;
}

/* ------------------------------------------------------------- AES-ECB */

typedef struct { const uint32_t *rk; int nr; int mode; int be32; } ecb_op;

static int ecb_chunk(void *arg, size_t idx, size_t off, size_t n, void *d, cudaStream_t st)
{
    ecb_op *o = (ecb_op *)arg;
    (void)idx; (void)off;
    return (o->be32 ? bc_aes_ecb_be32_dev : bc_aes_ecb_dev)(o->rk, o->nr, o->mode, d, d, n, st);
}

static int ecb_host(int be32, const uint32_t *rk, int nr, int mode, const void *in, void *out, size_t nbytes)
{
    int rc = check_host_args(rk, nr, in, out, nbytes);
    if (rc) return rc;
    nbytes -= nbytes % 16;                                    /* whole blocks only */
    if (nbytes == 0) return BC_OK;
    bc_dev *d;
    if ((rc = dev_prepare(&d))) return rc;
    if (classify(in) == 2 && classify(out) == 2) {
        rc = (be32 ? bc_aes_ecb_be32_dev : bc_aes_ecb_dev)(rk, nr, mode, in, out, nbytes, NULL);
        return rc ? rc : bc_device_sync();
    }
    reg_hold h;
    hold_begin(&h, in, nbytes, out, nbytes, NULL, 0);
    pthread_mutex_lock(&d->lock);
    if ((rc = dev_ensure_slots(d)) == BC_OK) {
        ecb_op o = { rk, nr, mode, be32 };
        tls_ramp = 1;
        rc = run_pipeline(d, (const unsigned char *)in, (unsigned char *)out, nbytes, pick_chunk(nbytes, in, out), ecb_chunk, &o);
        tls_ramp = 0;
    }
    pthread_mutex_unlock(&d->lock);
    hold_end(&h);
    return rc;
}

int bc_aes_ecb_host(const uint32_t *rk, int nr, int mode, const void *in, void *out, size_t nbytes)
{
    return ecb_host(0, rk, nr, mode, in, out, nbytes);
}

int bc_aes_ecb_be32_host(const uint32_t *rk, int nr, int mode, const void *in, void *out, size_t nbytes)
{
    return ecb_host(1, rk, nr, mode, in, out, nbytes);
}

/* ------------------------------------------------------ AES-CBC decrypt */

typedef struct { const uint32_t *rk; int nr; const unsigned char *ivs; int cfb; } cbcd_op;

static int cbcd_chunk(void *arg, size_t idx, size_t off, size_t n, void *d, cudaStream_t st)
{
    cbcd_op *o = (cbcd_op *)arg;
    (void)off;
    return (o->cfb ? bc_aes_cfb128_decrypt_dev : bc_aes_cbc_decrypt_dev)(o->rk, o->nr, o->ivs + 16 * idx, d, d, n, st);
}

static int chained_decrypt_host(int cfb, const uint32_t *rk, int nr, const unsigned char iv[16],
                                const void *in, void *out, size_t nbytes)
{
    int rc = check_host_args(rk, nr, in, out, nbytes);
    if (rc) return rc;
    if (!iv) { bc_set_error("iv is NULL"); return BC_ERR_BAD_ARG; }
    if (nbytes % 16) { bc_set_error("%s length %zu is not a multiple of 16", cfb ? "CFB128" : "CBC", nbytes); return BC_ERR_BAD_ARG; }
    if (nbytes == 0) return BC_OK;
    bc_dev *d;
    if ((rc = dev_prepare(&d))) return rc;
    if (classify(in) == 2 && classify(out) == 2) {
        rc = (cfb ? bc_aes_cfb128_decrypt_dev : bc_aes_cbc_decrypt_dev)(rk, nr, iv, in, out, nbytes, NULL);
        return rc ? rc : bc_device_sync();
    }
    /* Chunk k chains from the last ciphertext block of chunk k-1.  Collect those
     * blocks first: with in == out the pipeline's D2H would overwrite them. */
    reg_hold h;
    hold_begin(&h, in, nbytes, out, nbytes, NULL, 0);
    const size_t chunk = pick_chunk(nbytes, in, out);
    const size_t n_chunks = (nbytes + chunk - 1) / chunk;
    unsigned char *ivs = (unsigned char *)malloc(16 * n_chunks);
    if (!ivs) { hold_end(&h); bc_set_error("out of host memory"); return BC_ERR_BAD_ARG; }
    memcpy(ivs, iv, 16);
    for (size_t k = 1; k < n_chunks; k++) memcpy(ivs + 16 * k, (const unsigned char *)in + k * chunk - 16, 16);
    pthread_mutex_lock(&d->lock);
    if ((rc = dev_ensure_slots(d)) == BC_OK) {
        cbcd_op o = { rk, nr, ivs, cfb };
        rc = run_pipeline(d, (const unsigned char *)in, (unsigned char *)out, nbytes, chunk, cbcd_chunk, &o);
    }
    pthread_mutex_unlock(&d->lock);
    free(ivs);
    hold_end(&h);
    return rc;
}

int bc_aes_cbc_decrypt_host(const uint32_t *rk, int nr, const unsigned char iv[16],
                            const void *in, void *out, size_t nbytes)
{
    return chained_decrypt_host(0, rk, nr, iv, in, out, nbytes);
}

int bc_aes_cfb128_decrypt_host(const uint32_t *rk, int nr, const unsigned char iv[16],
                               const void *in, void *out, size_t nbytes)
{
    return chained_decrypt_host(1, rk, nr, iv, in, out, nbytes);
}

/* Small per-item side data of the batched calls (IVs, keys): moved to / from the device in ONE
 * copy around the pipeline.  Copying it chunk by chunk would queue each small copy behind the
 * data copies already enqueued for later chunks on the same DMA engine and stall the kernels. */
static int side_upload(bc_dev *d, void *key, const void *host, size_t bytes, void **dev)
{
    int rc = bc_scratch_get(key, bytes, dev);
    if (rc) return rc;
    BC_CUDA(cudaMemcpyAsync(*dev, host, bytes, cudaMemcpyHostToDevice, d->s_k), "cudaMemcpyAsync side data");
    BC_CUDA(cudaStreamSynchronize(d->s_k), "cudaStreamSynchronize");
    return BC_OK;
}

static int side_download(bc_dev *d, void *host, const void *dev, size_t bytes)
{
    BC_CUDA(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, d->s_k), "cudaMemcpyAsync side data");
    BC_CUDA(cudaStreamSynchronize(d->s_k), "cudaStreamSynchronize");
    return BC_OK;
}

/* ------------------------------------------- AES-CBC encrypt, batched msgs */

typedef struct {
    const uint32_t *rk; int nr; size_t msg_len; size_t msgs_per_chunk;
    const unsigned char *ivs; unsigned char *ivs_out; unsigned char *d_ivs;   /* [BC_SLOTS][msgs_per_chunk][16] */
    size_t n_msgs; int cfb;
    unsigned char *d_ivs_all;   /* [n_msgs][16] resident for the whole call (in and out), or NULL */
} cbce_op;

static int cbce_chunk(void *arg, size_t idx, size_t off, size_t n, void *d, cudaStream_t st)
{
    cbce_op *o = (cbce_op *)arg;
    const size_t m0 = off / o->msg_len, nm = n / o->msg_len;
    if (o->d_ivs_all) {
        unsigned char *v = o->d_ivs_all + 16 * m0;
        return (o->cfb ? bc_aes_cfb128_encrypt_batch_dev : bc_aes_cbc_encrypt_batch_dev)(
            o->rk, o->nr, nm, o->msg_len, v, o->ivs_out ? v : NULL, d, d, st);
    }
    unsigned char *d_iv = o->d_ivs + (idx % BC_SLOTS) * o->msgs_per_chunk * 16;
    BC_CUDA(cudaMemcpyAsync(d_iv, o->ivs + 16 * m0, 16 * nm, cudaMemcpyHostToDevice, st), "cudaMemcpyAsync ivs");
    int rc = (o->cfb ? bc_aes_cfb128_encrypt_batch_dev : bc_aes_cbc_encrypt_batch_dev)(
        o->rk, o->nr, nm, o->msg_len, d_iv, o->ivs_out ? d_iv : NULL, d, d, st);
    if (rc) return rc;
    if (o->ivs_out)
        BC_CUDA(cudaMemcpyAsync(o->ivs_out + 16 * m0, d_iv, 16 * nm, cudaMemcpyDeviceToHost, st), "cudaMemcpyAsync ivs_out");
    return BC_OK;
}

static int chained_encrypt_batch_host(int cfb, const uint32_t *rk, int nr, size_t n_msgs, size_t msg_len,
                                      const void *ivs, void *ivs_out, const void *in, void *out)
{
    int rc = check_host_args(rk, nr, in, out, n_msgs * msg_len);
    if (rc) return rc;
    if (msg_len % 16) { bc_set_error("message length %zu is not a multiple of 16", msg_len); return BC_ERR_BAD_ARG; }
    if (n_msgs == 0 || msg_len == 0) return BC_OK;
    if (!ivs) { bc_set_error("ivs is NULL"); return BC_ERR_BAD_ARG; }
    bc_dev *d;
    if ((rc = dev_prepare(&d))) return rc;
    if ((classify(in) == 2) != (classify(ivs) == 2) || (ivs_out && (classify(in) == 2) != (classify(ivs_out) == 2))) {
        bc_set_error("mixed buffers: ivs / ivs_out must be on the same side (host or device) as the data");
        return BC_ERR_BAD_ARG;
    }
    if (classify(in) == 2 && classify(out) == 2 && classify(ivs) == 2) {
        rc = (cfb ? bc_aes_cfb128_encrypt_batch_dev : bc_aes_cbc_encrypt_batch_dev)(rk, nr, n_msgs, msg_len, ivs, ivs_out, in, out, NULL);
        return rc ? rc : bc_device_sync();
    }
    reg_hold h;
    hold_begin(&h, in, n_msgs * msg_len, out, n_msgs * msg_len, ivs, 16 * n_msgs);
    hold_add(&h, ivs_out, 16 * n_msgs);
    pthread_mutex_lock(&d->lock);
    if ((rc = dev_ensure_slots(d)) == BC_OK) {
        if (msg_len <= BC_CHUNK_BYTES) {
            cbce_op o = { rk, nr, msg_len, BC_CHUNK_BYTES / msg_len, (const unsigned char *)ivs,
                          (unsigned char *)ivs_out, NULL, n_msgs, cfb, NULL };
            void *all = NULL;
            if ((rc = side_upload(d, (void *)&d->aux_cap, ivs, 16 * n_msgs, &all)) == BC_OK) {
                o.d_ivs_all = (unsigned char *)all;
                rc = run_pipeline(d, (const unsigned char *)in, (unsigned char *)out, n_msgs * msg_len,
                                  o.msgs_per_chunk * msg_len, cbce_chunk, &o);
                if (rc == BC_OK && ivs_out) rc = side_download(d, ivs_out, all, 16 * n_msgs);
            }
        } else {
            /* a message longer than one pipeline buffer: walk it in pieces, the chain value
             * travels through ivs_out of one piece into ivs of the next */
            unsigned char iv[16];
            for (size_t m = 0; m < n_msgs && rc == BC_OK; m++) {
                memcpy(iv, (const unsigned char *)ivs + 16 * m, 16);
                for (size_t off = 0; off < msg_len && rc == BC_OK; off += BC_CHUNK_BYTES) {
                    size_t n = msg_len - off < BC_CHUNK_BYTES ? msg_len - off : BC_CHUNK_BYTES;
                    cbce_op o = { rk, nr, n, 1, iv, iv, NULL, 1, cfb, NULL };
                    if ((rc = dev_ensure_aux(d, BC_SLOTS * 16)) != BC_OK) break;
                    o.d_ivs = d->d_aux;
                    rc = run_pipeline(d, (const unsigned char *)in + m * msg_len + off,
                                      (unsigned char *)out + m * msg_len + off, n, n, cbce_chunk, &o);
                }
                if (ivs_out) memcpy((unsigned char *)ivs_out + 16 * m, iv, 16);
            }
        }
    }
    pthread_mutex_unlock(&d->lock);
    hold_end(&h);
    return rc;
}

int bc_aes_cbc_encrypt_batch_host(const uint32_t *rk, int nr, size_t n_msgs, size_t msg_len,
                                  const void *ivs, void *ivs_out, const void *in, void *out)
{
    return chained_encrypt_batch_host(0, rk, nr, n_msgs, msg_len, ivs, ivs_out, in, out);
}

int bc_aes_cfb128_encrypt_batch_host(const uint32_t *rk, int nr, size_t n_msgs, size_t msg_len,
                                     const void *ivs, void *ivs_out, const void *in, void *out)
{
    return chained_encrypt_batch_host(1, rk, nr, n_msgs, msg_len, ivs, ivs_out, in, out);
}

/* ---------------------------------------------------------- RC4 batched */

#define BC_ARC4_CTX_BYTES 264   /* sizeof(arc4_context): int x, int y, m[256] (arc4.h:35-41) */

int bc_xor_dev(const void *d_a, const void *d_b, void *d_out, size_t nbytes, void *stream);

typedef struct {
    size_t length; unsigned keylen; size_t streams_per_chunk; int keystream_only;
    const unsigned char *keys; const unsigned char *ctx_in; unsigned char *ctx_out;
    unsigned char *d_keys;      /* [BC_SLOTS][streams_per_chunk][keylen] */
    unsigned char *d_ctx;       /* [BC_SLOTS][streams_per_chunk][264] when contexts travel */
    unsigned char *d_keys_all;  /* [n_streams][keylen] resident for the whole call, or NULL */
} rc4_op;

/* streams [s0, s0+ns) of the call, data (if any) already in `d` */
static int rc4_issue(rc4_op *o, size_t slot, size_t s0, size_t ns, void *d, cudaStream_t st)
{
    unsigned char *d_keys = o->d_keys + slot * o->streams_per_chunk * o->keylen;
    unsigned char *d_ctx = o->d_ctx ? o->d_ctx + slot * o->streams_per_chunk * BC_ARC4_CTX_BYTES : NULL;
    if (o->ctx_in)
        BC_CUDA(cudaMemcpyAsync(d_ctx, o->ctx_in + s0 * BC_ARC4_CTX_BYTES, ns * BC_ARC4_CTX_BYTES,
                                cudaMemcpyHostToDevice, st), "cudaMemcpyAsync ctx_in");
    else if (o->d_keys_all)
        d_keys = o->d_keys_all + s0 * o->keylen;
    else
        BC_CUDA(cudaMemcpyAsync(d_keys, o->keys + s0 * o->keylen, ns * o->keylen, cudaMemcpyHostToDevice, st),
                "cudaMemcpyAsync keys");
    int rc = bc_arc4_batch_dev(ns, d_keys, o->keylen, o->length, o->keystream_only ? NULL : d, d,
                               o->ctx_in ? d_ctx : NULL, o->ctx_out ? d_ctx : NULL, st);
    if (rc) return rc;
    if (o->ctx_out)
        BC_CUDA(cudaMemcpyAsync(o->ctx_out + s0 * BC_ARC4_CTX_BYTES, d_ctx, ns * BC_ARC4_CTX_BYTES,
                                cudaMemcpyDeviceToHost, st), "cudaMemcpyAsync ctx_out");
    return BC_OK;
}

static int rc4_chunk(void *arg, size_t idx, size_t off, size_t n, void *d, cudaStream_t st)
{
    rc4_op *o = (rc4_op *)arg;
    return rc4_issue(o, idx % BC_SLOTS, off / o->length, n / o->length, d, st);
}

/* in == NULL asks for the raw keystream (what arc4_prep produces, arc4.c:72-97). */
int bc_arc4_batch_host(size_t n_streams, const void *keys, unsigned keylen, size_t length,
                       const void *in, void *out, const void *ctx_in, void *ctx_out)
{
    if (n_streams == 0) return BC_OK;
    if ((!keys && !ctx_in) || (!ctx_in && (keylen < 1 || keylen > 256))) {
        bc_set_error("RC4 needs keys with 1 <= keylen <= 256 (got %u) or contexts to resume", keylen);
        return BC_ERR_BAD_ARG;
    }
    if (length && !out) { bc_set_error("NULL output pointer"); return BC_ERR_BAD_ARG; }
    bc_dev *d;
    int rc = dev_prepare(&d);
    if (rc) return rc;
    if (!keylen) keylen = 1;
    const int with_ctx = (ctx_in || ctx_out);
    {
        /* all operands in device memory: launch in place, no staging; a mix is refused */
        const void *ops[5] = { keys, in, out, ctx_in, ctx_out };
        int n_dev = 0, n_host = 0;
        for (int i = 0; i < 5; i++) if (ops[i]) { if (classify(ops[i]) == 2) n_dev++; else n_host++; }
        if (n_dev && n_host) {
            bc_set_error("mixed buffers: keys / data / contexts must all be host memory or all device memory");
            return BC_ERR_BAD_ARG;
        }
        if (n_dev) {
            rc = bc_arc4_batch_dev(n_streams, keys, keylen, length, in, out, ctx_in, ctx_out, NULL);
            return rc ? rc : bc_device_sync();
        }
    }
    reg_hold h;
    hold_begin(&h, in, n_streams * length, out, n_streams * length, keys, n_streams * (size_t)keylen);
    hold_add(&h, ctx_in, n_streams * BC_ARC4_CTX_BYTES);
    hold_add(&h, ctx_out, n_streams * BC_ARC4_CTX_BYTES);
    pthread_mutex_lock(&d->lock);
    if ((rc = dev_ensure_slots(d)) == BC_OK) {
        if (length <= BC_CHUNK_BYTES) {
            /* whole streams per chunk */
            size_t spc = length ? BC_CHUNK_BYTES / length : 4096;
            if (spc > 65536) spc = 65536;
            rc4_op o = { length, keylen, spc, in == NULL, (const unsigned char *)keys, (const unsigned char *)ctx_in,
                         (unsigned char *)ctx_out, NULL, NULL, NULL };
            if (!ctx_in) {
                void *all = NULL;
                if ((rc = side_upload(d, (void *)&d->aux_cap, keys, n_streams * keylen, &all)) != BC_OK) goto rc4_done;
                o.d_keys_all = (unsigned char *)all;
            }
            size_t need_keys = (BC_SLOTS * spc * keylen + 255) & ~(size_t)255;
            size_t need_ctx = with_ctx ? BC_SLOTS * spc * BC_ARC4_CTX_BYTES : 0;
            if ((rc = dev_ensure_aux(d, need_keys + need_ctx)) == BC_OK) {
                o.d_keys = d->d_aux;
                o.d_ctx = with_ctx ? d->d_aux + need_keys : NULL;
                if (length) {
                    rc = run_pipeline(d, (const unsigned char *)in, (unsigned char *)out, n_streams * length,
                                      spc * length, rc4_chunk, &o);
                } else {
                    /* key setup only (arc4_setup): nothing streams, the contexts come back */
                    bc_slot *sl = &d->slot[0];
                    for (size_t s0 = 0; s0 < n_streams && rc == BC_OK; s0 += spc) {
                        size_t ns = n_streams - s0 < spc ? n_streams - s0 : spc;
                        rc = rc4_issue(&o, 0, s0, ns, sl->d_buf, sl->stream);
                        if (rc == BC_OK) {
                            cudaError_t e = cudaStreamSynchronize(sl->stream);
                            if (e != cudaSuccess) rc = bc_cuda_fail((int)e, "cudaStreamSynchronize");
                        }
                    }
                }
            }
        } else {
            /* a stream longer than one pipeline buffer: pieces chained through its context */
            unsigned char ctx[BC_ARC4_CTX_BYTES];
            for (size_t s = 0; s < n_streams && rc == BC_OK; s++) {
                int have_ctx = 0;
                if (ctx_in) { memcpy(ctx, (const unsigned char *)ctx_in + s * BC_ARC4_CTX_BYTES, BC_ARC4_CTX_BYTES); have_ctx = 1; }
                for (size_t off = 0; off < length && rc == BC_OK; off += BC_CHUNK_BYTES) {
                    size_t n = length - off < BC_CHUNK_BYTES ? length - off : BC_CHUNK_BYTES;
                    rc4_op o = { n, keylen, 1, in == NULL, keys ? (const unsigned char *)keys + s * keylen : NULL,
                                 have_ctx ? ctx : NULL, ctx, NULL, NULL, NULL };
                    size_t need_keys = (BC_SLOTS * (size_t)keylen + 255) & ~(size_t)255;
                    if ((rc = dev_ensure_aux(d, need_keys + BC_SLOTS * BC_ARC4_CTX_BYTES)) != BC_OK) break;
                    o.d_keys = d->d_aux;
                    o.d_ctx = d->d_aux + need_keys;
                    rc = run_pipeline(d, in ? (const unsigned char *)in + s * length + off : NULL,
                                      (unsigned char *)out + s * length + off, n, n, rc4_chunk, &o);
                    have_ctx = 1;
                }
                if (ctx_out) memcpy((unsigned char *)ctx_out + s * BC_ARC4_CTX_BYTES, ctx, BC_ARC4_CTX_BYTES);
            }
        }
    }
rc4_done:
    pthread_mutex_unlock(&d->lock);
    hold_end(&h);
    return rc;
}

/* out = a ^ b on host buffers (arc4_crypt, arc4.c:101-112).  `a` rides the pipeline
 * buffer, `b` a second device buffer per slot. */
typedef struct { const unsigned char *b; unsigned char *d_b; bc_dev *dev; int b_pageable; } xor_op;

static int xor_chunk(void *arg, size_t idx, size_t off, size_t n, void *d, cudaStream_t st)
{
    xor_op *o = (xor_op *)arg;
    const int sl = (int)(idx % BC_SLOTS);
    unsigned char *d_b = o->d_b + (size_t)sl * BC_CHUNK_BYTES;
    if (st == o->dev->s_k) {
        /* enqueue-ahead pipeline: the second operand rides the H2D stream right behind the first
         * (on the kernel stream it would queue behind the next chunks' copies), an event orders it */
        BC_CUDA(cudaMemcpyAsync(d_b, o->b + off, n, cudaMemcpyHostToDevice, o->dev->s_h2d), "cudaMemcpyAsync b");
        BC_CUDA(cudaEventRecord(o->dev->e_h2d[sl], o->dev->s_h2d), "cudaEventRecord");
        BC_CUDA(cudaStreamWaitEvent(st, o->dev->e_h2d[sl], 0), "cudaStreamWaitEvent");
    } else {
        /* host-in-the-loop pipeline: same stream as the kernel.  A pageable second operand is staged
         * like the first one — through a pinned buffer of the slot, filled by the parallel copy pool
         * (the runtime's own pageable path stages with one thread) — and the slot was retired before
         * this chunk was issued, so its previous copy out of pin_b has completed. */
        const unsigned char *src = o->b + off;
        if (o->b_pageable) {
            bc_slot *s = &o->dev->slot[sl];
            if (!s->pin_b) BC_CUDA(cudaHostAlloc((void **)&s->pin_b, BC_CHUNK_BYTES, cudaHostAllocDefault), "cudaHostAlloc(staging b)");
            par_memcpy(s->pin_b, src, n);
            src = s->pin_b;
        }
        BC_CUDA(cudaMemcpyAsync(d_b, src, n, cudaMemcpyHostToDevice, st), "cudaMemcpyAsync b");
    }
    return bc_xor_dev(d, d_b, d, n, st);
}

int bc_xor_host(const void *a, const void *b, void *out, size_t nbytes)
{
    if (nbytes == 0) return BC_OK;
    if (!a || !b || !out) { bc_set_error("NULL pointer"); return BC_ERR_BAD_ARG; }
    bc_dev *d;
    int rc = dev_prepare(&d);
    if (rc) return rc;
    {
        const int ca = classify(a) == 2, cb = classify(b) == 2, co = classify(out) == 2;
        if (ca != cb || cb != co) {
            bc_set_error("mixed buffers: a, b and out must all be host memory or all device memory");
            return BC_ERR_BAD_ARG;
        }
        if (ca) { rc = bc_xor_dev(a, b, out, nbytes, NULL); return rc ? rc : bc_device_sync(); }
    }
    reg_hold h;
    hold_begin(&h, a, nbytes, b, nbytes, out, nbytes);
    pthread_mutex_lock(&d->lock);
    if ((rc = dev_ensure_slots(d)) == BC_OK) {
        void *d_b = NULL;
        if ((rc = bc_scratch_get((void *)d, BC_SLOTS * BC_CHUNK_BYTES, &d_b)) == BC_OK) {
            xor_op o = { (const unsigned char *)b, (unsigned char *)d_b, d, classify(b) == 0 };
            size_t chunk = pick_chunk(nbytes, a, out);
            if (o.b_pageable && chunk > g_pageable_chunk) chunk = g_pageable_chunk;
            rc = run_pipeline(d, (const unsigned char *)a, (unsigned char *)out, nbytes, chunk, xor_chunk, &o);
        }
    }
    pthread_mutex_unlock(&d->lock);
    hold_end(&h);
    return rc;
}

/* ------------------------------------------------------------- shutdown */

/* Release everything the library owns on every device (pipeline buffers, staging, streams,
 * events, scratch).  Optional: the resources are process-lifetime singletons; this exists for
 * hosts that unload the library or run leak checkers.  No call may be in flight; later calls
 * simply allocate again. */
int bc_shutdown(void)
{
    pthread_once(&g_once, init_locks);
    int prev = -1, rc = BC_OK;
    cudaGetDevice(&prev);
    int n = bc_device_count();
    for (int dev = 0; dev < n && dev < BC_MAX_DEVICES; dev++) {
        bc_dev *d = &g_dev[dev];
        pthread_mutex_lock(&d->lock);
        int touched = d->ready || d->d_aux || d->pool;
        for (int i = 0; i < BC_MAX_SCRATCH; i++) touched |= d->scratch[i].ptr != NULL;
        if (touched && cudaSetDevice(dev) == cudaSuccess) {
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess && rc == BC_OK) rc = bc_cuda_fail((int)e, "cudaDeviceSynchronize");
            if (d->ready) {
                for (int i = 0; i < BC_SLOTS; i++) {
                    bc_slot *s = &d->slot[i];
                    cudaFree(s->d_buf);
                    if (s->pin_in) cudaFreeHost(s->pin_in);
                    if (s->pin_out) cudaFreeHost(s->pin_out);
                    if (s->pin_b) cudaFreeHost(s->pin_b);
                    cudaStreamDestroy(s->stream);
                    cudaEventDestroy(s->done);
                    cudaEventDestroy(d->e_h2d[i]);
                    cudaEventDestroy(d->e_k[i]);
                    cudaEventDestroy(d->e_d2h[i]);
                    memset(s, 0, sizeof *s);
                }
                cudaStreamDestroy(d->s_h2d);
                cudaStreamDestroy(d->s_k);
                cudaStreamDestroy(d->s_d2h);
                d->ready = 0;
            }
            if (d->d_aux) { cudaFree(d->d_aux); d->d_aux = NULL; d->aux_cap = 0; }
            pthread_mutex_lock(&d->scratch_lock);
            if (d->pool) { cudaMemPoolDestroy(d->pool); d->pool = NULL; }
            for (int i = 0; i < BC_MAX_SCRATCH; i++)
                if (d->scratch[i].ptr) { cudaFree(d->scratch[i].ptr); memset(&d->scratch[i], 0, sizeof d->scratch[i]); }
            pthread_mutex_unlock(&d->scratch_lock);
        }
        pthread_mutex_unlock(&d->lock);
    }
    reg_shutdown();
    if (prev >= 0) cudaSetDevice(prev);
    return rc;
}

/* ------------------------------------------------------------ multi-GPU */

/* Host-side locality: put the calling thread on the CPUs of the NUMA node the GPU hangs off and
 * prefer that node for the pages it touches next (pinned staging buffers are allocated by the
 * thread that first needs them).  On single-node boxes, or where sysfs does not say, this is a
 * no-op.  Never fails the caller: locality is an optimisation, not a requirement. */

static int parse_cpulist(const char *s, cpu_set_t *set)
{
    int n = 0;
    CPU_ZERO(set);
    while (*s) {
        char *end;
        long a = strtol(s, &end, 10), b = a;
        if (end == s) break;
        if (*end == '-') { s = end + 1; b = strtol(s, &end, 10); if (end == s) break; }
        for (long c = a; c <= b && c < CPU_SETSIZE; c++) { CPU_SET((int)c, set); n++; }
        s = (*end == ',') ? end + 1 : end;
        if (*end != ',') break;
    }
    return n;
}

int bc_device_numa_node(int device)
{
    char bus[32] = { 0 }, path[128];
    if (cudaDeviceGetPCIBusId(bus, (int)sizeof bus, device) != cudaSuccess) { cudaGetLastError(); return -1; }
    for (char *c = bus; *c; c++) if (*c >= 'A' && *c <= 'Z') *c = (char)(*c - 'A' + 'a');
    snprintf(path, sizeof path, "/sys/bus/pci/devices/%s/numa_node", bus);
    FILE *f = fopen(path, "r");
    if (!f) return -1;
    int node = -1;
    if (fscanf(f, "%d", &node) != 1) node = -1;
    fclose(f);
    return node;
}

int bc_bind_host_thread_to_device(int device)
{
    const int node = bc_device_numa_node(device);
    if (node < 0) return BC_OK;
    char path[128], buf[4096];
    snprintf(path, sizeof path, "/sys/devices/system/node/node%d/cpulist", node);
    FILE *f = fopen(path, "r");
    if (!f) return BC_OK;
    if (!fgets(buf, (int)sizeof buf, f)) { fclose(f); return BC_OK; }
    fclose(f);
    cpu_set_t want, have, both;
    if (parse_cpulist(buf, &want) <= 0) return BC_OK;
    if (sched_getaffinity(0, sizeof have, &have) != 0) return BC_OK;
    CPU_AND(&both, &want, &have);
    if (CPU_COUNT(&both) == 0) return BC_OK;              /* the node's CPUs are outside our cgroup: leave things be */
    sched_setaffinity(0, sizeof both, &both);
#ifdef SYS_set_mempolicy
    if (node < 64) {
        unsigned long mask = 1ul << node;
        syscall(SYS_set_mempolicy, 1 /* MPOL_PREFERRED */, &mask, 65ul);
    }
#endif
    return BC_OK;
}

/* One host thread per GPU: the reference's slice drivers (aes-modes/test.c:28-44,
 * 194-206, 273-285; test.c:44-58) with a device per slice instead of a core per slice.  No
 * collective: every shard is an ordinary bc_*_host call on its own device. */
typedef enum { SH_CTR, SH_ECB, SH_CBC_DEC, SH_CFB_DEC, SH_CBC_ENC, SH_CFB_ENC, SH_RC4 } shard_kind;

typedef struct {
    int device; shard_kind kind; int rc; char err[256];
    const uint32_t *rk; int nr; int mode_or_bits;
    const unsigned char *counter; uint64_t first_block;        /* CTR */
    unsigned char iv[16];                                      /* chained decrypt: this shard's chaining block */
    size_t n_items, item_len; unsigned keylen;                 /* batched ops: messages / streams */
    const unsigned char *side_in; unsigned char *side_out;     /* ivs / keys in, ivs_out out */
    const unsigned char *ctx_in; unsigned char *ctx_out;       /* RC4 contexts */
    const unsigned char *in; unsigned char *out; size_t nbytes;
} shard_job;

static void *shard_main(void *p)
{
    shard_job *j = (shard_job *)p;
    j->rc = bc_set_device(j->device);
    if (j->rc == BC_OK) {
        bc_bind_host_thread_to_device(j->device);
        switch (j->kind) {
        case SH_CTR:
            j->rc = bc_aes_ctr_host(j->rk, j->nr, j->counter, j->mode_or_bits, j->first_block, j->in, j->out, j->nbytes, NULL);
            break;
        case SH_ECB:
            j->rc = bc_aes_ecb_host(j->rk, j->nr, j->mode_or_bits, j->in, j->out, j->nbytes);
            break;
        case SH_CBC_DEC:
            j->rc = bc_aes_cbc_decrypt_host(j->rk, j->nr, j->iv, j->in, j->out, j->nbytes);
            break;
        case SH_CFB_DEC:
            j->rc = bc_aes_cfb128_decrypt_host(j->rk, j->nr, j->iv, j->in, j->out, j->nbytes);
            break;
        case SH_CBC_ENC:
            j->rc = bc_aes_cbc_encrypt_batch_host(j->rk, j->nr, j->n_items, j->item_len, j->side_in, j->side_out, j->in, j->out);
            break;
        case SH_CFB_ENC:
            j->rc = bc_aes_cfb128_encrypt_batch_host(j->rk, j->nr, j->n_items, j->item_len, j->side_in, j->side_out, j->in, j->out);
            break;
        case SH_RC4:
            j->rc = bc_arc4_batch_host(j->n_items, j->side_in, j->keylen, j->item_len, j->in, j->out, j->ctx_in, j->ctx_out);
            break;
        }
    }
    if (j->rc != BC_OK) snprintf(j->err, sizeof j->err, "GPU %d: %s", j->device, bc_last_error());
    return NULL;
}

static int shards_check(int n_gpus)
{
    if (n_gpus < 1 || n_gpus > BC_MAX_DEVICES) { bc_set_error("n_gpus=%d out of range", n_gpus); return BC_ERR_BAD_ARG; }
    int have = bc_device_count();
    if (have < n_gpus) { bc_set_error("asked for %d GPUs, %d visible", n_gpus, have); return BC_ERR_NO_DEVICE; }
    return BC_OK;
}

/* Run jobs[0..n) on devices 0..n-1, one thread each; shards with nothing to do are skipped. */
static int shards_run(shard_job *jobs, int n)
{
    pthread_t th[BC_MAX_DEVICES];
    int started[BC_MAX_DEVICES] = { 0 }, rc = BC_OK;
    for (int g = 0; g < n; g++) {
        jobs[g].device = g; jobs[g].rc = BC_OK; jobs[g].err[0] = 0;
        const int empty = (jobs[g].kind == SH_CBC_ENC || jobs[g].kind == SH_CFB_ENC || jobs[g].kind == SH_RC4)
                              ? jobs[g].n_items == 0 : jobs[g].nbytes == 0;
        if (empty) continue;
        if (pthread_create(&th[g], NULL, shard_main, &jobs[g]) != 0) { bc_set_error("pthread_create failed"); rc = BC_ERR_BAD_ARG; break; }
        started[g] = 1;
    }
    for (int g = 0; g < n; g++) {
        if (!started[g]) continue;
        pthread_join(th[g], NULL);
        if (jobs[g].rc != BC_OK && rc == BC_OK) { rc = jobs[g].rc; bc_set_error("%s", jobs[g].err); }
    }
    return rc;
}

/* Equal contiguous ranges of `total` items, remainder to the last GPU (the reference's slice
 * rule, aes-modes/test.c:33-35). */
static void shard_range(size_t total, int n, int g, size_t *first, size_t *count)
{
    const size_t per = total / (size_t)n;
    *first = per * (size_t)g;
    *count = (g == n - 1) ? total - *first : per;
}

/* CTR: unlike the reference harness, which restarts the counter on every slice
 * (aes-modes/test.c:273-285), shard g starts its counter at its first block index. */
int bc_aes_ctr_host_multi(int n_gpus, const uint32_t *rk, int nr, const unsigned char counter[16],
                          int counter_bits, const void *in, void *out, size_t nbytes)
{
    int rc = shards_check(n_gpus);
    if (rc) return rc;
    shard_job jobs[BC_MAX_DEVICES];
    memset(jobs, 0, sizeof jobs);
    const size_t total_blocks = (nbytes + 15) / 16;
    for (int g = 0; g < n_gpus; g++) {
        size_t b0, nb;
        shard_range(total_blocks, n_gpus, g, &b0, &nb);
        const size_t byte0 = b0 * 16;
        const size_t byte1 = (g == n_gpus - 1) ? nbytes : (b0 + nb) * 16;
        shard_job *j = &jobs[g];
        j->kind = SH_CTR; j->rk = rk; j->nr = nr; j->counter = counter; j->mode_or_bits = counter_bits;
        j->first_block = b0;
        j->in = (const unsigned char *)in + byte0;
        j->out = (unsigned char *)out + byte0;
        j->nbytes = byte1 > byte0 ? byte1 - byte0 : 0;
    }
    return shards_run(jobs, n_gpus);
}

/* ECB: whole blocks, contiguous block ranges (aes-modes/test.c:28-44, :194-206). */
int bc_aes_ecb_host_multi(int n_gpus, const uint32_t *rk, int nr, int mode, const void *in, void *out, size_t nbytes)
{
    int rc = shards_check(n_gpus);
    if (rc) return rc;
    shard_job jobs[BC_MAX_DEVICES];
    memset(jobs, 0, sizeof jobs);
    for (int g = 0; g < n_gpus; g++) {
        size_t b0, nb;
        shard_range(nbytes / 16, n_gpus, g, &b0, &nb);
        shard_job *j = &jobs[g];
        j->kind = SH_ECB; j->rk = rk; j->nr = nr; j->mode_or_bits = mode;
        j->in = (const unsigned char *)in + b0 * 16;
        j->out = (unsigned char *)out + b0 * 16;
        j->nbytes = nb * 16;
    }
    return shards_run(jobs, n_gpus);
}

/* CBC / CFB128 decrypt: shard g > 0 chains from the last ciphertext block of shard g-1
 * (aes.c:784-797: P_i = D(C_i) ^ C_{i-1}), which is input data — read here, before any shard
 * runs, because with in == out the neighbour's output would overwrite it. */
static int chained_decrypt_host_multi(int cfb, int n_gpus, const uint32_t *rk, int nr, const unsigned char iv[16],
                                      const void *in, void *out, size_t nbytes)
{
    int rc = shards_check(n_gpus);
    if (rc) return rc;
    if (!iv) { bc_set_error("iv is NULL"); return BC_ERR_BAD_ARG; }
    if (nbytes % 16) { bc_set_error("%s length %zu is not a multiple of 16", cfb ? "CFB128" : "CBC", nbytes); return BC_ERR_BAD_ARG; }
    if (nbytes && (!in || !out)) { bc_set_error("NULL data pointer with nbytes=%zu", nbytes); return BC_ERR_BAD_ARG; }
    if (nbytes && (classify(in) == 2 || classify(out) == 2)) {
        bc_set_error("the *_host_multi calls take HOST buffers (a device buffer lives on one GPU)");
        return BC_ERR_BAD_ARG;
    }
    shard_job jobs[BC_MAX_DEVICES];
    memset(jobs, 0, sizeof jobs);
    for (int g = 0; g < n_gpus; g++) {
        size_t b0, nb;
        shard_range(nbytes / 16, n_gpus, g, &b0, &nb);
        shard_job *j = &jobs[g];
        j->kind = cfb ? SH_CFB_DEC : SH_CBC_DEC; j->rk = rk; j->nr = nr;
        if (nb) memcpy(j->iv, b0 == 0 ? iv : (const unsigned char *)in + b0 * 16 - 16, 16);
        j->in = (const unsigned char *)in + b0 * 16;
        j->out = (unsigned char *)out + b0 * 16;
        j->nbytes = nb * 16;
    }
    return shards_run(jobs, n_gpus);
}

int bc_aes_cbc_decrypt_host_multi(int n_gpus, const uint32_t *rk, int nr, const unsigned char iv[16],
                                  const void *in, void *out, size_t nbytes)
{
    return chained_decrypt_host_multi(0, n_gpus, rk, nr, iv, in, out, nbytes);
}

int bc_aes_cfb128_decrypt_host_multi(int n_gpus, const uint32_t *rk, int nr, const unsigned char iv[16],
                                     const void *in, void *out, size_t nbytes)
{
    return chained_decrypt_host_multi(1, n_gpus, rk, nr, iv, in, out, nbytes);
}

/* Batched CBC / CFB128 encrypt: by message index; a single message does not shard. */
static int chained_encrypt_batch_host_multi(int cfb, int n_gpus, const uint32_t *rk, int nr, size_t n_msgs, size_t msg_len,
                                            const void *ivs, void *ivs_out, const void *in, void *out)
{
    int rc = shards_check(n_gpus);
    if (rc) return rc;
    shard_job jobs[BC_MAX_DEVICES];
    memset(jobs, 0, sizeof jobs);
    for (int g = 0; g < n_gpus; g++) {
        size_t m0, nm;
        shard_range(n_msgs, n_gpus, g, &m0, &nm);
        shard_job *j = &jobs[g];
        j->kind = cfb ? SH_CFB_ENC : SH_CBC_ENC; j->rk = rk; j->nr = nr;
        j->n_items = nm; j->item_len = msg_len;
        j->side_in = ivs ? (const unsigned char *)ivs + 16 * m0 : NULL;
        j->side_out = ivs_out ? (unsigned char *)ivs_out + 16 * m0 : NULL;
        j->in = (const unsigned char *)in + m0 * msg_len;
        j->out = (unsigned char *)out + m0 * msg_len;
    }
    return shards_run(jobs, n_gpus);
}

int bc_aes_cbc_encrypt_batch_host_multi(int n_gpus, const uint32_t *rk, int nr, size_t n_msgs, size_t msg_len,
                                        const void *ivs, void *ivs_out, const void *in, void *out)
{
    return chained_encrypt_batch_host_multi(0, n_gpus, rk, nr, n_msgs, msg_len, ivs, ivs_out, in, out);
}

int bc_aes_cfb128_encrypt_batch_host_multi(int n_gpus, const uint32_t *rk, int nr, size_t n_msgs, size_t msg_len,
                                           const void *ivs, void *ivs_out, const void *in, void *out)
{
    return chained_encrypt_batch_host_multi(1, n_gpus, rk, nr, n_msgs, msg_len, ivs, ivs_out, in, out);
}

/* Batched RC4: by stream index (test.c:44-58 slices ONE stream's XOR; whole streams are the unit
 * here because a single RC4 stream is a serial chain). */
int bc_arc4_batch_host_multi(int n_gpus, size_t n_streams, const void *keys, unsigned keylen, size_t length,
                             const void *in, void *out, const void *ctx_in, void *ctx_out)
{
    int rc = shards_check(n_gpus);
    if (rc) return rc;
    shard_job jobs[BC_MAX_DEVICES];
    memset(jobs, 0, sizeof jobs);
    for (int g = 0; g < n_gpus; g++) {
        size_t s0, ns;
        shard_range(n_streams, n_gpus, g, &s0, &ns);
        shard_job *j = &jobs[g];
        j->kind = SH_RC4; j->n_items = ns; j->item_len = length; j->keylen = keylen;
        j->side_in = keys ? (const unsigned char *)keys + s0 * keylen : NULL;
        j->ctx_in = ctx_in ? (const unsigned char *)ctx_in + s0 * BC_ARC4_CTX_BYTES : NULL;
        j->ctx_out = ctx_out ? (unsigned char *)ctx_out + s0 * BC_ARC4_CTX_BYTES : NULL;
        j->in = in ? (const unsigned char *)in + s0 * length : NULL;
        j->out = out ? (unsigned char *)out + s0 * length : NULL;
    }
    return shards_run(jobs, n_gpus);
}
