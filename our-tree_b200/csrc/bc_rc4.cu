/*
 * bc_rc4.cu — batched RC4 / ARC4 for sm_100a: many independent key streams,
 * one thread per stream, KSA + PRGA + XOR fused, each stream's 256-byte
 * permutation resident in shared memory.  Also the two-buffer XOR that
 * arc4_crypt is (arc4.c:101-112).
 *
 * Follows the reference's arc4.c (KSA :43-67, PRGA :72-97) = rc4.c's intended
 * semantics (rc4.c:56-99); bit-exact against oracle/ in tests/.
 *
 * Layout: a CTA of 256 threads owns 64 KB of shared memory viewed as 64 rows of
 * 256 words; byte i of thread t's permutation is byte (i&3) of word
 * [i>>2][t].  Thread t therefore only ever touches bank t%32: every warp-wide
 * LDS/STS is one conflict-free wavefront although the indices are data-
 * dependent.  The address of byte i is ((i*0x101) & 0xFC03) | t*4 — one PRMT
 * (which also reduces i mod 256 for free) and one LOP3.
 * Three CTAs (768 streams, 192 KB) are resident per SM.
 *
 * Roofline: 5 shared accesses per keystream byte (3 loads, 2 stores) on a pipe
 * that serves 32 lanes per clock per SM -> 6.4 bytes/clk/SM; HBM needs only
 * 2 bytes of traffic per byte.  The issue bound binds.
 */
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "bulkcipher.h"
#include "bc_internal.h"
#include "bc_tma.cuh"

namespace {

constexpr int kRc4Threads = 256;
constexpr int kRc4Smem    = kRc4Threads * 256;   /* 64 KB */
constexpr int kCtxBytes   = 264;                 /* arc4_context: int x, int y, m[256] (arc4.h:35-41) */

struct Rc4Args {
    uint64_t n_streams;
    const unsigned char *keys;
    uint32_t keylen;
    uint64_t length;
    const unsigned char *in;     /* NULL: emit the raw keystream (arc4_prep) */
    unsigned char *out;
    const unsigned char *ctx_in; /* NULL: run the KSA from keys */
    unsigned char *ctx_out;
};

/* `base` = shared-window address of the CTA's array (a ptxas-time constant that
 * folds into the LDS/STS immediate); `t4` = thread*4 selects the column. */
struct Sm { uint32_t base, t4; };

__device__ __forceinline__ uint32_t sbox_addr(uint32_t i, const Sm sm)
{
    /* PRMT copies byte 0 of i into bytes 0 and 1 (so bits above 7 of i are ignored) */
    return sm.base + ((__byte_perm(i, 0, 0x4400) & 0xFC03u) | sm.t4);
}

__device__ __forceinline__ uint32_t lds8(uint32_t a)
{
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}

__device__ __forceinline__ void sts8(uint32_t a, uint32_t v)
{
    asm volatile("st.shared.u8 [%0], %1;" :: "r"(a), "r"(v) : "memory");
}

__device__ __forceinline__ void sts32(uint32_t a, uint32_t v)
{
    asm volatile("st.shared.u32 [%0], %1;" :: "r"(a), "r"(v) : "memory");
}

__device__ __forceinline__ uint32_t lds32u(uint32_t a)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}

/* one PRGA step (arc4.c:82-91); returns the keystream byte */
__device__ __forceinline__ uint32_t prga_step(uint32_t &x, uint32_t &y, const Sm sbase)
{
    x = x + 1;
    const uint32_t ax = sbox_addr(x, sbase);
    const uint32_t a = lds8(ax);
    y = y + a;
    const uint32_t ay = sbox_addr(y, sbase);
    const uint32_t b = lds8(ay);
    sts8(ax, b);
    sts8(ay, a);
    return lds8(sbox_addr(a + b, sbase));
}

template <bool VEC, bool KEY16>
__global__ void __launch_bounds__(kRc4Threads, 3)
rc4_batch_kernel(const __grid_constant__ Rc4Args a)
{
    extern __shared__ __align__(1024) uint32_t smem[];
    /* shared-window address of this thread's column: row stride is 1024 bytes */
    const Sm sbase = { (uint32_t)__cvta_generic_to_shared(smem), threadIdx.x * 4u };
    const uint32_t col = sbase.base + sbase.t4;
    const uint64_t n_threads = (uint64_t)gridDim.x * kRc4Threads;

    for (uint64_t s = (uint64_t)blockIdx.x * kRc4Threads + threadIdx.x; s < a.n_streams; s += n_threads) {
        uint32_t x = 0, y = 0;
        if (a.ctx_in) {
            /* resume: x, y, m[] exactly as arc4_prep left them (arc4.c:93-94) */
            const unsigned char *c = a.ctx_in + s * kCtxBytes;
            x = *(const uint32_t *)c;
            y = *(const uint32_t *)(c + 4);
            const uint2 *m = (const uint2 *)(c + 8);
            for (int w = 0; w < 32; w++) {
                const uint2 v = m[w];
                sts32(col + (2 * w) * 1024u, v.x);
                sts32(col + (2 * w + 1) * 1024u, v.y);
            }
        } else {
            /* KSA (arc4.c:43-67): identity, then 256 swaps driven by the key */
            for (int w = 0; w < 64; w++) sts32(col + w * 1024u, 0x03020100u + 0x04040404u * w);
            uint32_t j = 0;
            if (KEY16) {
                const uint4 kv = *(const uint4 *)(a.keys + s * 16);
                const uint32_t kw[4] = { kv.x, kv.y, kv.z, kv.w };
                for (int i0 = 0; i0 < 256; i0 += 16) {
#pragma unroll
                    for (int e = 0; e < 16; e++) {
                        const uint32_t ai = sbox_addr(i0 + e, sbase);
                        const uint32_t v = lds8(ai);
                        j = j + v + ((kw[e >> 2] >> (8 * (e & 3))) & 0xffu);
                        const uint32_t aj = sbox_addr(j, sbase);
                        const uint32_t u = lds8(aj);
                        sts8(ai, u);
                        sts8(aj, v);
                    }
                }
            } else {
                const unsigned char *key = a.keys + s * a.keylen;
                uint32_t k = 0;
                for (int i = 0; i < 256; i++) {
                    if (k >= a.keylen) k = 0;                       /* arc4.c:60 */
                    const uint32_t ai = sbox_addr(i, sbase);
                    const uint32_t v = lds8(ai);
                    j = j + v + key[k++];
                    const uint32_t aj = sbox_addr(j, sbase);
                    const uint32_t u = lds8(aj);
                    sts8(ai, u);
                    sts8(aj, v);
                }
            }
        }

        /* PRGA + XOR */
        const unsigned char *pin = a.in ? a.in + s * a.length : nullptr;
        unsigned char *pout = a.out + s * a.length;
        uint64_t pos = 0;
        if (VEC) {
            for (; pos + 16 <= a.length; pos += 16) {
                uint32_t w[4];
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    uint32_t acc = prga_step(x, y, sbase);
                    acc |= prga_step(x, y, sbase) << 8;
                    acc |= prga_step(x, y, sbase) << 16;
                    acc |= prga_step(x, y, sbase) << 24;
                    w[q] = acc;
                }
                uint4 p = make_uint4(0, 0, 0, 0);
                if (pin) p = *(const uint4 *)(pin + pos);
                p.x ^= w[0]; p.y ^= w[1]; p.z ^= w[2]; p.w ^= w[3];
                *(uint4 *)(pout + pos) = p;
            }
        }
        for (; pos < a.length; pos++) {
            const uint32_t k = prga_step(x, y, sbase);
            pout[pos] = (unsigned char)((pin ? pin[pos] : 0) ^ k);
        }

        if (a.ctx_out) {
            unsigned char *c = a.ctx_out + s * kCtxBytes;
            *(uint32_t *)c = x & 0xffu;
            *(uint32_t *)(c + 4) = y & 0xffu;
            uint2 *m = (uint2 *)(c + 8);
            for (int w = 0; w < 32; w++)
                m[w] = make_uint2(lds32u(col + (2 * w) * 1024u), lds32u(col + (2 * w + 1) * 1024u));
        }
    }
}

/* ======================================================================
 * TMA-staged variant (the fast path: length % 16 == 0, 16-byte aligned buffers)
 *
 * The plain kernel above lets every lane read/write its own stream with 16-byte
 * global accesses: one LDG.128 touches 32 different lines and costs 32 data-pipe
 * wavefronts (profiles/r1_summary.md: 40 % of the pipe).  Here the stream data
 * moves as 2-D TMA boxes of 32 streams x 64 bytes ([n_streams][length] u8
 * tensor, SWIZZLE_64B), one bulk-tensor load and one bulk-tensor store per warp
 * per 64 keystream bytes, double buffered.  A lane then reaches its own row
 * with four LDS.128 / STS.128 that the swizzle keeps conflict free, XORs the
 * keystream in place and the box goes back with a TMA store.  Rows past
 * n_streams and bytes past `length` are clipped by the TMA unit.
 *
 * Permutation layout (64 KB per 256-thread CTA): byte i of thread t lives at
 * i*256 + b0(t), b0 = (t>>7)*128 + (t&31)*4 + ((t>>5)&3): the four warps of a
 * half-CTA share words, lane L still only touches bank L, and the address is a
 * single PRMT (index byte -> address byte 1, b0 -> byte 0).
 * The m[x+1] load of the next step is issued one step early and patched when
 * this step's swap lands on it, which halves the dependent chain per byte.
 * ==================================================================== */

/* How a warp's 32 x 64-byte box meets the data: IO_LOAD = TMA-load the input box, XOR in shared memory,
 * TMA-store (out of place); IO_XOR = in == out, store the keystream box and let the TMA unit XOR it into
 * HBM; IO_KS = no input, the keystream itself is the output (arc4_prep).  A template parameter, so each
 * case is straight-line code (round 2: shared-memory reductions instead of LDS.128 + STS.128 for IO_LOAD
 * were tried and lose, 1 269 vs 1 354 GB/s). */
enum { IO_LOAD = 0, IO_XOR = 1, IO_KS = 2 };
#ifndef BC_RC4_CUNROLL
#define BC_RC4_CUNROLL 4      /* 16-byte chunks of a 64-byte slab unrolled in the PRGA loop (A/B knob) */
#endif
constexpr int kCUnroll = BC_RC4_CUNROLL;
constexpr int kTmaWarps    = kRc4Threads / 32;
constexpr int kTmaStageOff = kRc4Smem;              /* [warp][2][kBoxBytes] */
constexpr int kTmaBarOff   = kTmaStageOff + kTmaWarps * 2 * kBoxBytes;
constexpr int kTmaSmem     = kTmaBarOff + kTmaWarps * 2 * 8;

struct Rc4TmaArgs {
    uint64_t n_streams;
    const unsigned char *keys;
    uint32_t keylen;
    uint64_t length;
    uint32_t has_in;
    uint32_t xor_in_place;       /* in == out: keystream boxes are XORed into global memory by the TMA unit */
    const unsigned char *ctx_in;
    unsigned char *ctx_out;
};

__device__ __forceinline__ uint32_t paddr(uint32_t i, uint32_t b0) { return __byte_perm(i, b0, 0x5504); }

/* PRGA step with the next step's m[x+1] in flight (arc4.c:82-91).
 * ax = address of m[x] for THIS step (already advanced), a = m[x]. */
struct Prga { uint32_t x, y, a, ax; };

__device__ __forceinline__ uint32_t prga_next(Prga &p, uint32_t base, uint32_t b0)
{
    const uint32_t a = p.a, ax = p.ax;
    p.y += a;
    const uint32_t ay = base + paddr(p.y, b0);
    p.x += 1;                                        /* index of the NEXT step */
    const uint32_t axn = base + paddr(p.x, b0);
    uint32_t an = lds8(axn);                         /* m[x+1], one step early */
    const uint32_t b = lds8(ay);
    sts8(ax, b);
    sts8(ay, a);
    if (ay == axn) an = a;                           /* this step's swap just wrote that slot */
    const uint32_t k = lds8(base + paddr(a + b, b0));
    p.a = an;
    p.ax = axn;
    return k;
}

template <bool KEY16, int IO>
__global__ void __launch_bounds__(kRc4Threads, 2)
rc4_batch_tma_kernel(const __grid_constant__ Rc4TmaArgs a, const __grid_constant__ CUtensorMap tm_in,
                     const __grid_constant__ CUtensorMap tm_out)
{
    extern __shared__ __align__(1024) uint32_t smem[];
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(smem);
    const uint32_t t = threadIdx.x, lane = t & 31u, warp = t >> 5;
    const uint32_t b0 = ((t >> 7) << 7) | (lane << 2) | (warp & 3u);
    const uint32_t stage = base + kTmaStageOff + warp * (2 * kBoxBytes);
    const uint32_t bar = base + kTmaBarOff + warp * 16;
    /* this lane's row inside a box, SWIZZLE_64B: 16-byte chunk c sits at chunk c ^ ((row>>1)&3) */
    const uint32_t row_off = lane * kSlab, swz = (lane >> 1) & 3u;

    if (lane == 0) { mbar_init(bar, 1); mbar_init(bar + 8, 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();

    uint32_t phases = 0u;                            /* bit b = parity to wait for on buffer b */
    const uint64_t n_groups = (a.n_streams + 31) / 32;
    const uint64_t n_warps = (uint64_t)gridDim.x * kTmaWarps;
    const uint32_t n_slabs = (uint32_t)((a.length + kSlab - 1) / kSlab);

    for (uint64_t sg = blockIdx.x + (uint64_t)gridDim.x * warp; sg < n_groups; sg += n_warps) {   /* round-robin over CTAs first */
        const uint64_t s = sg * 32 + lane;
        const bool live = s < a.n_streams;
        const uint32_t row0 = (uint32_t)(sg * 32);

        /* first box on its way while the key schedule runs */
        if (IO == IO_LOAD && n_slabs && lane == 0) {
            mbar_expect_tx(bar, kBoxBytes);
            tma_load_2d(stage, &tm_in, 0, row0, bar);
        }

        Prga p = { 0u, 0u, 0u, 0u };
        if (a.ctx_in) {
            /* resume: x, y, m[] exactly as arc4_prep left them (arc4.c:93-94); idle lanes get an identity */
            const unsigned char *c = a.ctx_in + (live ? s : 0) * kCtxBytes;
            if (live) { p.x = *(const uint32_t *)c; p.y = *(const uint32_t *)(c + 4); }
            for (int i = 0; i < 256; i++) sts8(base + paddr(i, b0), live ? (uint32_t)c[8 + i] : (uint32_t)i);
        } else {
            for (int i = 0; i < 256; i++) sts8(base + paddr(i, b0), i);
            uint32_t j = 0;
            if (KEY16) {
                uint4 kv = make_uint4(0, 0, 0, 0);
                if (live) kv = *(const uint4 *)(a.keys + s * 16);
                const uint32_t kw[4] = { kv.x, kv.y, kv.z, kv.w };
                for (int i0 = 0; i0 < 256; i0 += 16) {
#pragma unroll
                    for (int e = 0; e < 16; e++) {
                        const uint32_t ai = base + paddr(i0 + e, b0);
                        const uint32_t v = lds8(ai);
                        j = j + v + ((kw[e >> 2] >> (8 * (e & 3))) & 0xffu);
                        const uint32_t aj = base + paddr(j, b0);
                        const uint32_t u = lds8(aj);
                        sts8(ai, u);
                        sts8(aj, v);
                    }
                }
            } else {
                const unsigned char *key = a.keys + (live ? s : 0) * a.keylen;
                uint32_t k = 0;
                for (int i = 0; i < 256; i++) {
                    if (k >= a.keylen) k = 0;
                    const uint32_t ai = base + paddr(i, b0);
                    const uint32_t v = lds8(ai);
                    j = j + v + key[k++];
                    const uint32_t aj = base + paddr(j, b0);
                    const uint32_t u = lds8(aj);
                    sts8(ai, u);
                    sts8(aj, v);
                }
            }
        }
        /* prime the one-step-ahead load: m[x+1] */
        p.x += 1;
        p.ax = base + paddr(p.x, b0);
        p.a = lds8(p.ax);

        for (uint32_t k = 0; k < n_slabs; k++) {
            const uint32_t buf = k & 1u;
            const uint32_t sbuf = stage + buf * kBoxBytes;
            if (lane == 0 && k + 1 < n_slabs) {
                tma_store_wait_read();               /* the other buffer's last store has left shared memory */
                if (IO == IO_LOAD) {
                    mbar_expect_tx(bar + 8 * (buf ^ 1u), kBoxBytes);
                    tma_load_2d(stage + (buf ^ 1u) * kBoxBytes, &tm_in, (k + 1) * kSlab, row0, bar + 8 * (buf ^ 1u));
                }
            }
            __syncwarp();
            if (IO == IO_LOAD) { mbar_wait(bar + 8 * buf, (phases >> buf) & 1u); phases ^= 1u << buf; }
            const uint32_t valid = (uint32_t)(a.length - (uint64_t)k * kSlab < kSlab ? a.length - (uint64_t)k * kSlab : kSlab);
#pragma unroll (kCUnroll)
            for (uint32_t c = 0; c < 4; c++) {
                if (c * 16 < valid) {                /* length is a multiple of 16: whole chunks only */
                    uint32_t w[4];
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        uint32_t acc = prga_next(p, base, b0);
                        acc = __byte_perm(acc, prga_next(p, base, b0), 0x3240);
                        acc = __byte_perm(acc, prga_next(p, base, b0), 0x3410);
                        acc = __byte_perm(acc, prga_next(p, base, b0), 0x4210);
                        w[q] = acc;
                    }
                    const uint32_t sa = sbuf + row_off + ((c ^ swz) << 4);
                    if (IO == IO_LOAD) {
                        uint4 d = lds128(sa);
                        d.x ^= w[0]; d.y ^= w[1]; d.z ^= w[2]; d.w ^= w[3];
                        sts128(sa, d);
                    } else {
                        sts128(sa, make_uint4(w[0], w[1], w[2], w[3]));      /* keystream only (IO_XOR: the TMA unit XORs it into HBM) */
                    }
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                if (IO == IO_XOR) tma_reduce_xor_2d(&tm_out, k * (kSlab / 4), row0, sbuf);
                else tma_store_2d(&tm_out, k * kSlab, row0, sbuf);
            }
        }
        if (lane == 0) tma_store_wait_read();        /* both buffers reusable by the next stream group */
        __syncwarp();

        if (a.ctx_out && live) {
            /* undo the one-step lookahead: the context holds the x of the last byte produced */
            unsigned char *c = a.ctx_out + s * kCtxBytes;
            *(uint32_t *)c = (p.x - 1) & 0xffu;
            *(uint32_t *)(c + 4) = p.y & 0xffu;
            for (int i = 0; i < 256; i++) c[8 + i] = (unsigned char)lds8(base + paddr(i, b0));
        }
    }
    if (lane == 0) tma_store_wait_all();
}

template <bool KEY16, int IO>
int launch_rc4_tma(const Rc4TmaArgs &a, const CUtensorMap &tin, const CUtensorMap &tout, cudaStream_t st)
{
    int rc = set_smem<rc4_batch_tma_kernel<KEY16, IO>>(kTmaSmem);
    if (rc) return rc;
    const int sms = bc_cached_sm_count();
    if (sms <= 0) return BC_ERR_CUDA;
    const uint64_t groups = (a.n_streams + 31) / 32;
    const uint64_t ctas = (groups + kTmaWarps - 1) / kTmaWarps;
    const uint64_t cap = (uint64_t)sms * 2;
    const int grid = (int)(ctas < cap ? ctas : cap);
    rc4_batch_tma_kernel<KEY16, IO><<<grid, kRc4Threads, kTmaSmem, st>>>(a, tin, tout);
    BC_CUDA(cudaGetLastError(), "rc4_batch_tma_kernel launch");
    return BC_OK;
}

__global__ void xor_kernel(const uint4 *a, const uint4 *b, uint4 *o, uint64_t n16,
                           const unsigned char *ta, const unsigned char *tb, unsigned char *to, uint32_t tail)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
        const uint4 p = a[i], q = b[i];
        o[i] = make_uint4(p.x ^ q.x, p.y ^ q.y, p.z ^ q.z, p.w ^ q.w);
    }
    if (blockIdx.x == 0 && threadIdx.x < tail) to[threadIdx.x] = ta[threadIdx.x] ^ tb[threadIdx.x];
}

template <bool VEC, bool KEY16>
int launch_rc4(const Rc4Args &a, cudaStream_t st)
{
    int rc = set_smem<rc4_batch_kernel<VEC, KEY16>>(kRc4Smem);
    if (rc) return rc;
    const int sms = bc_cached_sm_count();
    if (sms <= 0) return BC_ERR_CUDA;
    uint64_t ctas = (a.n_streams + kRc4Threads - 1) / kRc4Threads;
    const uint64_t cap = (uint64_t)sms * 3;
    const int grid = (int)(ctas < cap ? ctas : cap);
    rc4_batch_kernel<VEC, KEY16><<<grid, kRc4Threads, kRc4Smem, st>>>(a);
    BC_CUDA(cudaGetLastError(), "rc4_batch_kernel launch");
    return BC_OK;
}

}  // namespace

extern "C" int bc_arc4_batch_dev(size_t n_streams, const void *d_keys, unsigned keylen, size_t length,
                                 const void *d_in, void *d_out, const void *d_ctx_in, void *d_ctx_out,
                                 void *stream)
{
    if (n_streams == 0) return BC_OK;
    if (!d_ctx_in && (!d_keys || keylen < 1 || keylen > 256)) {
        bc_set_error("RC4 needs keys with 1 <= keylen <= 256 (got %u) or contexts to resume", keylen);
        return BC_ERR_BAD_ARG;
    }
    if (length && !d_out) { bc_set_error("NULL output pointer"); return BC_ERR_BAD_ARG; }
    if ((((uintptr_t)d_ctx_in) | ((uintptr_t)d_ctx_out)) & 7u) {
        bc_set_error("arc4 contexts must be 8-byte aligned");
        return BC_ERR_BAD_ARG;
    }
    if (d_in && d_in != d_out) {
        const uintptr_t i = (uintptr_t)d_in, o = (uintptr_t)d_out;
        const size_t nb = n_streams * length;
        if (i < o + nb && o < i + nb) { bc_set_error("input and output overlap partially"); return BC_ERR_OVERLAP; }
    }
    Rc4Args a;
    a.n_streams = n_streams;
    a.keys = (const unsigned char *)d_keys;
    a.keylen = keylen;
    a.length = length;
    a.in = (const unsigned char *)d_in;
    a.out = (unsigned char *)d_out;
    a.ctx_in = (const unsigned char *)d_ctx_in;
    a.ctx_out = (unsigned char *)d_ctx_out;
    const bool vec = (length % 16 == 0) && !((((uintptr_t)d_in) | ((uintptr_t)d_out)) & 15u);
    const bool key16 = !d_ctx_in && keylen == 16 && !(((uintptr_t)d_keys) & 15u);
    cudaStream_t st = (cudaStream_t)stream;
    if (vec && length >= 16 && length < ((size_t)1 << 32) && n_streams < ((size_t)1 << 31)) {
        /* fast path: stream data staged through shared memory by TMA */
        Rc4TmaArgs t;
        t.n_streams = n_streams;
        t.keys = (const unsigned char *)d_keys;
        t.keylen = keylen;
        t.length = length;
        t.has_in = d_in != nullptr;
        t.xor_in_place = (d_in != nullptr && d_in == d_out && !getenv("BULKCIPHER_NO_TMA_REDUCE"));
        t.ctx_in = (const unsigned char *)d_ctx_in;
        t.ctx_out = (unsigned char *)d_ctx_out;
        CUtensorMap tin, tout;
        int rc = make_row_tensor_map(&tout, d_out, n_streams, length, t.xor_in_place != 0);
        if (rc) return rc;
        if ((rc = make_row_tensor_map(&tin, d_in ? d_in : d_out, n_streams, length))) return rc;
        const int io = !t.has_in ? IO_KS : t.xor_in_place ? IO_XOR : IO_LOAD;
        if (key16) return io == IO_LOAD ? launch_rc4_tma<true, IO_LOAD>(t, tin, tout, st)
                        : io == IO_XOR ? launch_rc4_tma<true, IO_XOR>(t, tin, tout, st) : launch_rc4_tma<true, IO_KS>(t, tin, tout, st);
        return io == IO_LOAD ? launch_rc4_tma<false, IO_LOAD>(t, tin, tout, st)
             : io == IO_XOR ? launch_rc4_tma<false, IO_XOR>(t, tin, tout, st) : launch_rc4_tma<false, IO_KS>(t, tin, tout, st);
    }
    if (vec) return key16 ? launch_rc4<true, true>(a, st) : launch_rc4<true, false>(a, st);
    return key16 ? launch_rc4<false, true>(a, st) : launch_rc4<false, false>(a, st);
}

/* out = a ^ b on device buffers: what arc4_crypt does with a prepared keystream
 * (arc4.c:101-112).  Any of the three may alias elementwise. */
extern "C" int bc_xor_dev(const void *d_a, const void *d_b, void *d_out, size_t nbytes, void *stream)
{
    if (nbytes == 0) return BC_OK;
    if (!d_a || !d_b || !d_out) { bc_set_error("NULL pointer"); return BC_ERR_BAD_ARG; }
    if ((((uintptr_t)d_a) | ((uintptr_t)d_b) | ((uintptr_t)d_out)) & 15u) {
        bc_set_error("device buffers must be 16-byte aligned");
        return BC_ERR_BAD_ARG;
    }
    const int sms = bc_cached_sm_count();
    if (sms <= 0) return BC_ERR_CUDA;
    const uint64_t n16 = nbytes / 16;
    const uint32_t tail = (uint32_t)(nbytes % 16);
    uint64_t want = (n16 + 255) / 256;
    if (want < 1) want = 1;
    const uint64_t cap = (uint64_t)sms * 8;
    const unsigned grid = (unsigned)(want < cap ? want : cap);
    const unsigned char *ta = (const unsigned char *)d_a + n16 * 16, *tb = (const unsigned char *)d_b + n16 * 16;
    xor_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const uint4 *)d_a, (const uint4 *)d_b, (uint4 *)d_out, n16,
                                                        ta, tb, (unsigned char *)d_out + n16 * 16, tail);
    BC_CUDA(cudaGetLastError(), "xor_kernel launch");
    return BC_OK;
}
