/*
 * bc_aes.cu — AES-128/192/256 ECB / CBC / CTR kernels for sm_100a and their
 * extern "C" launchers (the device-pointer half of include/bulkcipher.h).
 *
 * Written from scratch for B200; not a port of the reference's
 * aes-gpu/Source/AES.cu.  Results are bit-exact with the reference's aes.c /
 * aesni.c (checked against oracle/ in tests/).
 *
 * Design (DESIGN.md §3 has the numbers):
 *  - Integer / shared-memory work, no tensor cores.  One AES round of one
 *    block is 16 table lookups; the shared-memory pipe serves 32 lanes per
 *    clock per SM, which is the roofline that binds (not HBM).
 *  - T-tables FT0..FT3 (aes.c:409-422) live in shared memory replicated once
 *    per bank: word (i, table, lane) sits at byte  i*256 + (table&1)*128 +
 *    lane*4  of array table>>1, so lane L only ever touches bank L and every
 *    warp-wide lookup is one conflict-free wavefront.  With this layout a
 *    lookup address is ONE PRMT: the state byte lands in address byte 1, the
 *    lane offset in byte 0; the table select rides in the LDS immediate.
 *  - Same little-endian column words as aes.c (GET_ULONG_LE, aes.c:42-49): a
 *    128-bit global load is the state, no byte swaps.
 *  - Round keys arrive as a __grid_constant__ kernel argument (constant bank),
 *    so XORs take them as direct operands and launches are re-entrant.
 *  - CTR: counters are generated in registers from the block index.  Work is
 *    cut into groups of 256 consecutive counter values (only the last counter
 *    byte varies inside a group), so 15 of round 1's 16 lookups and 12 of
 *    round 2's 16 are per-group constants (AES_FROUND index pattern,
 *    aes.c:603-621), and round 1's 16th is a per-lane constant for the whole
 *    kernel: 132 instead of 160 lookups per AES-128 block.  Lane i of a warp
 *    precomputes the constants of the warp's i-th group; they are broadcast
 *    with shuffles.
 *  - Persistent grid: one 1024-thread CTA per SM (tables fill 128 KB; 32 warps
 *    at 64 registers hide the LDS latency), warps stride over 4 KiB
 *    row-groups; global traffic is 128-bit, coalesced (a warp row = 512
 *    contiguous bytes), streaming (evict-first) hints.
 *  - Batched CBC/CFB encrypt (one chain per lane) stages its rows through
 *    shared memory with 2-D TMA boxes (bc_tma.cuh).
 *
 * The BC_CTR_* / BC_BLK_THREADS macros are the knobs of the A/B runs recorded in
 * profiles/r1_summary.md; the defaults are the measured best.
 */
#include <cuda_runtime.h>
#include <stdint.h>

#include "bulkcipher.h"
#include "bc_internal.h"
#include "bc_tma.cuh"
#include "bc_sbox.inc"

namespace {

__constant__ uint8_t c_sbox[256]     = { BC_SBOX_INIT };
__constant__ uint8_t c_inv_sbox[256] = { BC_INV_SBOX_INIT };

#ifndef BC_BLK_THREADS
#define BC_BLK_THREADS 1024
#endif
constexpr int kThreads      = BC_BLK_THREADS;    /* ECB / CBC kernels: CTA size */
constexpr int kWarps        = kThreads / 32;
#ifndef BC_CTR_ILP
#define BC_CTR_ILP 2          /* rows in flight per thread in the CTR fast path */
#endif
#ifndef BC_CTR_PREFETCH
#define BC_CTR_PREFETCH 1     /* fetch the next trip's plaintext before computing this trip's keystream */
#endif
#ifndef BC_CTR_F3
#define BC_CTR_F3 1           /* keep round 1's eight per-lane lookups in registers for the whole kernel */
#endif
#ifndef BC_CTR_THREADS
#define BC_CTR_THREADS 1024
#endif
constexpr int kCtrThreads   = BC_CTR_THREADS;      /* CTR kernel CTA size (512 or 1024) */
constexpr int kArrayBytes   = 65536;           /* two interleaved tables */
constexpr int kEncSmemBytes = 2 * kArrayBytes; /* FT0|FT1, FT2|FT3 */
constexpr int kDecSmemBytes = 3 * kArrayBytes; /* RT0|RT1, RT2|RT3, RSb×4|unused */
constexpr int kRowsPerGroup = 8;               /* 8 rows × 32 lanes = 256 blocks = 4 KiB */
constexpr int kMaxGroupsPerChunk = 8;          /* CTR: groups handled per warp trip (1, 2, 4 or 8) */

struct AesSched { uint32_t rk[60]; };

/* byte offsets of the tables inside dynamic shared memory */
#ifndef BC_NO_REPLICATION
#define BC_NO_REPLICATION 0   /* A/B knob: plain 1 KB tables, 32 random indices collide in the banks */
#endif
#if BC_NO_REPLICATION
constexpr int OFF_T0 = 0, OFF_T1 = 1024, OFF_T2 = 2048, OFF_T3 = 3072;
constexpr int OFF_SB = 4096;
#else
constexpr int OFF_T0 = 0, OFF_T1 = 128, OFF_T2 = kArrayBytes, OFF_T3 = kArrayBytes + 128;
constexpr int OFF_SB = 2 * kArrayBytes;
#endif

__device__ __forceinline__ uint32_t rotl32(uint32_t v, int n) { return __funnelshift_l(v, v, n); }
__device__ __forceinline__ uint32_t xtime8(uint32_t x) { return ((x << 1) ^ ((x & 0x80u) ? 0x1bu : 0u)) & 0xffu; }

/* FT0[i] = {02,01,01,03}·S(i), low byte first (aes.c:415-418) */
__device__ __forceinline__ uint32_t ft0_word(uint32_t i)
{
    uint32_t s = c_sbox[i], d = xtime8(s);
    return d | (s << 8) | (s << 16) | ((d ^ s) << 24);
}

/* RT0[i] = {0E,09,0D,0B}·S^-1(i), low byte first (aes.c:426-429) */
__device__ __forceinline__ uint32_t rt0_word(uint32_t i)
{
    uint32_t q = c_inv_sbox[i], q2 = xtime8(q), q4 = xtime8(q2), q8 = xtime8(q4);
    return (q8 ^ q4 ^ q2) | ((q8 ^ q) << 8) | ((q8 ^ q4 ^ q) << 16) | ((q8 ^ q2 ^ q) << 24);
}

template <bool DEC>
__device__ __forceinline__ void fill_tables(uint32_t *sm)
{
#if BC_NO_REPLICATION
    for (uint32_t i = threadIdx.x; i < 256u; i += blockDim.x) {
        uint32_t w = DEC ? rt0_word(i) : ft0_word(i);
        sm[i] = w; sm[256 + i] = rotl32(w, 8); sm[512 + i] = rotl32(w, 16); sm[768 + i] = rotl32(w, 24);
        if (DEC) sm[1024 + i] = (uint32_t)c_inv_sbox[i] * 0x01010101u;
    }
    return;
#endif
    /* one thread writes four neighbouring banks' copies of an entry with one 128-bit store: a quarter
     * of the store instructions and of the table arithmetic of a word-per-thread fill (the shared-memory
     * wavefronts are the same 1 024 per 128 KB; tools/microbench/fill_probe.cu has the timings) */
    uint4 *sm4 = reinterpret_cast<uint4 *>(sm);
    for (uint32_t idx = threadIdx.x; idx < 256u * 8u; idx += blockDim.x) {
        const uint32_t i = idx >> 3, q = idx & 7u;          /* entry i, lanes 4q .. 4q+3 */
        const uint32_t w = DEC ? rt0_word(i) : ft0_word(i);
        const uint32_t w1 = rotl32(w, 8), w2 = rotl32(w, 16), w3 = rotl32(w, 24);
        sm4[i * 16 + q]            = make_uint4(w, w, w, w);
        sm4[i * 16 + 8 + q]        = make_uint4(w1, w1, w1, w1);
        sm4[4096 + i * 16 + q]     = make_uint4(w2, w2, w2, w2);
        sm4[4096 + i * 16 + 8 + q] = make_uint4(w3, w3, w3, w3);
        if (DEC) { const uint32_t r = (uint32_t)c_inv_sbox[i] * 0x01010101u; sm4[8192 + i * 16 + q] = make_uint4(r, r, r, r); }
    }
}

/* lookup address: byte K of x -> address byte 1, lane*4 -> address byte 0 */
template <int K>
__device__ __forceinline__ uint32_t baddr(uint32_t x, uint32_t lane4)
{
#if BC_NO_REPLICATION
    (void)lane4;
    return ((x >> (8 * K)) & 0xffu) << 2;
#else
    return __byte_perm(x, lane4, 0x5504 | (K << 4));
#endif
}

template <int OFF>
__device__ __forceinline__ uint32_t lds32(uint32_t addr)
{
    uint32_t v;
    asm("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(OFF));
    return v;
}

/* Shared-memory context of a thread: `base` is the shared-window address of
 * the tables (a ptxas-time constant, it folds into the LDS immediate together
 * with the table offset); `lane4` = lane*4 selects the thread's bank. */
struct Tb { uint32_t base, lane4; };

/* TB(OFF, K, x): table at byte offset OFF, indexed by byte K of x — one PRMT + one LDS */
#define TB(OFF, K, x) lds32<OFF>(tb.base + baddr<K>((x), tb.lane4))

/* TBL(k, K, x): forward/reverse table k (FTk / RTk) indexed by byte K of x.  Default: four
 * tables in shared memory.  BC_SINGLE_TABLE=1 is the "__byte_perm formulation" of the north
 * star: only table 0 is looked up and tables 1-3 are its byte rotations (aes.c:420-422), one
 * extra PRMT per lookup — kept as an A/B knob (profiles/r1_summary.md), it loses because the
 * ALU pipe then saturates before the shared-memory pipe. */
#ifndef BC_SINGLE_TABLE
#define BC_SINGLE_TABLE 0
#endif
template <int k> __device__ __forceinline__ uint32_t rot_table(uint32_t v)
{
    return k == 0 ? v : __byte_perm(v, 0, k == 1 ? 0x2103 : k == 2 ? 0x1032 : 0x0321);
}
#if BC_SINGLE_TABLE
#define TBL(k, K, x) rot_table<k>(TB(OFF_T0, K, x))
#else
#define TBL(k, K, x) TB((k) == 0 ? OFF_T0 : (k) == 1 ? OFF_T1 : (k) == 2 ? OFF_T2 : OFF_T3, K, x)
#endif

struct State { uint32_t a, b, c, d; };

/* one forward table round, AES_FROUND (aes.c:601-622) */
__device__ __forceinline__ State enc_round(const State s, const uint32_t *rk, const Tb tb)
{
    State t;
    t.a = TBL(0, 0, s.a) ^ TBL(1, 1, s.b) ^ TBL(2, 2, s.c) ^ TBL(3, 3, s.d) ^ rk[0];
    t.b = TBL(0, 0, s.b) ^ TBL(1, 1, s.c) ^ TBL(2, 2, s.d) ^ TBL(3, 3, s.a) ^ rk[1];
    t.c = TBL(0, 0, s.c) ^ TBL(1, 1, s.d) ^ TBL(2, 2, s.a) ^ TBL(3, 3, s.b) ^ rk[2];
    t.d = TBL(0, 0, s.d) ^ TBL(1, 1, s.a) ^ TBL(2, 2, s.b) ^ TBL(3, 3, s.c) ^ rk[3];
    return t;
}

/* one reverse table round, AES_RROUND (aes.c:624-645) */
__device__ __forceinline__ State dec_round(const State s, const uint32_t *rk, const Tb tb)
{
    State t;
    t.a = TBL(0, 0, s.a) ^ TBL(1, 1, s.d) ^ TBL(2, 2, s.c) ^ TBL(3, 3, s.b) ^ rk[0];
    t.b = TBL(0, 0, s.b) ^ TBL(1, 1, s.a) ^ TBL(2, 2, s.d) ^ TBL(3, 3, s.c) ^ rk[1];
    t.c = TBL(0, 0, s.c) ^ TBL(1, 1, s.b) ^ TBL(2, 2, s.a) ^ TBL(3, 3, s.d) ^ rk[2];
    t.d = TBL(0, 0, s.d) ^ TBL(1, 1, s.c) ^ TBL(2, 2, s.b) ^ TBL(3, 3, s.a) ^ rk[3];
    return t;
}

/* pick byte0 of w0, byte1 of w1, byte2 of w2, byte3 of w3 */
__device__ __forceinline__ uint32_t merge_bytes(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3)
{
    return __byte_perm(__byte_perm(w0, w1, 0x0050), __byte_perm(w2, w3, 0x7200), 0x7610);
}

/* Last forward round (aes.c:721-743).  Plain S(x) is read out of the T-tables:
 * byte 0 of FT2, byte 1 of FT3, byte 2 of FT0, byte 3 of FT1 are all S(x). */
__device__ __forceinline__ State enc_last(const State s, const uint32_t *rk, const Tb tb)
{
    State t;
    t.a = merge_bytes(TBL(2, 0, s.a), TBL(3, 1, s.b), TBL(0, 2, s.c), TBL(1, 3, s.d)) ^ rk[0];
    t.b = merge_bytes(TBL(2, 0, s.b), TBL(3, 1, s.c), TBL(0, 2, s.d), TBL(1, 3, s.a)) ^ rk[1];
    t.c = merge_bytes(TBL(2, 0, s.c), TBL(3, 1, s.d), TBL(0, 2, s.a), TBL(1, 3, s.b)) ^ rk[2];
    t.d = merge_bytes(TBL(2, 0, s.d), TBL(3, 1, s.a), TBL(0, 2, s.b), TBL(1, 3, s.c)) ^ rk[3];
    return t;
}

/* Last reverse round (aes.c:687-709) from the byte-replicated RSb table. */
__device__ __forceinline__ State dec_last(const State s, const uint32_t *rk, const Tb tb)
{
    State t;
    t.a = merge_bytes(TB(OFF_SB, 0, s.a), TB(OFF_SB, 1, s.d), TB(OFF_SB, 2, s.c), TB(OFF_SB, 3, s.b)) ^ rk[0];
    t.b = merge_bytes(TB(OFF_SB, 0, s.b), TB(OFF_SB, 1, s.a), TB(OFF_SB, 2, s.d), TB(OFF_SB, 3, s.c)) ^ rk[1];
    t.c = merge_bytes(TB(OFF_SB, 0, s.c), TB(OFF_SB, 1, s.b), TB(OFF_SB, 2, s.a), TB(OFF_SB, 3, s.d)) ^ rk[2];
    t.d = merge_bytes(TB(OFF_SB, 0, s.d), TB(OFF_SB, 1, s.c), TB(OFF_SB, 2, s.b), TB(OFF_SB, 3, s.a)) ^ rk[3];
    return t;
}

template <int NR, bool DEC>
__device__ __forceinline__ State cipher_block(State s, const AesSched &ks, const Tb tb)
{
    s.a ^= ks.rk[0]; s.b ^= ks.rk[1]; s.c ^= ks.rk[2]; s.d ^= ks.rk[3];
#pragma unroll
    for (int r = 1; r < NR; r++) s = DEC ? dec_round(s, ks.rk + 4 * r, tb) : enc_round(s, ks.rk + 4 * r, tb);
    return DEC ? dec_last(s, ks.rk + 4 * NR, tb) : enc_last(s, ks.rk + 4 * NR, tb);
}

__device__ __forceinline__ uint32_t smem_base_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ uint4 ld_stream(const uint4 *p) { return __ldcs(p); }
__device__ __forceinline__ void  st_stream(uint4 *p, uint4 v) { __stcs(p, v); }

/* ------------------------------------------------------------------- CTR */

struct CtrArgs {
    uint64_t base_hi, base_lo;   /* counter of virtual block 0 (low byte of base_lo is 0) */
    uint32_t c_lo;               /* low byte of the first real counter: virtual index of block 0 */
    uint32_t carry128;           /* 1: carry from bytes 8..15 into 0..7 (aes.c:884-888); 0: aesni.c:139 */
    const uint4 *in;
    uint4 *out;
    uint64_t nblocks;            /* whole 16-byte blocks */
    uint32_t tail;               /* bytes of a final partial block, 0..15 */
    uint64_t n_chunks;
    uint32_t gpc;                /* groups per chunk: 8 for big buffers, fewer so small ones still fill the SMs */
    uint4 *last_ks;              /* optional: keystream of the last block touched */
};

/* Keystream block for one counter value: rounds 1-2 from the group constants
 * (c0, d0..d3) and the lane's precomputed round-1 lookup f, then NR-3 full table
 * rounds and the last round.  4 + 16*(NR-3) + 16 lookups (132 for AES-128). */
template <int NR>
__device__ __forceinline__ State ctr_keystream(uint32_t c0, uint32_t d0, uint32_t d1, uint32_t d2, uint32_t d3,
                                               uint32_t f, const AesSched &ks, const Tb tb)
{
    const uint32_t x0 = c0 ^ f;
    State s;
    s.a = d0 ^ TBL(0, 0, x0);
    s.b = d1 ^ TBL(3, 3, x0);
    s.c = d2 ^ TBL(2, 2, x0);
    s.d = d3 ^ TBL(1, 1, x0);
#pragma unroll
    for (int r = 3; r < NR; r++) s = enc_round(s, ks.rk + 4 * r, tb);
    return enc_last(s, ks.rk + 4 * NR, tb);
}

__device__ __forceinline__ uint4 xor4(uint4 p, const State s)
{
    return make_uint4(p.x ^ s.a, p.y ^ s.b, p.z ^ s.c, p.w ^ s.d);
}

template <int NR, int THREADS>
__global__ void __launch_bounds__(THREADS, 1)
aes_ctr_kernel(const __grid_constant__ AesSched ks, const __grid_constant__ CtrArgs a)
{
    constexpr int WARPS = THREADS / 32;
    extern __shared__ __align__(1024) uint32_t smem[];
    fill_tables<false>(smem);
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31u;
    Tb tb = { smem_base_u32(smem), lane * 4u };
    asm volatile("" : "+r"(tb.lane4));    /* pin table reads below the barrier */

    /* Round 1's only varying lookup is FT3[j ^ k3] with j the last counter byte.  Row `it`
     * of a group gives this lane j = it*32 + lane for EVERY group, so the eight values are
     * fetched once per kernel and round 1 costs no lookup at all afterwards. */
    const uint32_t k3 = ks.rk[3] >> 24;
#if BC_CTR_F3
    uint32_t f3[kRowsPerGroup];
#pragma unroll
    for (int it = 0; it < kRowsPerGroup; it++)
        f3[it] = TBL(3, 0, ((uint32_t)it * 32u + lane) ^ k3);
    /* the fast path walks rows in order, so the values rotate through f3[0..] instead of being indexed */
#define F3(k) f3[k]
#define F3_ROTATE()                                                                      \
    do {                                                                                 \
        uint32_t t_[BC_CTR_ILP];                                                         \
        _Pragma("unroll") for (int k_ = 0; k_ < BC_CTR_ILP; k_++) t_[k_] = f3[k_];       \
        _Pragma("unroll") for (int k_ = 0; k_ < kRowsPerGroup - BC_CTR_ILP; k_++) f3[k_] = f3[k_ + BC_CTR_ILP]; \
        _Pragma("unroll") for (int k_ = 0; k_ < BC_CTR_ILP; k_++) f3[kRowsPerGroup - BC_CTR_ILP + k_] = t_[k_]; \
    } while (0)
#define F3_AT(it) f3[it]
#else
    const uint32_t a3_lane = (((lane ^ k3) & 0xffu) << 8) | tb.lane4;
    uint32_t f3_it = 0;                  /* row index << 13: flips the bits of j above bit 4 */
#define F3(k) lds32<OFF_T3>(tb.base + (a3_lane ^ (f3_it + ((uint32_t)(k) << 13))))
#define F3_ROTATE() do { f3_it = (f3_it + ((uint32_t)BC_CTR_ILP << 13)) & (7u << 13); } while (0)
#define F3_AT(it) lds32<OFF_T3>(tb.base + (a3_lane ^ ((uint32_t)(it) << 13)))
#endif

    const uint64_t nb_all = a.nblocks + (a.tail ? 1u : 0u);
    /* work items go round-robin over CTAs first, then over a CTA's warps: every SM gets the
     * same number of items to within one (the warps of an SM share one data pipe anyway) */
    const uint64_t warp_gid = blockIdx.x + (uint64_t)gridDim.x * (threadIdx.x >> 5);
    const uint64_t n_warps = (uint64_t)gridDim.x * WARPS;

    for (uint64_t chunk = warp_gid; chunk < a.n_chunks; chunk += n_warps) {
        /* ---- per-group constants: lane i works on group chunk*gpc + (i & (gpc-1)) ---- */
        uint32_t C0, D0, D1, D2, D3;
        {
            const uint64_t g = chunk * a.gpc + (lane & (a.gpc - 1));
            const uint64_t lo = a.base_lo + (g << 8);
            const uint64_t hi = a.base_hi + ((a.carry128 && lo < a.base_lo) ? 1u : 0u);
            const uint32_t s0 = __byte_perm((uint32_t)(hi >> 32), 0, 0x0123) ^ ks.rk[0];
            const uint32_t s1 = __byte_perm((uint32_t)hi, 0, 0x0123) ^ ks.rk[1];
            const uint32_t s2 = __byte_perm((uint32_t)(lo >> 32), 0, 0x0123) ^ ks.rk[2];
            const uint32_t s3 = __byte_perm((uint32_t)lo, 0, 0x0123) ^ ks.rk[3];   /* byte 3 varies per block */
            C0 = TBL(0, 0, s0) ^ TBL(1, 1, s1) ^ TBL(2, 2, s2) ^ ks.rk[4];
            const uint32_t X1 = TBL(0, 0, s1) ^ TBL(1, 1, s2) ^ TBL(2, 2, s3) ^ TBL(3, 3, s0) ^ ks.rk[5];
            const uint32_t X2 = TBL(0, 0, s2) ^ TBL(1, 1, s3) ^ TBL(2, 2, s0) ^ TBL(3, 3, s1) ^ ks.rk[6];
            const uint32_t X3 = TBL(0, 0, s3) ^ TBL(1, 1, s0) ^ TBL(2, 2, s1) ^ TBL(3, 3, s2) ^ ks.rk[7];
            D0 = TBL(1, 1, X1) ^ TBL(2, 2, X2) ^ TBL(3, 3, X3) ^ ks.rk[8];
            D1 = TBL(0, 0, X1) ^ TBL(1, 1, X2) ^ TBL(2, 2, X3) ^ ks.rk[9];
            D2 = TBL(0, 0, X2) ^ TBL(1, 1, X3) ^ TBL(3, 3, X1) ^ ks.rk[10];
            D3 = TBL(0, 0, X3) ^ TBL(2, 2, X1) ^ TBL(3, 3, X2) ^ ks.rk[11];
        }

        const uint64_t v0 = chunk * (uint64_t)(a.gpc * 256u);   /* first virtual index */
        const bool full = (v0 >= a.c_lo) && (v0 + a.gpc * 256u <= a.c_lo + a.nblocks);

        if (full) {
            /* fast path: two rows in flight per thread, plaintext fetched two rows ahead */
            const uint4 *pin = a.in + (v0 - a.c_lo) + lane;
            uint4 *pout = a.out + (v0 - a.c_lo) + lane;
#if BC_CTR_PREFETCH
            uint4 q0 = ld_stream(pin), q1 = ld_stream(pin + (BC_CTR_ILP == 2 ? 32 : 0));
#else
            uint4 q0, q1;
#endif
            (void)q1;
#pragma unroll 1
            for (int gi = 0; gi < (int)a.gpc; gi++) {
                const uint32_t c0 = __shfl_sync(0xffffffffu, C0, gi);
                const uint32_t d0 = __shfl_sync(0xffffffffu, D0, gi);
                const uint32_t d1 = __shfl_sync(0xffffffffu, D1, gi);
                const uint32_t d2 = __shfl_sync(0xffffffffu, D2, gi);
                const uint32_t d3 = __shfl_sync(0xffffffffu, D3, gi);
                const bool last_group = (gi + 1 == (int)a.gpc);
#pragma unroll 1
                for (int it = 0; it < kRowsPerGroup; it += BC_CTR_ILP) {
#if BC_CTR_ILP == 2
#if BC_CTR_PREFETCH
                    uint4 n0 = q0, n1 = q1;
                    if (!(last_group && it + 2 >= kRowsPerGroup)) {       /* rows it+2, it+3 (may be the next group's) */
                        n0 = ld_stream(pin + 64);
                        n1 = ld_stream(pin + 96);
                    }
#else
                    q0 = ld_stream(pin); q1 = ld_stream(pin + 32);
#endif
                    const State sA = ctr_keystream<NR>(c0, d0, d1, d2, d3, F3(0), ks, tb);
                    const State sB = ctr_keystream<NR>(c0, d0, d1, d2, d3, F3(1), ks, tb);
                    st_stream(pout, xor4(q0, sA));
                    st_stream(pout + 32, xor4(q1, sB));
#if BC_CTR_PREFETCH
                    q0 = n0; q1 = n1;
#endif
                    pin += 64; pout += 64;
#else
#if BC_CTR_PREFETCH
                    uint4 n0 = q0;
                    if (!(last_group && it + 1 >= kRowsPerGroup)) n0 = ld_stream(pin + 32);
#else
                    q0 = ld_stream(pin);
#endif
                    const State sA = ctr_keystream<NR>(c0, d0, d1, d2, d3, F3(0), ks, tb);
                    st_stream(pout, xor4(q0, sA));
#if BC_CTR_PREFETCH
                    q0 = n0;
#endif
                    pin += 32; pout += 32;
#endif
                    F3_ROTATE();
                }
            }
            continue;
        }

        /* edge chunks: per-lane bounds, ragged tail, last-keystream output */
#pragma unroll 1
        for (int gi = 0; gi < (int)a.gpc; gi++) {
            const uint32_t c0 = __shfl_sync(0xffffffffu, C0, gi);
            const uint32_t d0 = __shfl_sync(0xffffffffu, D0, gi);
            const uint32_t d1 = __shfl_sync(0xffffffffu, D1, gi);
            const uint32_t d2 = __shfl_sync(0xffffffffu, D2, gi);
            const uint32_t d3 = __shfl_sync(0xffffffffu, D3, gi);
            const uint64_t vg = v0 + (uint64_t)gi * 256 + lane;          /* this lane's row-0 virtual index */
#pragma unroll
            for (int it = 0; it < kRowsPerGroup; it++) {
                const uint64_t v = vg + (uint64_t)it * 32;
                const uint64_t b = v - a.c_lo;
                if (!(v >= a.c_lo && b < nb_all)) continue;
                const State s = ctr_keystream<NR>(c0, d0, d1, d2, d3, F3_AT(it), ks, tb);
                if (b < a.nblocks) {
                    st_stream(a.out + b, xor4(ld_stream(a.in + b), s));
                } else {                                       /* ragged tail, byte-wise (aes.c:879-895) */
                    const unsigned char *pi = (const unsigned char *)(a.in + b);
                    unsigned char *po = (unsigned char *)(a.out + b);
                    for (uint32_t i = 0; i < a.tail; i++) {
                        const uint32_t q = i >> 2;
                        const uint32_t w = q == 0 ? s.a : q == 1 ? s.b : q == 2 ? s.c : s.d;
                        po[i] = pi[i] ^ (unsigned char)(w >> (8 * (i & 3)));
                    }
                }
                if (a.last_ks && b == nb_all - 1) *a.last_ks = make_uint4(s.a, s.b, s.c, s.d);
            }
        }
    }
}

/* --------------------------------------------------- ECB enc/dec, CBC dec */

enum { MODE_ECB_ENC = 0, MODE_ECB_DEC = 1, MODE_CBC_DEC = 2, MODE_CFB_DEC = 3 };

struct BlkArgs {
    const uint4 *in;
    uint4 *out;
    uint64_t nblocks;
    uint64_t n_groups;       /* ceil(nblocks / 256) */
    uint4 iv;                /* CBC: block before block 0 */
    const uint4 *boundary;   /* CBC in-place: boundary[g] = ciphertext block 256*g - 1 (g >= 1) */
    uint32_t be32;           /* 1: blocks are four BIG-endian 32-bit words in memory (aes-gpu's uint[] interface) */
};

template <int NR, int MODE>
__global__ void __launch_bounds__(kThreads, 1)
aes_blocks_kernel(const __grid_constant__ AesSched ks, const __grid_constant__ BlkArgs a)
{
    constexpr bool DEC = (MODE == MODE_ECB_DEC || MODE == MODE_CBC_DEC);      /* inverse tables */
    constexpr bool CHAIN = (MODE == MODE_CBC_DEC || MODE == MODE_CFB_DEC);    /* needs C[i-1] */
    extern __shared__ __align__(1024) uint32_t smem[];
    fill_tables<DEC>(smem);
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31u;
    Tb tb = { smem_base_u32(smem), lane * 4u };
    asm volatile("" : "+r"(tb.lane4));

    const uint64_t warp_gid = blockIdx.x + (uint64_t)gridDim.x * (threadIdx.x >> 5);   /* round-robin over CTAs first */
    const uint64_t n_warps = (uint64_t)gridDim.x * kWarps;

    for (uint64_t g = warp_gid; g < a.n_groups; g += n_warps) {
        const uint64_t b0 = g * 256 + lane;
        const bool full = (g * 256 + 256 <= a.nblocks);
        uint4 carry = make_uint4(0, 0, 0, 0);      /* chained modes: lane 31's ciphertext of the previous row */
        if (CHAIN) {
            if (g == 0) carry = a.iv;
            else carry = a.boundary ? a.boundary[g] : a.in[g * 256 - 1];
        }
#pragma unroll 1
        for (int it = 0; it < kRowsPerGroup; it++) {
            const uint64_t b = b0 + (uint64_t)it * 32;
            const bool live = full || b < a.nblocks;
            uint4 c = make_uint4(0, 0, 0, 0);
            if (live) c = ld_stream(a.in + b);
            if (!CHAIN && a.be32) {                /* GETWORD/PUTWORD of aes-gpu/Source/AES.cu:42-43: swap to the byte stream */
                c.x = __byte_perm(c.x, 0, 0x0123); c.y = __byte_perm(c.y, 0, 0x0123);
                c.z = __byte_perm(c.z, 0, 0x0123); c.w = __byte_perm(c.w, 0, 0x0123);
            }
            uint4 prev = make_uint4(0, 0, 0, 0);
            if (CHAIN && a.in != a.out) {
                /* out of place nothing overwrites the input: C_{i-1} is one more (cache-resident) load,
                 * cheaper on the shared data pipe than eight shuffles per row */
                if (live) prev = (b == 0) ? a.iv : a.in[b - 1];
            } else if (CHAIN) {
                /* in place: a row's loads all come before its stores, so the lane to the left can
                 * still be read from memory; only lane 0 needs the previous row's last block, which
                 * that row's store has already overwritten — lane 0 reads it a row early into `carry` */
                if (live && lane != 0) prev = a.in[b - 1];
                if (lane == 0) {
                    prev = carry;
                    /* the next row's lane 0 needs this row's last ciphertext block, which lane 31 is
                     * about to overwrite: fetch it now (one single-lane load = one wavefront, where
                     * four warp-wide shuffles cost four on the same data pipe) */
                    if (it + 1 < kRowsPerGroup && (full || b + 31 < a.nblocks)) carry = a.in[b + 31];
                }
                /* lane L just read in[b-1] (lane 0: in[b+31]), blocks other lanes store below: order every
                 * lane's loads before any lane's store */
                __syncwarp();
            }
            State s;
            if (MODE == MODE_CFB_DEC) s = State{ prev.x, prev.y, prev.z, prev.w };   /* P_i = E(C_{i-1}) ^ C_i (aes.c:832-842) */
            else s = State{ c.x, c.y, c.z, c.w };
            s = cipher_block<NR, DEC>(s, ks, tb);
            if (MODE == MODE_CBC_DEC) {            /* P_i = D(C_i) ^ C_{i-1}  (aes.c:784-797) */
                s.a ^= prev.x; s.b ^= prev.y; s.c ^= prev.z; s.d ^= prev.w;
            } else if (MODE == MODE_CFB_DEC) {
                s.a ^= c.x; s.b ^= c.y; s.c ^= c.z; s.d ^= c.w;
            }
            if (!CHAIN && a.be32) {
                s.a = __byte_perm(s.a, 0, 0x0123); s.b = __byte_perm(s.b, 0, 0x0123);
                s.c = __byte_perm(s.c, 0, 0x0123); s.d = __byte_perm(s.d, 0, 0x0123);
            }
            if (live) st_stream(a.out + b, make_uint4(s.a, s.b, s.c, s.d));
        }
    }
}

/* CBC in-place pre-pass: save the ciphertext block in front of every 256-block
 * group before the main kernel overwrites it. */
__global__ void cbc_save_boundaries_kernel(const uint4 *in, uint4 *boundary, uint64_t n_groups)
{
    uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= 1 && g < n_groups) boundary[g] = in[g * 256 - 1];
}

/* ------------------------------------------- CBC encrypt, batched messages */

struct CbcEncArgs {
    const uint4 *in;
    uint4 *out;
    const uint4 *ivs;
    uint4 *ivs_out;
    uint64_t n_msgs;
    uint64_t blocks_per_msg;
};

/* One thread per message: the chain C_i = E(P_i ^ C_{i-1}) (aes.c:801-812) is
 * serial inside a message and independent across messages. */
template <int NR, bool CFB>
__global__ void __launch_bounds__(kThreads, 1)
aes_cbc_enc_batch_kernel(const __grid_constant__ AesSched ks, const __grid_constant__ CbcEncArgs a)
{
    extern __shared__ __align__(1024) uint32_t smem[];
    fill_tables<false>(smem);
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31u;
    Tb tb = { smem_base_u32(smem), lane * 4u };
    asm volatile("" : "+r"(tb.lane4));

    const uint64_t n_threads = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t m = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; m < a.n_msgs; m += n_threads) {
        uint4 c = a.ivs[m];
        const uint4 *pi = a.in + m * a.blocks_per_msg;
        uint4 *po = a.out + m * a.blocks_per_msg;
        /* CBC: C_i = E(P_i ^ C_{i-1}) (aes.c:801-812); CFB128: C_i = E(C_{i-1}) ^ P_i (aes.c:846-856) */
        for (uint64_t i = 0; i < a.blocks_per_msg; i++) {
            const uint4 p0 = pi[i];
            State s = CFB ? State{ c.x, c.y, c.z, c.w } : State{ p0.x ^ c.x, p0.y ^ c.y, p0.z ^ c.z, p0.w ^ c.w };
            s = cipher_block<NR, false>(s, ks, tb);
            c = CFB ? make_uint4(s.a ^ p0.x, s.b ^ p0.y, s.c ^ p0.z, s.d ^ p0.w) : make_uint4(s.a, s.b, s.c, s.d);
            po[i] = c;
        }
        if (a.ivs_out) a.ivs_out[m] = c;     /* iv <- last ciphertext block (aes.c:807) */
    }
}

/* TMA-staged form of the batch kernel (bc_tma.cuh): a warp owns 32 messages and moves
 * boxes of 32 messages x 64 bytes (4 blocks of each chain) with one bulk-tensor load
 * and one store, double buffered, so the global side is coalesced although every
 * lane walks its own message.  The chain value never leaves registers. */
constexpr int kCbcStages = 3;                /* boxes in flight per warp: no stall on the store of the previous slab */
constexpr int kCbcTmaMaxWarps = 16;          /* 128 KB tables + 16 x 6 KB staging fits 227 KB */
constexpr int kCbcWarpSmem = kCbcStages * kBoxBytes + kCbcStages * 8;

__device__ __forceinline__ void tma_store_wait_read_1()
{
    asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
}

template <int NR, bool CFB>
__global__ void __launch_bounds__(kCbcTmaMaxWarps * 32, 1)
aes_cbc_enc_batch_tma_kernel(const __grid_constant__ AesSched ks, const __grid_constant__ CbcEncArgs a,
                             const __grid_constant__ CUtensorMap tm_in, const __grid_constant__ CUtensorMap tm_out)
{
    extern __shared__ __align__(1024) uint32_t smem[];
    fill_tables<false>(smem);
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
    Tb tb = { smem_base_u32(smem), lane * 4u };
    const uint32_t stage = tb.base + kEncSmemBytes + warp * (kCbcStages * kBoxBytes);
    const uint32_t bar = tb.base + kEncSmemBytes + warps * (kCbcStages * kBoxBytes) + warp * (kCbcStages * 8);
    if (lane == 0) for (int i = 0; i < kCbcStages; i++) mbar_init(bar + 8 * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    asm volatile("" : "+r"(tb.lane4));

    uint32_t phases = 0u;
    const uint64_t msg_len = a.blocks_per_msg * 16;
    const uint32_t n_slabs = (uint32_t)((msg_len + kSlab - 1) / kSlab);
    const uint64_t n_groups = (a.n_msgs + 31) / 32;
    const uint64_t n_warps = (uint64_t)gridDim.x * warps;

    for (uint64_t mg = blockIdx.x + (uint64_t)gridDim.x * warp; mg < n_groups; mg += n_warps) {
        const uint64_t m = mg * 32 + lane;
        const bool live = m < a.n_msgs;
        const uint32_t row0 = (uint32_t)(mg * 32);
        if (lane == 0) {                                   /* two boxes ahead */
            mbar_expect_tx(bar, kBoxBytes);
            tma_load_2d(stage, &tm_in, 0, row0, bar);
            if (n_slabs > 1) {
                mbar_expect_tx(bar + 8, kBoxBytes);
                tma_load_2d(stage + kBoxBytes, &tm_in, kSlab, row0, bar + 8);
            }
        }
        uint4 c = make_uint4(0, 0, 0, 0);
        if (live) c = a.ivs[m];
        uint32_t buf = 0;
        for (uint32_t k = 0; k < n_slabs; k++) {
            const uint32_t sbuf = stage + buf * kBoxBytes;
            mbar_wait(bar + 8 * buf, (phases >> buf) & 1u);
            phases ^= 1u << buf;
            const uint32_t valid = (uint32_t)(msg_len - (uint64_t)k * kSlab < kSlab ? msg_len - (uint64_t)k * kSlab : kSlab);
#pragma unroll 1
            for (uint32_t q = 0; q < 4; q++) {
                if (q * 16 < valid) {
                    const uint32_t sa = sbuf + box_chunk_off(lane, q);
                    const uint4 pt = lds128(sa);
                    State s = CFB ? State{ c.x, c.y, c.z, c.w }
                                  : State{ pt.x ^ c.x, pt.y ^ c.y, pt.z ^ c.z, pt.w ^ c.w };    /* aes.c:803-804 / :849-853 */
                    s = cipher_block<NR, false>(s, ks, tb);
                    c = CFB ? make_uint4(s.a ^ pt.x, s.b ^ pt.y, s.c ^ pt.z, s.d ^ pt.w) : make_uint4(s.a, s.b, s.c, s.d);
                    sts128(sa, c);
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                tma_store_2d(&tm_out, k * kSlab, row0, sbuf);
                if (k + 2 < n_slabs) {
                    /* slab k+2 goes into the buffer slab k-1 was stored from: that store (issued a
                     * whole slab ago) must have left shared memory; the one just issued may still run */
                    const uint32_t nb = buf == 0 ? 2u : buf - 1u;
                    tma_store_wait_read_1();
                    mbar_expect_tx(bar + 8 * nb, kBoxBytes);
                    tma_load_2d(stage + nb * kBoxBytes, &tm_in, (k + 2) * kSlab, row0, bar + 8 * nb);
                }
            }
            buf = buf == kCbcStages - 1 ? 0u : buf + 1u;
        }
        if (lane == 0) tma_store_wait_read();              /* all buffers reusable by the next message group */
        __syncwarp();
        /* the next group starts at buffer 0 again; `phases` keeps each buffer's barrier parity */
        if (a.ivs_out && live) a.ivs_out[m] = c;     /* iv <- last ciphertext block (aes.c:807) */
    }
    if (lane == 0) tma_store_wait_all();
}

/* ------------------------------------------------------------ host helpers */

int check_common(const uint32_t *rk, int nr, const void *in, const void *out, size_t nbytes)
{
    if (!rk || (nr != 10 && nr != 12 && nr != 14)) {
        bc_set_error("bad key schedule (rk=%p, nr=%d; nr must be 10, 12 or 14)", (const void *)rk, nr);
        return BC_ERR_BAD_ARG;
    }
    if (nbytes && (!in || !out)) {
        bc_set_error("NULL data pointer with nbytes=%zu", nbytes);
        return BC_ERR_BAD_ARG;
    }
    if (((uintptr_t)in | (uintptr_t)out) & 15u) {
        bc_set_error("device buffers must be 16-byte aligned");
        return BC_ERR_BAD_ARG;
    }
    return BC_OK;
}

int check_overlap(const void *in, const void *out, size_t nbytes)
{
    const uintptr_t i = (uintptr_t)in, o = (uintptr_t)out;
    if (i != o && i < o + nbytes && o < i + nbytes) {
        bc_set_error("input and output overlap partially; they must be identical or disjoint");
        return BC_ERR_OVERLAP;
    }
    return BC_OK;
}

void load_sched(AesSched &ks, const uint32_t *rk, int nr)
{
    for (int i = 0; i < 60; i++) ks.rk[i] = (i < 4 * (nr + 1)) ? rk[i] : 0u;
}

/* One CTA per work item until every SM has one: small buffers spread over the
 * SMs instead of queueing on a few (items are dealt to CTAs round-robin). */
int grid_for(uint64_t items)
{
    int sms = bc_cached_sm_count();
    if (sms <= 0) return -1;
    if (items < 1) items = 1;
    return (int)(items < (uint64_t)sms ? items : (uint64_t)sms);
}

template <int NR, int THREADS>
int launch_ctr_t(const AesSched &ks, const CtrArgs &a, cudaStream_t st)
{
    int rc = set_smem<aes_ctr_kernel<NR, THREADS>>(kEncSmemBytes);
    if (rc) return rc;
    const int grid = grid_for(a.n_chunks);
    if (grid < 0) return BC_ERR_CUDA;
    aes_ctr_kernel<NR, THREADS><<<grid, THREADS, kEncSmemBytes, st>>>(ks, a);
    BC_CUDA(cudaGetLastError(), "aes_ctr_kernel launch");
    return BC_OK;
}

template <int NR>
int launch_ctr(const AesSched &ks, const CtrArgs &a, cudaStream_t st)
{
    return launch_ctr_t<NR, kCtrThreads>(ks, a, st);
}

template <int NR, int MODE>
int launch_blocks(const AesSched &ks, const BlkArgs &a, cudaStream_t st)
{
    constexpr int smem = (MODE == MODE_ECB_ENC || MODE == MODE_CFB_DEC) ? kEncSmemBytes : kDecSmemBytes;
    int rc = set_smem<aes_blocks_kernel<NR, MODE>>(smem);
    if (rc) return rc;
    int grid = grid_for(a.n_groups);
    if (grid < 0) return BC_ERR_CUDA;
    aes_blocks_kernel<NR, MODE><<<grid, kThreads, smem, st>>>(ks, a);
    BC_CUDA(cudaGetLastError(), "aes_blocks_kernel launch");
    return BC_OK;
}

template <int MODE>
int launch_blocks_nr(int nr, const AesSched &ks, const BlkArgs &a, cudaStream_t st)
{
    switch (nr) {
    case 10: return launch_blocks<10, MODE>(ks, a, st);
    case 12: return launch_blocks<12, MODE>(ks, a, st);
    default: return launch_blocks<14, MODE>(ks, a, st);
    }
}

template <int NR, bool CFB>
int launch_cbc_enc(const AesSched &ks, const CbcEncArgs &a, cudaStream_t st)
{
    const int sms = bc_cached_sm_count();
    if (sms <= 0) return BC_ERR_CUDA;
    const uint64_t msg_len = a.blocks_per_msg * 16;
    if (msg_len < ((uint64_t)1 << 32) && a.n_msgs < ((uint64_t)1 << 31)) {
        /* TMA-staged: warps per CTA chosen so the message groups spread evenly over the SMs */
        const uint64_t groups = (a.n_msgs + 31) / 32;
        uint64_t warps = (groups + sms - 1) / sms;
        if (warps > kCbcTmaMaxWarps) warps = kCbcTmaMaxWarps;
        if (warps < 1) warps = 1;
        const uint64_t ctas = (groups + warps - 1) / warps;
        const int grid = (int)(ctas < (uint64_t)sms ? ctas : (uint64_t)sms);
        const int smem = kEncSmemBytes + (int)warps * kCbcWarpSmem;
        CUtensorMap tin, tout;
        int rc = make_row_tensor_map(&tin, a.in, a.n_msgs, msg_len);
        if (rc) return rc;
        if ((rc = make_row_tensor_map(&tout, a.out, a.n_msgs, msg_len))) return rc;
        if ((rc = set_smem<aes_cbc_enc_batch_tma_kernel<NR, CFB>>(kEncSmemBytes + kCbcTmaMaxWarps * kCbcWarpSmem))) return rc;
        aes_cbc_enc_batch_tma_kernel<NR, CFB><<<grid, (int)warps * 32, smem, st>>>(ks, a, tin, tout);
        BC_CUDA(cudaGetLastError(), "aes_cbc_enc_batch_tma_kernel launch");
        return BC_OK;
    }
    int rc = set_smem<aes_cbc_enc_batch_kernel<NR, CFB>>(kEncSmemBytes);
    if (rc) return rc;
    /* spread the messages evenly: ceil(n/SMs) threads per CTA, whole warps, <= kThreads */
    uint64_t per = (a.n_msgs + sms - 1) / sms;
    per = ((per + 31) / 32) * 32;
    if (per > kThreads) per = kThreads;
    uint64_t ctas = (a.n_msgs + per - 1) / per;
    int grid = (int)(ctas < (uint64_t)sms ? ctas : (uint64_t)sms);
    aes_cbc_enc_batch_kernel<NR, CFB><<<grid, (int)per, kEncSmemBytes, st>>>(ks, a);
    BC_CUDA(cudaGetLastError(), "aes_cbc_enc_batch_kernel launch");
    return BC_OK;
}

}  // namespace

/* ------------------------------------------------------------- C-ABI (dev) */

extern "C" int bc_aes_ctr_dev(const uint32_t *rk, int nr, const unsigned char counter[16], int counter_bits,
                              uint64_t first_block, const void *d_in, void *d_out, size_t nbytes,
                              void *d_last_keystream, void *stream)
{
    int rc = check_common(rk, nr, d_in, d_out, nbytes);
    if (rc) return rc;
    if (!counter || (counter_bits != 128 && counter_bits != 64)) {
        bc_set_error("counter must be non-NULL and counter_bits 64 or 128 (got %d)", counter_bits);
        return BC_ERR_BAD_ARG;
    }
    if ((rc = check_overlap(d_in, d_out, nbytes))) return rc;
    if ((uintptr_t)d_last_keystream & 15u) {
        bc_set_error("d_last_keystream must be 16-byte aligned (it is written with one 128-bit store)");
        return BC_ERR_BAD_ARG;
    }
    if (nbytes == 0) return BC_OK;

    uint64_t hi = 0, lo = 0;
    for (int i = 0; i < 8; i++) { hi = (hi << 8) | counter[i]; lo = (lo << 8) | counter[8 + i]; }
    const uint64_t lo2 = lo + first_block;
    if (counter_bits == 128 && lo2 < lo) hi++;

    AesSched ks;
    load_sched(ks, rk, nr);
    CtrArgs a;
    a.c_lo = (uint32_t)(lo2 & 0xffu);
    a.base_lo = lo2 & ~(uint64_t)0xffu;
    a.base_hi = hi;
    a.carry128 = (counter_bits == 128);
    a.in = (const uint4 *)d_in;
    a.out = (uint4 *)d_out;
    a.nblocks = nbytes / 16;
    a.tail = (uint32_t)(nbytes % 16);
    const uint64_t nb_all = a.nblocks + (a.tail ? 1 : 0);
    /* 8 groups per warp trip amortise the per-group constants best; small buffers use
     * fewer so that there are still at least ~2 trips for every warp of a full grid */
    const uint64_t n_groups = (a.c_lo + nb_all + 255) / 256;
    const int sms = bc_cached_sm_count();
    if (sms <= 0) return BC_ERR_CUDA;
    a.gpc = kMaxGroupsPerChunk;
    while (a.gpc > 1 && n_groups / a.gpc < (uint64_t)sms * (kCtrThreads / 32) * 2) a.gpc >>= 1;
    const uint64_t per_chunk = (uint64_t)a.gpc * 256;
    a.n_chunks = (a.c_lo + nb_all + per_chunk - 1) / per_chunk;
    a.last_ks = (uint4 *)d_last_keystream;
    cudaStream_t st = (cudaStream_t)stream;
    switch (nr) {
    case 10: return launch_ctr<10>(ks, a, st);
    case 12: return launch_ctr<12>(ks, a, st);
    default: return launch_ctr<14>(ks, a, st);
    }
}

namespace {

int ecb_dev(const uint32_t *rk, int nr, int mode, const void *d_in, void *d_out, size_t nbytes, void *stream, uint32_t be32)
{
    int rc = check_common(rk, nr, d_in, d_out, nbytes);
    if (rc) return rc;
    if (mode != BC_ENCRYPT && mode != BC_DECRYPT) {
        bc_set_error("mode must be BC_ENCRYPT (1) or BC_DECRYPT (0), got %d", mode);
        return BC_ERR_BAD_ARG;
    }
    if ((rc = check_overlap(d_in, d_out, nbytes))) return rc;
    AesSched ks;
    load_sched(ks, rk, nr);
    BlkArgs a = {};
    a.in = (const uint4 *)d_in;
    a.out = (uint4 *)d_out;
    a.nblocks = nbytes / 16;            /* whole blocks only, like the harness loop */
    a.n_groups = (a.nblocks + 255) / 256;
    a.be32 = be32;
    if (a.nblocks == 0) return BC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    return mode == BC_ENCRYPT ? launch_blocks_nr<MODE_ECB_ENC>(nr, ks, a, st)
                              : launch_blocks_nr<MODE_ECB_DEC>(nr, ks, a, st);
}

}  // namespace

extern "C" int bc_aes_ecb_dev(const uint32_t *rk, int nr, int mode, const void *d_in, void *d_out,
                              size_t nbytes, void *stream)
{
    return ecb_dev(rk, nr, mode, d_in, d_out, nbytes, stream, 0u);
}

extern "C" int bc_aes_ecb_be32_dev(const uint32_t *rk, int nr, int mode, const void *d_in, void *d_out,
                                   size_t nbytes, void *stream)
{
    return ecb_dev(rk, nr, mode, d_in, d_out, nbytes, stream, 1u);
}

namespace {

/* shared body of the two block-parallel chained decrypts (CBC, CFB128) */
template <int MODE>
int chained_decrypt_dev(const uint32_t *rk, int nr, const unsigned char iv[16],
                        const void *d_in, void *d_out, size_t nbytes, void *stream, const char *what)
{
    int rc = check_common(rk, nr, d_in, d_out, nbytes);
    if (rc) return rc;
    if (!iv) { bc_set_error("iv is NULL"); return BC_ERR_BAD_ARG; }
    if (nbytes % 16) { bc_set_error("%s length %zu is not a multiple of 16", what, nbytes); return BC_ERR_BAD_ARG; }
    if ((rc = check_overlap(d_in, d_out, nbytes))) return rc;
    if (nbytes == 0) return BC_OK;
    AesSched ks;
    load_sched(ks, rk, nr);
    BlkArgs a = {};
    a.in = (const uint4 *)d_in;
    a.out = (uint4 *)d_out;
    a.nblocks = nbytes / 16;
    a.n_groups = (a.nblocks + 255) / 256;
    uint32_t w[4];
    for (int i = 0; i < 4; i++)
        w[i] = (uint32_t)iv[4 * i] | ((uint32_t)iv[4 * i + 1] << 8) | ((uint32_t)iv[4 * i + 2] << 16) |
               ((uint32_t)iv[4 * i + 3] << 24);
    a.iv = make_uint4(w[0], w[1], w[2], w[3]);
    cudaStream_t st = (cudaStream_t)stream;
    void *scratch = nullptr;
    if (d_in == d_out && a.n_groups > 1) {
        /* in place: the block in front of each group would be overwritten by a
         * neighbouring warp — keep copies in stream-ordered scratch first */
        if ((rc = bc_stream_scratch_alloc((void *)st, a.n_groups * sizeof(uint4), &scratch))) return rc;
        const int tpb = 256;
        const unsigned blocks = (unsigned)((a.n_groups + tpb - 1) / tpb);
        cbc_save_boundaries_kernel<<<blocks, tpb, 0, st>>>(a.in, (uint4 *)scratch, a.n_groups);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { bc_stream_scratch_free((void *)st, scratch); return bc_cuda_fail((int)e, "cbc_save_boundaries_kernel launch"); }
        a.boundary = (const uint4 *)scratch;
    }
    rc = launch_blocks_nr<MODE>(nr, ks, a, st);
    if (scratch) {                                   /* freed in stream order: after the kernel that reads it */
        const int rc2 = bc_stream_scratch_free((void *)st, scratch);
        if (rc == BC_OK) rc = rc2;
    }
    return rc;
}

template <bool CFB>
int chained_encrypt_batch_dev(const uint32_t *rk, int nr, size_t n_msgs, size_t msg_len,
                              const void *d_ivs, void *d_ivs_out, const void *d_in, void *d_out, void *stream)
{
    int rc = check_common(rk, nr, d_in, d_out, n_msgs * msg_len);
    if (rc) return rc;
    if (msg_len % 16) { bc_set_error("message length %zu is not a multiple of 16", msg_len); return BC_ERR_BAD_ARG; }
    if (n_msgs == 0 || msg_len == 0) return BC_OK;
    if (!d_ivs || ((uintptr_t)d_ivs & 15u) || ((uintptr_t)d_ivs_out & 15u)) {
        bc_set_error("d_ivs must be non-NULL and 16-byte aligned");
        return BC_ERR_BAD_ARG;
    }
    if ((rc = check_overlap(d_in, d_out, n_msgs * msg_len))) return rc;
    AesSched ks;
    load_sched(ks, rk, nr);
    CbcEncArgs a;
    a.in = (const uint4 *)d_in;
    a.out = (uint4 *)d_out;
    a.ivs = (const uint4 *)d_ivs;
    a.ivs_out = (uint4 *)d_ivs_out;
    a.n_msgs = n_msgs;
    a.blocks_per_msg = msg_len / 16;
    cudaStream_t st = (cudaStream_t)stream;
    switch (nr) {
    case 10: return launch_cbc_enc<10, CFB>(ks, a, st);
    case 12: return launch_cbc_enc<12, CFB>(ks, a, st);
    default: return launch_cbc_enc<14, CFB>(ks, a, st);
    }
}

}  // namespace

extern "C" int bc_aes_cbc_decrypt_dev(const uint32_t *rk, int nr, const unsigned char iv[16],
                                      const void *d_in, void *d_out, size_t nbytes, void *stream)
{
    return chained_decrypt_dev<MODE_CBC_DEC>(rk, nr, iv, d_in, d_out, nbytes, stream, "CBC");
}

extern "C" int bc_aes_cfb128_decrypt_dev(const uint32_t *rk, int nr, const unsigned char iv[16],
                                         const void *d_in, void *d_out, size_t nbytes, void *stream)
{
    return chained_decrypt_dev<MODE_CFB_DEC>(rk, nr, iv, d_in, d_out, nbytes, stream, "CFB128");
}

extern "C" int bc_aes_cbc_encrypt_batch_dev(const uint32_t *rk, int nr, size_t n_msgs, size_t msg_len,
                                            const void *d_ivs, void *d_ivs_out,
                                            const void *d_in, void *d_out, void *stream)
{
    return chained_encrypt_batch_dev<false>(rk, nr, n_msgs, msg_len, d_ivs, d_ivs_out, d_in, d_out, stream);
}

extern "C" int bc_aes_cfb128_encrypt_batch_dev(const uint32_t *rk, int nr, size_t n_msgs, size_t msg_len,
                                               const void *d_ivs, void *d_ivs_out,
                                               const void *d_in, void *d_out, void *stream)
{
    return chained_encrypt_batch_dev<true>(rk, nr, n_msgs, msg_len, d_ivs, d_ivs_out, d_in, d_out, stream);
}

/* ------------------------------------------------ allocation-free variant */

namespace {

/* Boundary-block scratch for bc_aes_cbc_decrypt_inplace_noalloc_dev: one 16-byte
 * ciphertext block per 4 KiB row-group.  It lives in the module image, so every
 * GPU that loads libbulkcipher.so carries this 1 MiB until the process exits. */
constexpr unsigned long long kNoAllocMaxGroups = 65536;                       /* 65536 x 4 KiB = 256 MiB per call */
constexpr unsigned long long kNoAllocMaxBytes  = kNoAllocMaxGroups * 4096ull;
static __device__ __align__(256) uint4 g_noalloc_boundary[kNoAllocMaxGroups];

constexpr int kNoAllocMaxDevices = 64;
uint4 *g_noalloc_dev_ptr[kNoAllocMaxDevices];   /* device address of the array, per device; 0 = not looked up yet */

/* The array's name is only a host-side handle; the address a kernel can use has
 * to be asked for, with the right device current. */
int noalloc_scratch(uint4 **out)
{
    int dev = -1;
    BC_CUDA(cudaGetDevice(&dev), "cudaGetDevice");
    if (dev < 0 || dev >= kNoAllocMaxDevices) {
        bc_set_error("device ordinal %d is outside the supported range 0..%d", dev, kNoAllocMaxDevices - 1);
        return BC_ERR_BAD_ARG;
    }
    uint4 *p = __atomic_load_n(&g_noalloc_dev_ptr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        void *sym = nullptr;
        BC_CUDA(cudaGetSymbolAddress(&sym, g_noalloc_boundary), "cudaGetSymbolAddress(g_noalloc_boundary)");
        p = (uint4 *)sym;
        __atomic_store_n(&g_noalloc_dev_ptr[dev], p, __ATOMIC_RELEASE);   /* racing threads store the same value */
    }
    *out = p;
    return BC_OK;
}

}  // namespace

/* See include/bulkcipher.h.  Same kernels as bc_aes_cbc_decrypt_dev with
 * d_in == d_out; the only difference is where the boundary blocks are parked. */
extern "C" int bc_aes_cbc_decrypt_inplace_noalloc_dev(const uint32_t *rk, int nr, const unsigned char iv[16],
                                                      void *d_buf, size_t nbytes, void *stream)
{
    int rc = check_common(rk, nr, d_buf, d_buf, nbytes);
    if (rc) return rc;
    if (!iv) { bc_set_error("iv is NULL"); return BC_ERR_BAD_ARG; }
    if (nbytes % 16) { bc_set_error("CBC length %zu is not a multiple of 16", nbytes); return BC_ERR_BAD_ARG; }
    if (nbytes > kNoAllocMaxBytes) {
        bc_set_error("bc_aes_cbc_decrypt_inplace_noalloc_dev handles at most %llu bytes per call (got %zu); "
                     "split the buffer on a 4096-byte multiple and chain the iv, or use bc_aes_cbc_decrypt_dev",
                     kNoAllocMaxBytes, nbytes);
        return BC_ERR_TOO_LARGE;
    }
    if (nbytes == 0) return BC_OK;
    uint4 *scratch = nullptr;
    if ((rc = noalloc_scratch(&scratch))) return rc;

    AesSched ks;
    load_sched(ks, rk, nr);
    BlkArgs a = {};
    a.in = (const uint4 *)d_buf;
    a.out = (uint4 *)d_buf;
    a.nblocks = nbytes / 16;
    a.n_groups = (a.nblocks + 255) / 256;
    uint32_t w[4];
    for (int i = 0; i < 4; i++)
        w[i] = (uint32_t)iv[4 * i] | ((uint32_t)iv[4 * i + 1] << 8) | ((uint32_t)iv[4 * i + 2] << 16) |
               ((uint32_t)iv[4 * i + 3] << 24);
    a.iv = make_uint4(w[0], w[1], w[2], w[3]);
    cudaStream_t st = (cudaStream_t)stream;
    if (a.n_groups > 1) {
        /* entries 1..n_groups-1 are rewritten here before the main kernel reads them, so
         * nothing left over from an earlier call is ever consumed and nothing needs zeroing */
        const int tpb = 256;
        const unsigned blocks = (unsigned)((a.n_groups + tpb - 1) / tpb);
        cbc_save_boundaries_kernel<<<blocks, tpb, 0, st>>>(a.in, scratch, a.n_groups);
        BC_CUDA(cudaGetLastError(), "cbc_save_boundaries_kernel launch");
        a.boundary = scratch;
    }
    return launch_blocks_nr<MODE_CBC_DEC>(nr, ks, a, st);
}
