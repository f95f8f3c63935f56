// fill_probe.cu — where does the ~10 us floor of a small AES launch go?  Times, per variant, a
// 1024-thread CTA with 128 KB of dynamic shared memory (grid 1 and grid 148):
//   0 empty kernel            launch + CTA start-up only
//   1 fill as shipped         FT0..FT3 x 32 banks from c_sbox + xtime (bc_aes.cu fill_tables)
//   2 fill from a 1 KB FT0    constant u32 table, rotations by PRMT
//   3 fill by 1-D TMA         cp.async.bulk of a prebuilt 128 KB image in global memory (L2 resident)
// Prints kernel time (back-to-back launches, CUDA events) and clock64 cycles spent in the fill.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../our-tree_b200/csrc -o fill_probe fill_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "bc_sbox.inc"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__constant__ uint8_t c_sbox[256] = { BC_SBOX_INIT };
__constant__ uint32_t c_ft0[256];
constexpr int SMEM = 131072;

__device__ __forceinline__ uint32_t rotl32(uint32_t v, int n) { return __funnelshift_l(v, v, n); }
__device__ __forceinline__ uint32_t xtime8(uint32_t x) { return ((x << 1) ^ ((x & 0x80u) ? 0x1bu : 0u)) & 0xffu; }
__device__ __forceinline__ uint32_t ft0_word(uint32_t i) { uint32_t s = c_sbox[i], d = xtime8(s); return d | (s << 8) | (s << 16) | ((d ^ s) << 24); }

template <int MODE>
__global__ void __launch_bounds__(1024, 1) k(const uint4 *image, uint32_t *out, long long *cyc)
{
    extern __shared__ __align__(1024) uint32_t sm[];
    const long long t0 = clock64();
    if (MODE == 1 || MODE == 2) {
        for (uint32_t idx = threadIdx.x; idx < 256u * 32u; idx += blockDim.x) {
            uint32_t i = idx >> 5, l = idx & 31u;
            uint32_t w = MODE == 1 ? ft0_word(i) : c_ft0[i];
            sm[i * 64 + l] = w; sm[i * 64 + 32 + l] = rotl32(w, 8);
            sm[16384 + i * 64 + l] = rotl32(w, 16); sm[16384 + i * 64 + 32 + l] = rotl32(w, 24);
        }
    } else if (MODE == 3) {
        __shared__ __align__(8) uint64_t bar;
        const uint32_t b = (uint32_t)__cvta_generic_to_shared(&bar), dst = (uint32_t)__cvta_generic_to_shared(sm);
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(b) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(b), "r"(SMEM) : "memory");
            for (int c = 0; c < 4; c++)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             :: "r"(dst + c * 32768), "l"((const char *)image + c * 32768), "r"(32768), "r"(b) : "memory");
        }
        __syncthreads();
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(b) : "memory");
    }
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    if (MODE != 0) out[blockIdx.x * 1024 + threadIdx.x] = sm[(threadIdx.x * 37u) & 32767u];
}

template <int MODE>
int run(const uint4 *image, uint32_t *out, long long *cyc, int grid, const char *name)
{
    CK(cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    for (int i = 0; i < 5; i++) k<MODE><<<grid, 1024, SMEM>>>(image, out, cyc);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const int reps = 200;
    CK(cudaEventRecord(e0));
    for (int i = 0; i < reps; i++) k<MODE><<<grid, 1024, SMEM>>>(image, out, cyc);
    CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    long long c; CK(cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost));
    printf("%-22s grid %3d: %.2f us per launch (back to back), fill = %lld clk in CTA 0\n", name, grid, ms * 1e3 / reps, c);
    return 0;
}

int main()
{
    uint32_t ft0[256];
    static const uint8_t sb[256] = { BC_SBOX_INIT };
    for (int i = 0; i < 256; i++) { uint32_t s = sb[i], d = ((s << 1) ^ ((s & 0x80) ? 0x1b : 0)) & 0xff; ft0[i] = d | (s << 8) | (s << 16) | ((d ^ s) << 24); }
    CK(cudaMemcpyToSymbol(c_ft0, ft0, sizeof ft0));
    uint4 *image; uint32_t *out; long long *cyc;
    CK(cudaMalloc(&image, SMEM)); CK(cudaMemset(image, 1, SMEM));
    CK(cudaMalloc(&out, 148 * 1024 * 4)); CK(cudaMalloc(&cyc, 148 * 8));
    for (int grid : { 1, 148 }) {
        if (run<0>(image, out, cyc, grid, "empty") || run<1>(image, out, cyc, grid, "fill (shipped)") ||
            run<2>(image, out, cyc, grid, "fill (const FT0)") || run<3>(image, out, cyc, grid, "fill (1-D TMA image)")) return 1;
    }
    return 0;
}
