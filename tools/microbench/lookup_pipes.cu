// lookup_pipes.cu — microbenchmark: can table lookups through the texture pipe run
// concurrently with shared-memory lookups (different data pipes?) on sm_100a.
// Each thread chases a dependent chain of lookups in a 256-entry u32 table.
//   mode 0: LDS only (bank-replicated table, conflict free)      N lookups
//   mode 1: TEX only (tex1Dfetch on a 1 KB linear texture)        N lookups
//   mode 2: LDS + TEX interleaved, R LDS per 1 TEX                N lookups total
//   mode 3: LDS + LDG(.nc, L1-resident) interleaved
// Prints lookups/clk/SM.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lookup_pipes lookup_pipes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int THREADS = 1024;
constexpr int ITERS = 2048;

template <int MODE, int RATIO>
__global__ void __launch_bounds__(THREADS, 1) k(cudaTextureObject_t tex, const uint32_t *gtab, uint32_t *out, long long *cycles)
{
    extern __shared__ uint32_t sm[];
    for (int i = threadIdx.x; i < 256 * 32; i += THREADS) sm[i] = gtab[i >> 5];
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31;
    uint32_t a = threadIdx.x * 2654435761u, b = a ^ 0x9e3779b9u, c = a + 77u, d = b + 99u;   // 4 independent chains
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int r = 0; r < RATIO; r++) {
            if (MODE != 1) {
                a ^= sm[((a >> 8) & 0xff) * 32 + lane];
                b ^= sm[((b >> 8) & 0xff) * 32 + lane];
                c ^= sm[((c >> 8) & 0xff) * 32 + lane];
                d ^= sm[((d >> 8) & 0xff) * 32 + lane];
            }
        }
        if (MODE == 1 || MODE == 2) {
            a ^= tex1Dfetch<unsigned>(tex, (a >> 16) & 0xff);
            b ^= tex1Dfetch<unsigned>(tex, (b >> 16) & 0xff);
            if (MODE == 1) {
                c ^= tex1Dfetch<unsigned>(tex, (c >> 16) & 0xff);
                d ^= tex1Dfetch<unsigned>(tex, (d >> 16) & 0xff);
            }
        }
        if (MODE == 3) {
            a ^= __ldg(gtab + ((a >> 16) & 0xff));
            b ^= __ldg(gtab + ((b >> 16) & 0xff));
        }
    }
    long long t1 = clock64();
    out[blockIdx.x * THREADS + threadIdx.x] = a ^ b ^ c ^ d;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

// LOP3 issue rate: 8 independent chains of 3-input logic ops per thread (what a bitsliced AES is made of)
__global__ void __launch_bounds__(THREADS, 1) k_lop3(uint32_t *out, long long *cycles, uint32_t seed)
{
    uint32_t r[8];
    for (int i = 0; i < 8; i++) r[i] = threadIdx.x * 2654435761u + seed * (i + 1);
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
#pragma unroll
            for (int i = 0; i < 8; i++)
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r[i]) : "r"(r[(i + 3) & 7]), "r"(r[(i + 5) & 7]));
        }
    }
    long long t1 = clock64();
    uint32_t x = 0;
    for (int i = 0; i < 8; i++) x ^= r[i];
    out[blockIdx.x * THREADS + threadIdx.x] = x;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

// Peak rate of conflict-free 32-bit shared-memory lookups: 16 independent lookups per trip from 16
// index registers (the shape of one AES round), results folded with LOP3.
__global__ void __launch_bounds__(THREADS, 1) k_lds_peak(const uint32_t *gtab, uint32_t *out, long long *cycles)
{
    extern __shared__ uint32_t sm[];
    for (int i = threadIdx.x; i < 256 * 32; i += THREADS) sm[i] = gtab[i >> 5];
    __syncthreads();
    // same layout and addressing as the AES kernels: entry i of lane L at byte i*128 + L*4 here
    // (one table), address = one PRMT (index byte -> bits 7.., needs a shift) -- use the 256-byte
    // row variant instead: rows of 256 B, only the first 128 B used, so the index lands in byte 1
    __syncthreads();
    for (int i = threadIdx.x; i < 256 * 32; i += THREADS) sm[(i >> 5) * 64 + (i & 31)] = gtab[i >> 5];
    __syncthreads();
    const uint32_t lane4 = (threadIdx.x & 31) * 4;
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(sm);
    uint32_t s[4] = { threadIdx.x * 2654435761u, threadIdx.x * 40503u + 1, threadIdx.x * 9176u + 7, threadIdx.x + 99u };
    auto look = [&](uint32_t x, int k) {
        uint32_t a = __byte_perm(x, lane4, 0x5504 | (k << 4)), v;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(base + a));
        return v;
    };
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
        uint32_t t[4];
#pragma unroll
        for (int c = 0; c < 4; c++)
            t[c] = look(s[c], 0) ^ look(s[(c + 1) & 3], 1) ^ look(s[(c + 2) & 3], 2) ^ look(s[(c + 3) & 3], 3);
#pragma unroll
        for (int c = 0; c < 4; c++) s[c] = t[c];
    }
    long long t1 = clock64();
    out[blockIdx.x * THREADS + threadIdx.x] = s[0] ^ s[1] ^ s[2] ^ s[3];
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE, int RATIO>
int run(const char *name, cudaTextureObject_t tex, uint32_t *gtab, uint32_t *out, long long *cyc, int sms)
{
    CK(cudaFuncSetAttribute(k<MODE, RATIO>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE, RATIO><<<sms, THREADS, 32768>>>(tex, gtab, out, cyc);
    CK(cudaDeviceSynchronize());
    cudaEventRecord(e0);
    k<MODE, RATIO><<<sms, THREADS, 32768>>>(tex, gtab, out, cyc);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long c0; cudaMemcpy(&c0, cyc, 8, cudaMemcpyDeviceToHost);
    double lds = (MODE == 1) ? 0 : 4.0 * RATIO, texn = (MODE == 1) ? 4 : (MODE == 2 ? 2 : 0), ldg = (MODE == 3) ? 2 : 0;
    double per_thread = (lds + texn + ldg) * ITERS;
    double lanes_per_clk = per_thread * THREADS / (double)c0;
    printf("%-34s %8.3f ms  cycles %10lld  lookups/clk/SM %6.2f  (lds %.0f tex %.0f ldg %.0f per trip)\n", name, ms, c0, lanes_per_clk, lds, texn, ldg);
    return 0;
}

int main()
{
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    uint32_t h[256];
    for (int i = 0; i < 256; i++) h[i] = i * 2654435761u ^ (i << 13);
    uint32_t *gtab, *out; long long *cyc;
    CK(cudaMalloc(&gtab, 1024)); CK(cudaMalloc(&out, (size_t)sms * THREADS * 4)); CK(cudaMalloc(&cyc, sms * 8));
    CK(cudaMemcpy(gtab, h, 1024, cudaMemcpyHostToDevice));
    cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeLinear; rd.res.linear.devPtr = gtab;
    rd.res.linear.desc = cudaCreateChannelDesc<unsigned>(); rd.res.linear.sizeInBytes = 1024;
    cudaTextureDesc td = {}; td.readMode = cudaReadModeElementType;
    cudaTextureObject_t tex; CK(cudaCreateTextureObject(&tex, &rd, &td, nullptr));
    k_lop3<<<sms, THREADS>>>(out, cyc, 1);
    CK(cudaDeviceSynchronize());
    k_lop3<<<sms, THREADS>>>(out, cyc, 2);
    CK(cudaDeviceSynchronize());
    {
        long long c0; cudaMemcpy(&c0, cyc, 8, cudaMemcpyDeviceToHost);
        double ops = 8.0 * 16 * ITERS * THREADS;
        printf("%-34s cycles %10lld  LOP3 lane-ops/clk/SM %6.2f\n", "LOP3 only (8 chains)", c0, ops / (double)c0);
    }
    CK(cudaFuncSetAttribute(k_lds_peak, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    k_lds_peak<<<sms, THREADS, 65536>>>(gtab, out, cyc);
    CK(cudaDeviceSynchronize());
    k_lds_peak<<<sms, THREADS, 65536>>>(gtab, out, cyc);
    CK(cudaDeviceSynchronize());
    {
        long long c0; cudaMemcpy(&c0, cyc, 8, cudaMemcpyDeviceToHost);
        printf("%-34s cycles %10lld  lookups/clk/SM %6.2f\n", "LDS peak (16 independent/trip)", c0, 16.0 * ITERS * THREADS / (double)c0);
    }
    run<0, 4>("LDS only (16/trip)", tex, gtab, out, cyc, sms);
    run<1, 1>("TEX only (4/trip)", tex, gtab, out, cyc, sms);
    run<2, 4>("LDS 16 + TEX 2", tex, gtab, out, cyc, sms);
    run<2, 2>("LDS 8 + TEX 2", tex, gtab, out, cyc, sms);
    run<2, 1>("LDS 4 + TEX 2", tex, gtab, out, cyc, sms);
    run<3, 4>("LDS 16 + LDG.nc 2", tex, gtab, out, cyc, sms);
    run<3, 1>("LDS 4 + LDG.nc 2", tex, gtab, out, cyc, sms);
    return 0;
}
