// bitslice_aes.cu — microbenchmark: the register-resident BITSLICED AES-128 core on sm_100a,
// measured instead of estimated.
//
// One thread holds 32 blocks as 128 x 32-bit registers: register [byte j][bit b] carries bit b of
// state byte j of all 32 blocks.  SubBytes is the Boyar-Peralta straight-line circuit (the 128-gate
// listing: 32 AND, 96 XOR/XNOR; checked exhaustively against the S-box on the host below) applied to
// the 16 byte positions; ShiftRows is a renaming; MixColumns and AddRoundKey are XORs (round-key
// bits are 0 / 0xFFFFFFFF words in constant memory, a free LOP3 operand).  ptxas fuses the gates
// into LOP3.  NO I/O and NO transposes inside the timed loop: the state is loaded once (already
// bitsliced by the host), encrypted ITERS times (each output is the next input), stored once.
// This is the best case for the formulation — a real CTR kernel would add the 128x32 bit transpose
// of the keystream (~60 more logic ops per block) and the XOR with the data.
//
// Rounds reproduced: AES_FROUND x 9 + last round (aes-modes/aes.c:601-622, 721-743); checked
// against FIPS-197 C.1 and 31 other blocks with a plain host AES.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o bitslice_aes bitslice_aes.cu
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

#ifndef BS_THREADS
#define BS_THREADS 128
#endif
#ifndef BS_MIN_BLOCKS
#define BS_MIN_BLOCKS 2
#endif

__constant__ uint32_t c_rk[11 * 128];      // round r, byte j, bit b -> 0 or 0xFFFFFFFF

// Boyar-Peralta S-box on 8 bit-planes.  u[7] = most significant bit (U0 of the paper) ... u[0] = LSB (U7).
__host__ __device__ __forceinline__ void sbox_bp(uint32_t *x)
{
    const uint32_t U0 = x[7], U1 = x[6], U2 = x[5], U3 = x[4], U4 = x[3], U5 = x[2], U6 = x[1], U7 = x[0];
    const uint32_t T1 = U0 ^ U3;
    const uint32_t T2 = U0 ^ U5;
    const uint32_t T3 = U0 ^ U6;
    const uint32_t T4 = U3 ^ U5;
    const uint32_t T5 = U4 ^ U6;
    const uint32_t T6 = T1 ^ T5;
    const uint32_t T7 = U1 ^ U2;
    const uint32_t T8 = U7 ^ T6;
    const uint32_t T9 = U7 ^ T7;
    const uint32_t T10 = T6 ^ T7;
    const uint32_t T11 = U1 ^ U5;
    const uint32_t T12 = U2 ^ U5;
    const uint32_t T13 = T3 ^ T4;
    const uint32_t T14 = T6 ^ T11;
    const uint32_t T15 = T5 ^ T11;
    const uint32_t T16 = T5 ^ T12;
    const uint32_t T17 = T9 ^ T16;
    const uint32_t T18 = U3 ^ U7;
    const uint32_t T19 = T7 ^ T18;
    const uint32_t T20 = T1 ^ T19;
    const uint32_t T21 = U6 ^ U7;
    const uint32_t T22 = T7 ^ T21;
    const uint32_t T23 = T2 ^ T22;
    const uint32_t T24 = T2 ^ T10;
    const uint32_t T25 = T20 ^ T17;
    const uint32_t T26 = T3 ^ T16;
    const uint32_t T27 = T1 ^ T12;
    const uint32_t M1 = T13 & T6;
    const uint32_t M2 = T23 & T8;
    const uint32_t M3 = T14 ^ M1;
    const uint32_t M4 = T19 & U7;
    const uint32_t M5 = M4 ^ M1;
    const uint32_t M6 = T3 & T16;
    const uint32_t M7 = T22 & T9;
    const uint32_t M8 = T26 ^ M6;
    const uint32_t M9 = T20 & T17;
    const uint32_t M10 = M9 ^ M6;
    const uint32_t M11 = T1 & T15;
    const uint32_t M12 = T4 & T27;
    const uint32_t M13 = M12 ^ M11;
    const uint32_t M14 = T2 & T10;
    const uint32_t M15 = M14 ^ M11;
    const uint32_t M16 = M3 ^ M2;
    const uint32_t M17 = M5 ^ T24;
    const uint32_t M18 = M8 ^ M7;
    const uint32_t M19 = M10 ^ M15;
    const uint32_t M20 = M16 ^ M13;
    const uint32_t M21 = M17 ^ M15;
    const uint32_t M22 = M18 ^ M13;
    const uint32_t M23 = M19 ^ T25;
    const uint32_t M24 = M22 ^ M23;
    const uint32_t M25 = M22 & M20;
    const uint32_t M26 = M21 ^ M25;
    const uint32_t M27 = M20 ^ M21;
    const uint32_t M28 = M23 ^ M25;
    const uint32_t M29 = M28 & M27;
    const uint32_t M30 = M26 & M24;
    const uint32_t M31 = M20 & M23;
    const uint32_t M32 = M27 & M31;
    const uint32_t M33 = M27 ^ M25;
    const uint32_t M34 = M21 & M22;
    const uint32_t M35 = M24 & M34;
    const uint32_t M36 = M24 ^ M25;
    const uint32_t M37 = M21 ^ M29;
    const uint32_t M38 = M32 ^ M33;
    const uint32_t M39 = M23 ^ M30;
    const uint32_t M40 = M35 ^ M36;
    const uint32_t M41 = M38 ^ M40;
    const uint32_t M42 = M37 ^ M39;
    const uint32_t M43 = M37 ^ M38;
    const uint32_t M44 = M39 ^ M40;
    const uint32_t M45 = M42 ^ M41;
    const uint32_t M46 = M44 & T6;
    const uint32_t M47 = M40 & T8;
    const uint32_t M48 = M39 & U7;
    const uint32_t M49 = M43 & T16;
    const uint32_t M50 = M38 & T9;
    const uint32_t M51 = M37 & T17;
    const uint32_t M52 = M42 & T15;
    const uint32_t M53 = M45 & T27;
    const uint32_t M54 = M41 & T10;
    const uint32_t M55 = M44 & T13;
    const uint32_t M56 = M40 & T23;
    const uint32_t M57 = M39 & T19;
    const uint32_t M58 = M43 & T3;
    const uint32_t M59 = M38 & T22;
    const uint32_t M60 = M37 & T20;
    const uint32_t M61 = M42 & T1;
    const uint32_t M62 = M45 & T4;
    const uint32_t M63 = M41 & T2;
    const uint32_t L0 = M61 ^ M62;
    const uint32_t L1 = M50 ^ M56;
    const uint32_t L2 = M46 ^ M48;
    const uint32_t L3 = M47 ^ M55;
    const uint32_t L4 = M54 ^ M58;
    const uint32_t L5 = M49 ^ M61;
    const uint32_t L6 = M62 ^ L5;
    const uint32_t L7 = M46 ^ L3;
    const uint32_t L8 = M51 ^ M59;
    const uint32_t L9 = M52 ^ M53;
    const uint32_t L10 = M53 ^ L4;
    const uint32_t L11 = M60 ^ L2;
    const uint32_t L12 = M48 ^ M51;
    const uint32_t L13 = M50 ^ L0;
    const uint32_t L14 = M52 ^ M61;
    const uint32_t L15 = M55 ^ L1;
    const uint32_t L16 = M56 ^ L0;
    const uint32_t L17 = M57 ^ L1;
    const uint32_t L18 = M58 ^ L8;
    const uint32_t L19 = M63 ^ L4;
    const uint32_t L20 = L0 ^ L1;
    const uint32_t L21 = L1 ^ L7;
    const uint32_t L22 = L3 ^ L12;
    const uint32_t L23 = L18 ^ L2;
    const uint32_t L24 = L15 ^ L9;
    const uint32_t L25 = L6 ^ L10;
    const uint32_t L26 = L7 ^ L9;
    const uint32_t L27 = L8 ^ L10;
    const uint32_t L28 = L11 ^ L14;
    const uint32_t L29 = L11 ^ L17;
    const uint32_t S0 = L6 ^ L24;
    const uint32_t S1 = ~(L16 ^ L26);
    const uint32_t S2 = ~(L19 ^ L28);
    const uint32_t S3 = L6 ^ L21;
    const uint32_t S4 = L20 ^ L22;
    const uint32_t S5 = L25 ^ L29;
    const uint32_t S6 = ~(L13 ^ L27);
    const uint32_t S7 = ~(L6 ^ L23);
    x[7] = S0; x[6] = S1; x[5] = S2; x[4] = S3; x[3] = S4; x[2] = S5; x[1] = S6; x[0] = S7;
}

// one full round on the bitsliced state: SubBytes, ShiftRows, MixColumns (unless LAST), AddRoundKey
template <bool LAST>
__device__ __forceinline__ void round_bs(uint32_t (&s)[128], const uint32_t *rk)
{
#pragma unroll
    for (int j = 0; j < 16; j++) sbox_bp(&s[8 * j]);
    uint32_t t[128];
#pragma unroll
    for (int c = 0; c < 4; c++)
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int b = 0; b < 8; b++) t[8 * (4 * c + r) + b] = s[8 * (4 * ((c + r) & 3) + r) + b];   // ShiftRows
    if (LAST) {
#pragma unroll
        for (int i = 0; i < 128; i++) s[i] = t[i] ^ rk[i];
        return;
    }
#pragma unroll
    for (int c = 0; c < 4; c++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const uint32_t *a0 = &t[8 * (4 * c + r)], *a1 = &t[8 * (4 * c + ((r + 1) & 3))];
            const uint32_t *a2 = &t[8 * (4 * c + ((r + 2) & 3))], *a3 = &t[8 * (4 * c + ((r + 3) & 3))];
            uint32_t x[8];
#pragma unroll
            for (int b = 0; b < 8; b++) x[b] = a0[b] ^ a1[b];                  // 2*a0 ^ 3*a1 = xtime(a0 ^ a1) ^ a1
            const uint32_t h = x[7];
            uint32_t *o = &s[8 * (4 * c + r)];
            const uint32_t *k = &rk[8 * (4 * c + r)];
            o[0] = h ^ a1[0] ^ a2[0] ^ a3[0] ^ k[0];
            o[1] = x[0] ^ h ^ a1[1] ^ a2[1] ^ a3[1] ^ k[1];
            o[2] = x[1] ^ a1[2] ^ a2[2] ^ a3[2] ^ k[2];
            o[3] = x[2] ^ h ^ a1[3] ^ a2[3] ^ a3[3] ^ k[3];
            o[4] = x[3] ^ h ^ a1[4] ^ a2[4] ^ a3[4] ^ k[4];
            o[5] = x[4] ^ a1[5] ^ a2[5] ^ a3[5] ^ k[5];
            o[6] = x[5] ^ a1[6] ^ a2[6] ^ a3[6] ^ k[6];
            o[7] = x[6] ^ a1[7] ^ a2[7] ^ a3[7] ^ k[7];
        }
    }
}

__global__ void __launch_bounds__(BS_THREADS, BS_MIN_BLOCKS)
bitslice_aes_kernel(const uint32_t *in, uint32_t *out, int iters, long long *cycles)
{
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, nt = gridDim.x * blockDim.x;
    uint32_t s[128];
#pragma unroll
    for (int i = 0; i < 128; i++) s[i] = in[(size_t)i * nt + tid];
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 128; i++) s[i] ^= c_rk[i];
#pragma unroll 1
        for (int r = 1; r < 10; r++) round_bs<false>(s, c_rk + 128 * r);
        round_bs<true>(s, c_rk + 1280);
    }
    const long long t1 = clock64();
#pragma unroll
    for (int i = 0; i < 128; i++) out[(size_t)i * nt + tid] = s[i];
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

// ------------------------------------------------------------------ host reference (plain AES-128)
static uint8_t SB[256];
static uint8_t gmul(uint8_t a, uint8_t b) { uint8_t r = 0; while (b) { if (b & 1) r ^= a; a = (uint8_t)((a << 1) ^ ((a & 0x80) ? 0x1b : 0)); b >>= 1; } return r; }
static void make_sbox()
{
    for (int x = 0; x < 256; x++) {
        uint8_t inv = 0;
        for (int y = 1; y < 256 && x; y++) if (gmul((uint8_t)x, (uint8_t)y) == 1) { inv = (uint8_t)y; break; }
        uint8_t r = 0;
        for (int i = 0; i < 8; i++)
            r |= (uint8_t)((((inv >> i) ^ (inv >> ((i + 4) & 7)) ^ (inv >> ((i + 5) & 7)) ^ (inv >> ((i + 6) & 7)) ^ (inv >> ((i + 7) & 7)) ^ (0x63 >> i)) & 1) << i);
        SB[x] = r;
    }
}
static void expand128(const uint8_t key[16], uint8_t rk[176])
{
    memcpy(rk, key, 16);
    uint8_t rcon = 1;
    for (int i = 16; i < 176; i += 4) {
        uint8_t t[4] = { rk[i - 4], rk[i - 3], rk[i - 2], rk[i - 1] };
        if (i % 16 == 0) { uint8_t u = t[0]; t[0] = SB[t[1]] ^ rcon; t[1] = SB[t[2]]; t[2] = SB[t[3]]; t[3] = SB[u]; rcon = gmul(rcon, 2); }
        for (int k = 0; k < 4; k++) rk[i + k] = rk[i - 16 + k] ^ t[k];
    }
}
static void encrypt_ref(const uint8_t rk[176], uint8_t b[16])
{
    for (int i = 0; i < 16; i++) b[i] ^= rk[i];
    for (int r = 1; r <= 10; r++) {
        uint8_t t[16];
        for (int c = 0; c < 4; c++) for (int row = 0; row < 4; row++) t[4 * c + row] = SB[b[4 * ((c + row) & 3) + row]];
        if (r < 10) {
            for (int c = 0; c < 4; c++) {
                uint8_t *a = &t[4 * c], o[4];
                for (int row = 0; row < 4; row++)
                    o[row] = gmul(a[row], 2) ^ gmul(a[(row + 1) & 3], 3) ^ a[(row + 2) & 3] ^ a[(row + 3) & 3];
                memcpy(a, o, 4);
            }
        }
        for (int i = 0; i < 16; i++) b[i] = t[i] ^ rk[16 * r + i];
    }
}

int main(int argc, char **argv)
{
    const int iters = argc > 1 ? atoi(argv[1]) : 256;
    make_sbox();
    // the circuit against the S-box, exhaustively (host evaluation of the same code the device runs)
    for (int x = 0; x < 256; x++) {
        uint32_t p[8];
        for (int b = 0; b < 8; b++) p[b] = ((x >> b) & 1) ? 0xFFFFFFFFu : 0u;
        sbox_bp(p);
        int y = 0;
        for (int b = 0; b < 8; b++) y |= (p[b] & 1) << b;
        if (y != SB[x]) { printf("S-box circuit wrong at %02x\n", x); return 1; }
    }
    int dev = 0, sms = 0, khz = 0;
    CK(cudaGetDevice(&dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, bitslice_aes_kernel));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, bitslice_aes_kernel, BS_THREADS, 0));
    const int grid = sms * occ, nt = grid * BS_THREADS;
    printf("bitslice_aes: %d regs/thread, %zu B local (spills), %d threads/CTA, %d CTAs/SM, grid %d\n",
           fa.numRegs, (size_t)fa.localSizeBytes, BS_THREADS, occ, grid);

    uint8_t key[16], rk[176];
    for (int i = 0; i < 16; i++) key[i] = (uint8_t)i;                       // FIPS-197 C.1 key
    expand128(key, rk);
    std::vector<uint32_t> rkb(11 * 128);
    for (int r = 0; r < 11; r++) for (int j = 0; j < 16; j++) for (int b = 0; b < 8; b++)
        rkb[128 * r + 8 * j + b] = ((rk[16 * r + j] >> b) & 1) ? 0xFFFFFFFFu : 0u;
    CK(cudaMemcpyToSymbol(c_rk, rkb.data(), rkb.size() * 4));

    // thread t, block k: the C.1 plaintext with bytes 12..15 ^= (t*32+k) — block 0 of thread 0 IS the C.1 vector
    std::vector<uint32_t> h_in((size_t)128 * nt, 0), h_out((size_t)128 * nt);
    auto plain = [&](int t, int k, uint8_t p[16]) {
        for (int i = 0; i < 16; i++) p[i] = (uint8_t)(i * 0x11);
        const uint32_t v = (uint32_t)t * 32 + k;
        p[12] ^= (uint8_t)(v >> 24); p[13] ^= (uint8_t)(v >> 16); p[14] ^= (uint8_t)(v >> 8); p[15] ^= (uint8_t)v;
    };
    for (int t = 0; t < nt; t++) for (int k = 0; k < 32; k++) {
        uint8_t p[16]; plain(t, k, p);
        for (int j = 0; j < 16; j++) for (int b = 0; b < 8; b++)
            if ((p[j] >> b) & 1) h_in[(size_t)(8 * j + b) * nt + t] |= 1u << k;
    }
    uint32_t *d_in, *d_out; long long *d_cyc;
    CK(cudaMalloc(&d_in, h_in.size() * 4)); CK(cudaMalloc(&d_out, h_out.size() * 4)); CK(cudaMalloc(&d_cyc, grid * 8));
    CK(cudaMemcpy(d_in, h_in.data(), h_in.size() * 4, cudaMemcpyHostToDevice));

    bitslice_aes_kernel<<<grid, BS_THREADS>>>(d_in, d_out, 1, d_cyc);          // one encryption: parity
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(h_out.data(), d_out, h_out.size() * 4, cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int t = 0; t < nt && bad < 5; t += (t < 4 ? 1 : 997)) for (int k = 0; k < 32; k++) {
        uint8_t p[16], c[16] = { 0 }; plain(t, k, p); encrypt_ref(rk, p);
        for (int j = 0; j < 16; j++) for (int b = 0; b < 8; b++) c[j] |= (uint8_t)(((h_out[(size_t)(8 * j + b) * nt + t] >> k) & 1) << b);
        if (memcmp(c, p, 16)) { bad++; printf("MISMATCH thread %d block %d\n", t, k); }
        if (t == 0 && k == 0) {
            static const uint8_t c1[16] = { 0x69,0xc4,0xe0,0xd8,0x6a,0x7b,0x04,0x30,0xd8,0xcd,0xb7,0x80,0x70,0xb4,0xc5,0x5a };
            if (memcmp(c, c1, 16)) { bad++; printf("FIPS-197 C.1 MISMATCH\n"); } else printf("FIPS-197 C.1 ok\n");
        }
    }
    if (bad) return 1;
    printf("parity ok (sampled threads x 32 blocks vs host AES)\n");

    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e9f; long long cyc = 0;
    for (int rep = 0; rep < 5; rep++) {
        CK(cudaEventRecord(e0));
        bitslice_aes_kernel<<<grid, BS_THREADS>>>(d_in, d_out, iters, d_cyc);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) { best = ms; CK(cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost)); }
    }
    const double blocks = (double)nt * 32 * iters;
    const double clk_per_s = khz * 1e3;                                      // nominal SM clock (1.965 GHz on B200)
    const double blk_per_clk_sm = blocks / (best * 1e-3) / clk_per_s / sms;
    printf("time %.3f ms -> %.4f blocks/clk/SM = %.3f B/clk/SM at the nominal %.0f MHz (CTA 0 counted %lld clock64 ticks)\n",
           best, blk_per_clk_sm, 16 * blk_per_clk_sm, khz / 1e3, cyc);
    printf("core-only throughput %.1f GB/s (16 B x %.3g blocks / %.3f ms); table kernel: 3.5 B/clk/SM, 1024 GB/s end to end\n",
           blocks * 16 / (best * 1e6), blocks, best);
    return 0;
}
