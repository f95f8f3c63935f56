/*
 * bc_util.cu — harness-side device helpers of libbulkcipher.so for sm_100a.
 *
 * bc_fill_splitmix64_dev is the device counterpart of the reference harness's
 * input generator (`srand(1337); msg[i] = rand() % 255`, aes-modes/test.c:54-56,
 * :354): BASELINE config 5 (256 GiB) cannot be generated on the host and copied,
 * so the buffer is produced in HBM by a counter-based generator that the host can
 * replay for any window (SURVEY §7 "Verification at 256 GiB"): 64-bit word i of
 * the LOGICAL buffer is mix64(seed + i), mix64 being the splitmix64 output
 * function.  Pure HBM-write work: one 16-byte store per thread per trip,
 * coalesced, grid = a multiple of the SM count.
 */
#include <cuda_runtime.h>
#include <stdint.h>

#include "bulkcipher.h"
#include "bc_internal.h"

namespace {

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x)
{
    uint64_t z = x + 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}

__global__ void __launch_bounds__(256)
fill_splitmix64_kernel(uint4 *out, uint64_t n16, uint64_t seed_plus_first, unsigned char *tail_out, uint32_t tail)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
        const uint64_t a = mix64(seed_plus_first + 2 * i), b = mix64(seed_plus_first + 2 * i + 1);
        __stcs(out + i, make_uint4((uint32_t)a, (uint32_t)(a >> 32), (uint32_t)b, (uint32_t)(b >> 32)));
    }
    if (blockIdx.x == 0 && threadIdx.x < tail) {
        /* last partial 16 bytes: little-endian bytes of the two words that cover them */
        const uint64_t w = mix64(seed_plus_first + 2 * n16 + (threadIdx.x >> 3));
        tail_out[threadIdx.x] = (unsigned char)(w >> (8 * (threadIdx.x & 7u)));
    }
}

}  // namespace

extern "C" int bc_fill_splitmix64_dev(void *d_out, size_t nbytes, uint64_t seed, uint64_t first_word, void *stream)
{
    if (nbytes == 0) return BC_OK;
    if (!d_out || ((uintptr_t)d_out & 15u)) {
        bc_set_error("bc_fill_splitmix64_dev: d_out must be non-NULL and 16-byte aligned");
        return BC_ERR_BAD_ARG;
    }
    const int sms = bc_cached_sm_count();
    if (sms <= 0) return BC_ERR_CUDA;
    const uint64_t n16 = nbytes / 16;
    uint64_t want = (n16 + 255) / 256;
    if (want < 1) want = 1;
    const uint64_t cap = (uint64_t)sms * 8;
    const unsigned grid = (unsigned)(want < cap ? want : cap);
    fill_splitmix64_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
        (uint4 *)d_out, n16, seed + first_word, (unsigned char *)d_out + n16 * 16, (uint32_t)(nbytes % 16));
    BC_CUDA(cudaGetLastError(), "fill_splitmix64_kernel launch");
    return BC_OK;
}
