/*
 * bc_tma.cuh — TMA plumbing shared by the batched kernels (RC4 streams, CBC-encrypt
 * messages): every lane of a warp walks its own row of a [rows][row_bytes] byte
 * matrix, which as plain per-lane 16-byte accesses costs one data-pipe wavefront
 * per lane.  Instead a warp moves boxes of 32 rows x 64 bytes with one bulk-tensor
 * copy each way (cp.async.bulk.tensor.2d, SWIZZLE_64B so a lane's LDS.128/STS.128
 * on its row are conflict free); rows/bytes outside the matrix are clipped by the
 * TMA unit.  sm_100a only.
 */
#ifndef BC_TMA_CUH
#define BC_TMA_CUH

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "bulkcipher.h"
#include "bc_internal.h"

namespace {

/* Raise the kernel's dynamic shared-memory limit, once per device and kernel
 * (cudaFuncSetAttribute costs microseconds, which matters for small buffers). */
template <auto Kernel>
int set_smem(int bytes)
{
    static int done_bytes[64];
    int dev = 0;
    BC_CUDA(cudaGetDevice(&dev), "cudaGetDevice");
    if (dev >= 0 && dev < 64 && __atomic_load_n(&done_bytes[dev], __ATOMIC_ACQUIRE) >= bytes) return BC_OK;
    BC_CUDA(cudaFuncSetAttribute(Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes),
            "cudaFuncSetAttribute(MaxDynamicSharedMemorySize)");
    if (dev >= 0 && dev < 64) __atomic_store_n(&done_bytes[dev], bytes, __ATOMIC_RELEASE);
    return BC_OK;
}

constexpr int kSlab     = 64;            /* bytes of one row per box */
constexpr int kBoxBytes = 32 * kSlab;    /* 2 KB per box */

/* byte offset of 16-byte chunk c (0..3) of row `lane` inside a SWIZZLE_64B box */
__device__ __forceinline__ uint32_t box_chunk_off(uint32_t lane, uint32_t c)
{
    return lane * kSlab + ((c ^ ((lane >> 1) & 3u)) << 4);
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}

__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}

/* Wait for the phase with the given parity.  try_wait suspends the thread for a hardware time
 * slice per attempt; a box that never arrives (a bad tensor map, say) traps after a few seconds
 * instead of hanging the GPU. */
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    for (uint32_t spins = 0;; spins++) {
        uint32_t done;
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (done) return;
        if (spins > (1u << 24)) asm volatile("trap;");
    }
}

__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *tm, uint32_t c0, uint32_t c1, uint32_t bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 :: "r"(dst), "l"(tm), "r"(c0), "r"(c1), "r"(bar) : "memory");
}

__device__ __forceinline__ void tma_store_2d(const CUtensorMap *tm, uint32_t c0, uint32_t c1, uint32_t src)
{
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                 :: "l"(tm), "r"(c0), "r"(c1), "r"(src) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

/* global[box] ^= shared[box]: the XOR happens in the memory system (L2), nothing is loaded */
__device__ __forceinline__ void tma_reduce_xor_2d(const CUtensorMap *tm, uint32_t c0, uint32_t c1, uint32_t src)
{
    asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.xor.tile.bulk_group [%0, {%1, %2}], [%3];"
                 :: "l"(tm), "r"(c0), "r"(c1), "r"(src) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

__device__ __forceinline__ void tma_store_wait_read()
{
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

__device__ __forceinline__ void tma_store_wait_all()
{
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

__device__ __forceinline__ uint4 lds128(uint32_t a)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
    return v;
}

__device__ __forceinline__ void sts128(uint32_t a, uint4 v)
{
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" :: "r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int make_row_tensor_map(CUtensorMap *tm, const void *ptr, uint64_t n_rows, uint64_t row_bytes, bool as_u32 = false)
{
    static EncodeTiledFn encode = nullptr;
    if (!encode) {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        BC_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q), "cudaGetDriverEntryPoint");
        if (!fn || q != cudaDriverEntryPointSuccess) { bc_set_error("cuTensorMapEncodeTiled is not available"); return BC_ERR_CUDA; }
        encode = (EncodeTiledFn)fn;
    }
    /* as_u32: same bytes described as 32-bit elements (bitwise TMA reductions need a 32-bit type) */
    const cuuint64_t dims[2] = { (cuuint64_t)(as_u32 ? row_bytes / 4 : row_bytes), (cuuint64_t)n_rows };
    const cuuint64_t strides[1] = { (cuuint64_t)row_bytes };
    const cuuint32_t box[2] = { (cuuint32_t)(as_u32 ? kSlab / 4 : kSlab), 32u };
    const cuuint32_t estr[2] = { 1u, 1u };
    CUresult r = encode(tm, as_u32 ? CU_TENSOR_MAP_DATA_TYPE_UINT32 : CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(ptr), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { bc_set_error("cuTensorMapEncodeTiled failed (CUresult %d)", (int)r); return BC_ERR_CUDA; }
    return BC_OK;
}

}  // namespace

#endif
