/* bc_internal.h — shared between the CUDA translation units and the C host
 * layer of libbulkcipher.so.  Not installed; the public surface is include/. */
#ifndef BC_INTERNAL_H
#define BC_INTERNAL_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* printf-style, stores into the calling thread's error slot (bc_last_error). */
void bc_set_error(const char *fmt, ...);
/* Records "<what>: <cudaGetErrorString>" and returns BC_ERR_CUDA. */
int  bc_cuda_fail(int cuda_err, const char *what);
/* Number of SMs of the current device (cached per device). */
int  bc_cached_sm_count(void);
/* Grow-only per-device scratch buffer in HBM, owned by the library; one buffer per `key`.
 * For the host pipeline's own fixed keys only (called under the device lock). */
int  bc_scratch_get(void *key, size_t bytes, void **d_ptr);
/* Stream-ordered scratch from a library-owned memory pool: valid for the work enqueued on
 * `stream` between the two calls.  Any number of streams / host threads, no device sync. */
int  bc_stream_scratch_alloc(void *stream, size_t bytes, void **d_ptr);
int  bc_stream_scratch_free(void *stream, void *d_ptr);

#define BC_CUDA(call, what)                                        \
    do {                                                           \
        cudaError_t e__ = (call);                                  \
        if (e__ != cudaSuccess) return bc_cuda_fail((int)e__, what); \
    } while (0)

#ifdef __cplusplus
}
#endif
#endif
