// mst_device.cuh — shared device/host helpers for the sm_100a kernels of libmst_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/mst_b200.h"

// ---------------------------------------------------------------- host-side error plumbing
void mst_set_error(const char *fmt, ...);

#define MST_CUDA_CHECK(expr)                                                             \
    do {                                                                                 \
        cudaError_t _e = (expr);                                                         \
        if (_e != cudaSuccess) {                                                         \
            mst_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return MST_E_CUDA;                                                           \
        }                                                                                \
    } while (0)

#define MST_LAUNCH_CHECK(name)                                                           \
    do {                                                                                 \
        cudaError_t _e = cudaGetLastError();                                             \
        if (_e != cudaSuccess) {                                                         \
            mst_set_error("launch of %s failed: %s", name, cudaGetErrorString(_e));      \
            return MST_E_CUDA;                                                           \
        }                                                                                \
    } while (0)

// mst_encode_batch_device with a hook that enqueues the producers of d_events / d_songs on the front-end stream (internal)
int mst_encode_batch_device_front(const MstEvent *d_events, uint64_t n_events_total, const MstSongIn *d_songs, uint32_t n_songs,
                                  const MstFilter *h_filter, MstSongInfo *d_info, float *d_arena, uint64_t arena_floats,
                                  void *d_workspace, size_t workspace_bytes, void *stream,
                                  int (*front_hook)(void *ctx, cudaStream_t front), void *hook_ctx);

static inline size_t mst_align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---------------------------------------------------------------- device helpers
#ifdef __CUDACC__
__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ unsigned lanemask_lt() {
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}
__device__ __forceinline__ unsigned lanemask_gt() {
    unsigned m;
    asm("mov.u32 %0, %%lanemask_gt;" : "=r"(m));
    return m;
}

__device__ __forceinline__ int warp_reduce_min(int v) {
    for (int o = 16; o; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ int warp_reduce_max(int v) {
    for (int o = 16; o; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ unsigned warp_inclusive_scan(unsigned v) {
    unsigned lane = lane_id();
    for (int o = 1; o < 32; o <<= 1) {
        unsigned t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= (unsigned)o) v += t;
    }
    return v;
}

// Block-wide exclusive scan of one value per thread (blockDim.x <= 1024, multiple of 32).
// `warp_sums` is 33 words of shared memory.  Returns the exclusive prefix; *total gets the block sum.
__device__ __forceinline__ unsigned block_exclusive_scan(unsigned v, unsigned *warp_sums, unsigned *total) {
    unsigned lane = lane_id(), warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    unsigned inc = warp_inclusive_scan(v);
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        unsigned s = lane < nwarps ? warp_sums[lane] : 0;
        unsigned si = warp_inclusive_scan(s);
        warp_sums[lane] = si - s;
        if (lane == 31) warp_sums[32] = si;
    }
    __syncthreads();
    unsigned res = warp_sums[warp] + inc - v;
    *total = warp_sums[32];
    __syncthreads();
    return res;
}

// ---- bulk async copy (TMA 1-D) shared -> global, bulk-group completion
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void bulk_store_s2g(void *gdst, const void *ssrc, uint32_t bytes) {
    uint32_t s = (uint32_t)__cvta_generic_to_shared(ssrc);
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 :: "l"(gdst), "r"(s), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait() {
    asm volatile("cp.async.bulk.wait_group %0;" :: "n"(N) : "memory");
}
#endif  // __CUDACC__
