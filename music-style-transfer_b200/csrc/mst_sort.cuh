// mst_sort.cuh — stable LSD radix sort of (key, value) pairs by one 1024-thread CTA, in global memory.
// Used for merge_tracks' stable sort by tick (encode) and for create_midi's stable sort by time (decode).
#pragma once
#include "mst_device.cuh"

constexpr int kSortThreads = 1024;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortBatch = 8;       // strips whose keys a warp fetches together

// lanes of `act` whose 8-bit digit equals mine.  Eight ballots instead of match.any: the hardware match instruction
// serialises over the distinct values in the warp (32 for the low tick bytes), which made it the top stall of both
// sort kernels (ncu: the instruction consuming its result held 40 % of k_sort's samples).  All 32 lanes must call.
__device__ __forceinline__ unsigned match_digit8(unsigned act, uint32_t d) {
    unsigned peers = act;
#pragma unroll
    for (int b = 0; b < 8; b++) {
        const unsigned m = __ballot_sync(0xffffffffu, (d >> b) & 1u);
        peers &= ((d >> b) & 1u) ? m : ~m;
    }
    return peers;
}

// One 8-bit pass.  Every warp owns a contiguous segment of the input, so the relative order of equal
// digits is preserved: warp-private digit offsets come from a per-digit scan over the warps, ranks
// inside a 32-element strip from match_digit8.  whist: [kSortWarps][256] words of shared memory.
__device__ inline void radix_pass(const uint32_t *kin, const uint32_t *vin, uint32_t *kout, uint32_t *vout, uint32_t n,
                           int shift, uint32_t (*whist)[256], uint32_t *scan_tmp, int *skip_flag) {
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (unsigned i = tid; i < kSortWarps * 256; i += kSortThreads) (&whist[0][0])[i] = 0;
    if (tid == 0) *skip_flag = 0;
    __syncthreads();
    const uint32_t per = ((n + kSortThreads - 1) / kSortThreads) * 32;
    const uint32_t lo = min(n, warp * per), hi = min(n, lo + per);
    // strips are handled in batches of kSortBatch: the batch's keys are fetched first (independent loads in flight), so
    // a warp pays the L2 / HBM latency once per batch instead of once per strip
    for (uint32_t b0 = lo; b0 < hi; b0 += 32 * kSortBatch) {
        uint32_t kk[kSortBatch];
#pragma unroll
        for (int u = 0; u < kSortBatch; u++) { const uint32_t i = b0 + u * 32 + lane; kk[u] = i < hi ? kin[i] : 0u; }
#pragma unroll
        for (int u = 0; u < kSortBatch; u++) {
            const uint32_t i = b0 + u * 32 + lane;
            if (b0 + u * 32 >= hi) break;
            const bool valid = i < hi;
            const unsigned act = __ballot_sync(0xffffffffu, valid);
            const uint32_t d = (kk[u] >> shift) & 255u;
            const unsigned peers = match_digit8(act, d);
            if (valid && (peers & lanemask_lt()) == 0) whist[warp][d] += __popc(peers);
            __syncwarp();
        }
    }
    __syncthreads();
    uint32_t total = 0;
    if (tid < 256) {
        for (int w = 0; w < kSortWarps; w++) {
            uint32_t t = whist[w][tid];
            whist[w][tid] = total;
            total += t;
        }
        if (total == n) *skip_flag = 1;
    }
    uint32_t grand;
    uint32_t base = block_exclusive_scan(tid < 256 ? total : 0u, scan_tmp, &grand);
    if (tid < 256)
        for (int w = 0; w < kSortWarps; w++) whist[w][tid] += base;
    __syncthreads();
    if (*skip_flag) {   // every key has the same digit: the pass is the identity
        for (uint32_t i = tid; i < n; i += kSortThreads) { kout[i] = kin[i]; vout[i] = vin[i]; }
        __syncthreads();
        return;
    }
    for (uint32_t b0 = lo; b0 < hi; b0 += 32 * kSortBatch) {
        uint32_t kk[kSortBatch], vv[kSortBatch];
#pragma unroll
        for (int u = 0; u < kSortBatch; u++) {
            const uint32_t i = b0 + u * 32 + lane;
            kk[u] = i < hi ? kin[i] : 0u;
            vv[u] = i < hi ? vin[i] : 0u;
        }
#pragma unroll
        for (int u = 0; u < kSortBatch; u++) {
            const uint32_t i = b0 + u * 32 + lane;
            if (b0 + u * 32 >= hi) break;
            const bool valid = i < hi;
            const unsigned act = __ballot_sync(0xffffffffu, valid);
            const uint32_t d = (kk[u] >> shift) & 255u;
            const unsigned peers = match_digit8(act, d);
            if (valid) {
                const uint32_t dst = whist[warp][d] + __popc(peers & lanemask_lt());
                kout[dst] = kk[u];
                vout[dst] = vv[u];
            }
            __syncwarp();
            if (valid && (peers & lanemask_lt()) == 0) whist[warp][d] += __popc(peers);
            __syncwarp();
        }
    }
    __syncthreads();
}

