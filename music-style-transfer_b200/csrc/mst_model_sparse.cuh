// mst_model_sparse.cuh — the pitched input grid is a piano roll: 0.1-0.6 % of its cells hold a note.
// Every layer that consumes the raw grid (beats_conv model.py:46-52, PitchedRhythmEncoder.channels_linear :318,
// MelodyEncoder.channels_linear :236) is therefore evaluated as "value at an empty cell" + corrections at the
// occupied cells.  The occupied cells are listed once per training step in CSR form, grouped by beat row
// (n, c, b, t) and ordered by cell index, so every consumer is deterministic and never re-reads the dense grid.
// The result is exact for ANY grid (a dense grid just makes long lists); nothing is approximated.
#pragma once
#include <cuda_bf16.h>
#include "mst_model_kernels2.cuh"

namespace mstm {

struct __align__(8) CellRec {
    uint32_t cell;     // f * 56 + note inside the beat row
    float p[5];        // duration, velocity, flat, natural, sharp
};
static_assert(sizeof(CellRec) == 24, "CellRec layout");

// ------------------------------------------------------------------------------------------------ CSR build
// warp per beat row: count (pass 0) or fill (pass 1) the occupied cells of P[row][560][5]
template <int FILL>
__global__ void __launch_bounds__(256)
k_cells_scan(const float *__restrict__ P, long long rows, uint32_t *__restrict__ cnt, const uint32_t *__restrict__ off,
             CellRec *__restrict__ recs) {
    const int lane = threadIdx.x & 31;
    const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long n_warps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long row = warp0; row < rows; row += n_warps) {
        const float *pr = P + row * 2800;
        uint32_t base = FILL ? off[row] : 0u, total = 0;
        for (int c0 = 0; c0 < 560; c0 += 32) {
            const int cell = c0 + lane;
            float p[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
            if (cell < 560) {
#pragma unroll
                for (int e = 0; e < 5; e++) p[e] = pr[cell * 5 + e];
            }
            const bool hit = (p[0] != 0.f) | (p[1] != 0.f) | (p[2] != 0.f) | (p[3] != 0.f) | (p[4] != 0.f);
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (FILL && hit) {
                CellRec r;
                r.cell = (uint32_t)cell;
#pragma unroll
                for (int e = 0; e < 5; e++) r.p[e] = p[e];
                recs[base + total + __popc(m & ((1u << lane) - 1u))] = r;
            }
            total += __popc(m);
        }
        if (!FILL && lane == 0) cnt[row] = total;
    }
}

// exclusive scan of cnt[rows] -> off[rows + 1] (single CTA; rows is a few hundred thousand at most)
__global__ void __launch_bounds__(1024, 1)
k_cells_offsets(const uint32_t *__restrict__ cnt, uint32_t *__restrict__ off, long long rows) {
    __shared__ unsigned warp_sums[33];
    __shared__ unsigned carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (long long base = 0; base < rows; base += 1024) {
        const long long i = base + threadIdx.x;
        const unsigned v = i < rows ? cnt[i] : 0u;
        unsigned total;
        const unsigned ex = block_exclusive_scan(v, warp_sums, &total);
        const unsigned carry = carry_s;
        if (i < rows) off[i] = carry + ex;
        __syncthreads();
        if (threadIdx.x == 0) carry_s = carry + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) off[rows] = carry_s;
}

// ------------------------------------------------------------------------------------------------ beats_conv
// conv1d(50 -> 57, kernel 14, stride 7, padding 4) over the 56 notes, 8 output positions ("octaves"):
// out[row][oct][o] = leaky(bias[o] + sum over occupied cells (f, note) with tap = note - 7 oct + 4 in [0, 14) of
//                    sum_feat WT[f * 70 + tap * 5 + feat][o] * p[feat]).      WT is the (700, 57) transposed weight.
// warp per beat row, accumulators in shared memory, no atomics.
__global__ void __launch_bounds__(256)
k_conv_sparse_fwd(const uint32_t *__restrict__ off, const CellRec *__restrict__ recs, const float *__restrict__ WT,
                  const float *__restrict__ bias, float *__restrict__ out, long long rows, __nv_bfloat16 *__restrict__ out16) {
    __shared__ float acc_s[8][456];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float *acc = acc_s[warp];
    const long long warp0 = (long long)blockIdx.x * 8 + warp, n_warps = (long long)gridDim.x * 8;
    for (long long row = warp0; row < rows; row += n_warps) {
        const uint32_t lo = off[row], hi = off[row + 1];
        float *dst = out + row * 456;
        __nv_bfloat16 *dst16 = out16 ? out16 + row * 456 : nullptr;      // bf16 dense mode: the TMA-fed GEMM reads this copy
        if (lo == hi) {
            for (int i = lane; i < 456; i += 32) {
                const float v = leaky(bias[i % 57]);
                dst[i] = v;
                if (dst16) dst16[i] = __float2bfloat16_rn(v);
            }
            continue;
        }
        __syncwarp();
        for (int i = lane; i < 456; i += 32) acc[i] = bias[i % 57];
        __syncwarp();
        for (uint32_t k = lo; k < hi; k++) {
            const CellRec r = recs[k];
            const int f = (int)r.cell / 56, note = (int)r.cell - f * 56;
            // windows: oct with 0 <= note - 7 oct + 4 < 14  <=>  oct in [ceil((note - 9) / 7), floor((note + 4) / 7)]
            const int o_hi = min(7, (note + 4) / 7), o_lo = max(0, (note - 9 + 6) / 7);
            for (int oct = o_lo; oct <= o_hi; oct++) {
                const int tap = note - 7 * oct + 4;
                const float *w = WT + (long long)(f * 70 + tap * 5) * 57;
                for (int o = lane; o < 57; o += 32) {
                    float v = acc[oct * 57 + o];
#pragma unroll
                    for (int e = 0; e < 5; e++) v = fmaf(w[e * 57 + o], r.p[e], v);
                    acc[oct * 57 + o] = v;
                }
            }
            __syncwarp();
        }
        for (int i = lane; i < 456; i += 32) {
            const float v = leaky(acc[i]);
            dst[i] = v;
            if (dst16) dst16[i] = __float2bfloat16_rn(v);
        }
    }
}

// weight gradient: gT[f * 70 + tap * 5 + feat][o] += dpre[row][oct][o] * p[feat] over the occupied cells
__global__ void __launch_bounds__(256)
k_conv_sparse_wgrad(const uint32_t *__restrict__ off, const CellRec *__restrict__ recs, const float *__restrict__ dpre,
                    float *__restrict__ gT, long long rows) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long warp0 = (long long)blockIdx.x * 8 + warp, n_warps = (long long)gridDim.x * 8;
    for (long long row = warp0; row < rows; row += n_warps) {
        const uint32_t lo = off[row], hi = off[row + 1];
        const float *d = dpre + row * 456;
        for (uint32_t k = lo; k < hi; k++) {
            const CellRec r = recs[k];
            const int f = (int)r.cell / 56, note = (int)r.cell - f * 56;
            const int o_hi = min(7, (note + 4) / 7), o_lo = max(0, (note - 9 + 6) / 7);
            for (int oct = o_lo; oct <= o_hi; oct++) {
                const int tap = note - 7 * oct + 4;
                float *g = gT + (long long)(f * 70 + tap * 5) * 57;
                for (int o = lane; o < 57; o += 32) {
                    const float dv = d[oct * 57 + o];
                    if (dv == 0.f) continue;
#pragma unroll
                    for (int e = 0; e < 5; e++) if (r.p[e] != 0.f) atomicAdd(g + e * 57 + o, dv * r.p[e]);
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ channels_linear (280 -> 16)
// PitchedRhythmEncoder (model.py:318,358): a3[row, f][j] = leaky(b[j] + sum_{note, feat} W[j][note * 5 + feat] P[row, f, note, feat]).
// WT is the (280, 16) transposed weight.  warp per beat row = 10 fraction rows x 16 outputs.
__global__ void __launch_bounds__(256)
k_chlin_sparse_fwd(const uint32_t *__restrict__ off, const CellRec *__restrict__ recs, const float *__restrict__ WT,
                   const float *__restrict__ bias, float *__restrict__ out, long long rows) {
    __shared__ float acc_s[8][160];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float *acc = acc_s[warp];
    const long long warp0 = (long long)blockIdx.x * 8 + warp, n_warps = (long long)gridDim.x * 8;
    const int j = lane & 15, half = lane >> 4;
    for (long long row = warp0; row < rows; row += n_warps) {
        const uint32_t lo = off[row], hi = off[row + 1];
        float *dst = out + row * 160;
        if (lo == hi) {
            for (int i = lane; i < 160; i += 32) dst[i] = leaky(bias[i & 15]);
            continue;
        }
        __syncwarp();
        for (int i = lane; i < 160; i += 32) acc[i] = bias[i & 15];
        __syncwarp();
        // two cells per step (one per half warp); cells are ordered by index, so equal fractions may collide: atomics
        for (uint32_t k = lo + half; k < hi; k += 2) {
            const CellRec r = recs[k];
            const int f = (int)r.cell / 56, note = (int)r.cell - f * 56;
            float v = 0.f;
#pragma unroll
            for (int e = 0; e < 5; e++) v = fmaf(WT[(note * 5 + e) * 16 + j], r.p[e], v);
            atomicAdd(&acc[f * 16 + j], v);
        }
        __syncwarp();
        for (int i = lane; i < 160; i += 32) dst[i] = leaky(acc[i]);
    }
}

// weight gradient: gW[j][note * 5 + feat] += dpre[row, f][j] * p[feat]
__global__ void __launch_bounds__(256)
k_chlin_sparse_wgrad(const uint32_t *__restrict__ off, const CellRec *__restrict__ recs, const float *__restrict__ dpre,
                     float *__restrict__ gW, long long rows) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long warp0 = (long long)blockIdx.x * 8 + warp, n_warps = (long long)gridDim.x * 8;
    const int j = lane & 15, half = lane >> 4;
    for (long long row = warp0; row < rows; row += n_warps) {
        const uint32_t lo = off[row], hi = off[row + 1];
        for (uint32_t k = lo + half; k < hi; k += 2) {
            const CellRec r = recs[k];
            const int f = (int)r.cell / 56, note = (int)r.cell - f * 56;
            const float dv = dpre[(row * 10 + f) * 16 + j];
            if (dv == 0.f) continue;
#pragma unroll
            for (int e = 0; e < 5; e++) if (r.p[e] != 0.f) atomicAdd(gW + j * 280 + note * 5 + e, dv * r.p[e]);
        }
    }
}

}  // namespace mstm
