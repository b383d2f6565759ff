// mst_model_me.cuh — MelodyEncoder tail (model.py:284-297), forward and backward, restructured around two facts:
//   (1) the octave / degree vectors are per BEAT (n, c, b, t): the outer add a[q] = leaky(oct[o][q] + deg[d][q]) does not
//       depend on the beat fraction f, so for an EMPTY cell (P = 0) the whole 15 -> 8 linear gives the same vector
//       xe[note] for all 10 fractions of the beat;
//   (2) the grid is a piano roll: occupied cells are rare and listed in CSR form (mst_model_sparse.cuh).
// So every quantity is "10 x (per-note value of the empty cell)" + corrections at the occupied cells — exact for any
// grid.  Work drops from one 15 -> 8 linear per (fraction, note, channel) cell to one per (note, channel) + the list.
//   in15 = [a(8); bf(7)],  bf = leaky(bc + Wc.p5)  (= bfe = leaky(bc) when p5 = 0),  x = leaky(bl + Wl.in15)
//   melody[n,b,t,f,note,:] = sum_c x_c w_c / sum_c w_c,  w_c = sqrt(1 + sum over the channel of x^2)   (combine, :796)
#pragma once
#include "mst_model_loss.cuh"
#include "mst_model_sparse.cuh"

namespace mstm {

constexpr int kMeWc = 0, kMeBc = 35, kMeWl = 42, kMeBl = 162, kMeConst = 176;
__constant__ float c_me[kMeConst];      // channels_linear (7,5) + bias (7), linear (8,15) + bias (8)

struct MeArgs2 {
    const float *od;                    // (N, C, B, T, 120) post-activation: 64 octave values then 56 degree values
    const uint32_t *off; const CellRec *recs;   // occupied cells per beat row (n, c, b, t)
    float *ss;                          // (N, C) sum of squares (zeroed before k_me_sumsq2)
    float *melody;                      // (N, B, T, 10, 56, 8)
    const float *dmelody;               // backward: gradient w.r.t. melody
    float *dsum;                        // backward: (N, B, T, 56, 8) sum over the 10 fractions of dmelody
    float *stats;                       // backward: (N, C + 1): A_c = sum dmel.x_c, then B = sum dmel.melody (zeroed)
    float *dod;                         // backward: (N, C, B, T, 120) pre-activation gradient, written
    float *gWc, *gbc, *gWl, *gbl;       // parameter gradients (atomic accumulate)
    int N, C, B, T;
};

struct MeConsts { float bfe[7], kb[8]; };   // leaky(bc) and bl + Wl[:, 8:].bfe
__device__ __forceinline__ MeConsts me_consts() {
    MeConsts k;
#pragma unroll
    for (int q = 0; q < 7; q++) k.bfe[q] = leaky(c_me[kMeBc + q]);
#pragma unroll
    for (int e = 0; e < 8; e++) {
        float v = c_me[kMeBl + e];
#pragma unroll
        for (int q = 0; q < 7; q++) v = fmaf(c_me[kMeWl + e * 15 + 8 + q], k.bfe[q], v);
        k.kb[e] = v;
    }
    return k;
}
// empty-cell value at one note: ae[8] and the pre-activation xe_pre[8]
__device__ __forceinline__ void me_note(const MeConsts &k, const float *od, int note, float *ae, float *xpre) {
    const int o = note / 7, d = note - o * 7;
#pragma unroll
    for (int q = 0; q < 8; q++) ae[q] = leaky(od[o * 8 + q] + od[64 + d * 8 + q]);
#pragma unroll
    for (int e = 0; e < 8; e++) {
        float v = k.kb[e];
#pragma unroll
        for (int q = 0; q < 8; q++) v = fmaf(c_me[kMeWl + e * 15 + q], ae[q], v);
        xpre[e] = v;
    }
}
// occupied cell: bf[7] and the pre-activation x_pre[8] from the empty-cell pre-activation
__device__ __forceinline__ void me_occupied(const MeConsts &k, const float *p5, const float *xe_pre, float *bf, float *xpre) {
#pragma unroll
    for (int q = 0; q < 7; q++) {
        float v = c_me[kMeBc + q];
#pragma unroll
        for (int p = 0; p < 5; p++) v = fmaf(c_me[kMeWc + q * 5 + p], p5[p], v);
        bf[q] = leaky(v);
    }
#pragma unroll
    for (int e = 0; e < 8; e++) {
        float v = xe_pre[e];
#pragma unroll
        for (int q = 0; q < 7; q++) v = fmaf(c_me[kMeWl + e * 15 + 8 + q], bf[q] - k.bfe[q], v);
        xpre[e] = v;
    }
}

__device__ __forceinline__ float warp_sum(float v) {
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ------------------------------------------------------------------------------------------------ forward
// warp per beat row (n, c, b, t): ss[n, c] += 10 sum_note |xe|^2 + sum_occupied (|x|^2 - |xe|^2)
__global__ void __launch_bounds__(256)
k_me_sumsq2(MeArgs2 a) {
    __shared__ float od_s[8][120];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const MeConsts k = me_consts();
    const long long BT = (long long)a.B * a.T, rows = (long long)a.N * a.C * BT;
    for (long long row = (long long)blockIdx.x * 8 + warp; row < rows; row += (long long)gridDim.x * 8) {
        __syncwarp();
        for (int i = lane; i < 120; i += 32) od_s[warp][i] = a.od[row * 120 + i];
        __syncwarp();
        float s = 0.f;
        for (int note = lane; note < 56; note += 32) {
            float ae[8], xp[8];
            me_note(k, od_s[warp], note, ae, xp);
#pragma unroll
            for (int e = 0; e < 8; e++) { const float x = leaky(xp[e]); s = fmaf(10.f * x, x, s); }
        }
        for (uint32_t i = a.off[row] + lane; i < a.off[row + 1]; i += 32) {
            const CellRec r = a.recs[i];
            const int note = (int)(r.cell % 56);
            float ae[8], xp[8], bf[7], xt[8];
            me_note(k, od_s[warp], note, ae, xp);
            me_occupied(k, r.p, xp, bf, xt);
#pragma unroll
            for (int e = 0; e < 8; e++) { const float x = leaky(xt[e]), xe = leaky(xp[e]); s += x * x - xe * xe; }
        }
        s = warp_sum(s);
        if (lane == 0) atomicAdd(a.ss + row / BT, s);
    }
}

// warp per (n, b, t): melody tile (10, 56, 8) = (sum_c w_c xe_c) / wsum for every fraction, then the occupied cells
__global__ void __launch_bounds__(256)
k_me_fwd2(MeArgs2 a) {
    __shared__ float od_s[8][120];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const MeConsts k = me_consts();
    const long long BT = (long long)a.B * a.T, units = (long long)a.N * BT;
    for (long long u = (long long)blockIdx.x * 8 + warp; u < units; u += (long long)gridDim.x * 8) {
        const int n = (int)(u / BT);
        const long long bt = u - (long long)n * BT;
        float wsum = 0.f;
        for (int c = 0; c < a.C; c++) wsum += sqrtf(1.f + a.ss[n * a.C + c]);
        float base[2][8];
#pragma unroll
        for (int r = 0; r < 2; r++)
#pragma unroll
            for (int e = 0; e < 8; e++) base[r][e] = 0.f;
        for (int c = 0; c < a.C; c++) {
            const long long row = ((long long)n * a.C + c) * BT + bt;
            const float wc = sqrtf(1.f + a.ss[n * a.C + c]);
            __syncwarp();
            for (int i = lane; i < 120; i += 32) od_s[warp][i] = a.od[row * 120 + i];
            __syncwarp();
#pragma unroll
            for (int r = 0; r < 2; r++) {
                const int note = lane + 32 * r;
                if (note < 56) {
                    float ae[8], xp[8];
                    me_note(k, od_s[warp], note, ae, xp);
#pragma unroll
                    for (int e = 0; e < 8; e++) base[r][e] = fmaf(leaky(xp[e]), wc, base[r][e]);
                }
            }
        }
        float *tile = a.melody + u * 4480;
#pragma unroll
        for (int r = 0; r < 2; r++) {
            const int note = lane + 32 * r;
            if (note < 56) {
                const float4 v0 = make_float4(base[r][0] / wsum, base[r][1] / wsum, base[r][2] / wsum, base[r][3] / wsum);
                const float4 v1 = make_float4(base[r][4] / wsum, base[r][5] / wsum, base[r][6] / wsum, base[r][7] / wsum);
                for (int f = 0; f < 10; f++) {
                    float4 *dst = reinterpret_cast<float4 *>(tile + (f * 56 + note) * 8);
                    dst[0] = v0; dst[1] = v1;
                }
            }
        }
        // corrections at the occupied cells, channel by channel (cells of one channel are distinct)
        for (int c = 0; c < a.C; c++) {
            const long long row = ((long long)n * a.C + c) * BT + bt;
            const uint32_t lo = a.off[row], hi = a.off[row + 1];
            if (lo == hi) continue;
            const float wc = sqrtf(1.f + a.ss[n * a.C + c]) / wsum;
            __syncwarp();
            for (int i = lane; i < 120; i += 32) od_s[warp][i] = a.od[row * 120 + i];
            __syncwarp();
            for (uint32_t i = lo + lane; i < hi; i += 32) {
                const CellRec r = a.recs[i];
                const int note = (int)(r.cell % 56);
                float ae[8], xp[8], bf[7], xt[8];
                me_note(k, od_s[warp], note, ae, xp);
                me_occupied(k, r.p, xp, bf, xt);
                float *m = tile + r.cell * 8;
#pragma unroll
                for (int e = 0; e < 8; e++) m[e] += wc * (leaky(xt[e]) - leaky(xp[e]));
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ backward
// combine backward (see k_combine_bwd): dx_c = dmel * w_c / S + (A_c - Bs) / (S w_c) * x_c,
//   A_c = sum over the channel of dmel . x_c,  Bs = sum over the sample of dmel . melody.
// warp per (n, b, t): dsum[note] = sum_f dmel[f][note]; Bs partial; A_c partial = sum_note dsum.xe_c + occupied corrections
__global__ void __launch_bounds__(256)
k_me_bwd_stats2(MeArgs2 a) {
    __shared__ float od_s[8][120];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const MeConsts k = me_consts();
    const long long BT = (long long)a.B * a.T, units = (long long)a.N * BT;
    for (long long u = (long long)blockIdx.x * 8 + warp; u < units; u += (long long)gridDim.x * 8) {
        const int n = (int)(u / BT);
        const long long bt = u - (long long)n * BT;
        const float *dm = a.dmelody + u * 4480, *mel = a.melody + u * 4480;
        float ds[2][8], bs = 0.f;
#pragma unroll
        for (int r = 0; r < 2; r++) {
#pragma unroll
            for (int e = 0; e < 8; e++) ds[r][e] = 0.f;
            const int note = lane + 32 * r;
            if (note < 56) {
                for (int f = 0; f < 10; f++) {
                    const float4 *dp = reinterpret_cast<const float4 *>(dm + (f * 56 + note) * 8);
                    const float4 *mp = reinterpret_cast<const float4 *>(mel + (f * 56 + note) * 8);
                    const float4 d0 = dp[0], d1 = dp[1], m0 = mp[0], m1 = mp[1];
                    ds[r][0] += d0.x; ds[r][1] += d0.y; ds[r][2] += d0.z; ds[r][3] += d0.w;
                    ds[r][4] += d1.x; ds[r][5] += d1.y; ds[r][6] += d1.z; ds[r][7] += d1.w;
                    bs += d0.x * m0.x + d0.y * m0.y + d0.z * m0.z + d0.w * m0.w + d1.x * m1.x + d1.y * m1.y + d1.z * m1.z + d1.w * m1.w;
                }
                float4 *out = reinterpret_cast<float4 *>(a.dsum + (u * 56 + note) * 8);
                out[0] = make_float4(ds[r][0], ds[r][1], ds[r][2], ds[r][3]);
                out[1] = make_float4(ds[r][4], ds[r][5], ds[r][6], ds[r][7]);
            }
        }
        bs = warp_sum(bs);
        if (lane == 0) atomicAdd(a.stats + n * (a.C + 1) + a.C, bs);
        for (int c = 0; c < a.C; c++) {
            const long long row = ((long long)n * a.C + c) * BT + bt;
            __syncwarp();
            for (int i = lane; i < 120; i += 32) od_s[warp][i] = a.od[row * 120 + i];
            __syncwarp();
            float s = 0.f;
#pragma unroll
            for (int r = 0; r < 2; r++) {
                const int note = lane + 32 * r;
                if (note < 56) {
                    float ae[8], xp[8];
                    me_note(k, od_s[warp], note, ae, xp);
#pragma unroll
                    for (int e = 0; e < 8; e++) s = fmaf(ds[r][e], leaky(xp[e]), s);
                }
            }
            for (uint32_t i = a.off[row] + lane; i < a.off[row + 1]; i += 32) {
                const CellRec r = a.recs[i];
                const int note = (int)(r.cell % 56);
                float ae[8], xp[8], bf[7], xt[8];
                me_note(k, od_s[warp], note, ae, xp);
                me_occupied(k, r.p, xp, bf, xt);
                const float *d = dm + r.cell * 8;
#pragma unroll
                for (int e = 0; e < 8; e++) s = fmaf(d[e], leaky(xt[e]) - leaky(xp[e]), s);
            }
            s = warp_sum(s);
            if (lane == 0) atomicAdd(a.stats + n * (a.C + 1) + c, s);
        }
    }
}

struct MeBwdWarpSmem { float od[120], dxp[56 * 8], dxa[56 * 8]; };

// warp per beat row (n, c, b, t).  Per note (all 10 fractions treated as empty cells):
//   dxs = (k1 dsum + 10 k2 xe) leaky'(xe);  occupied cells replace their empty share by the true one.
//   dxp multiplies the empty input [ae; bfe], dxa = dxp + sum of the true cells' dx multiplies ae (shared by both).
// The weight gradients that only the occupied cells feed (channels_linear, linear.weight[:, 8:]) are k_me_bwd_cells'.
__global__ void __launch_bounds__(256)
k_me_bwd2(MeArgs2 a) {
    __shared__ MeBwdWarpSmem sm[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    MeBwdWarpSmem &s = sm[warp];
    const MeConsts k = me_consts();
    const long long BT = (long long)a.B * a.T, rows = (long long)a.N * a.C * BT;
    float gWa[64], Sp[8], Sa[8];       // linear.weight[:, :8] gradient, sum of dxp, sum of dxa — reduced once at the end
#pragma unroll
    for (int i = 0; i < 64; i++) gWa[i] = 0.f;
#pragma unroll
    for (int e = 0; e < 8; e++) { Sp[e] = 0.f; Sa[e] = 0.f; }
    for (long long row = (long long)blockIdx.x * 8 + warp; row < rows; row += (long long)gridDim.x * 8) {
        const long long nc = row / BT, bt = row - nc * BT;
        const int n = (int)(nc / a.C), c = (int)(nc - (long long)n * a.C);
        float wsum = 0.f;
        for (int j = 0; j < a.C; j++) wsum += sqrtf(1.f + a.ss[n * a.C + j]);
        const float wc = sqrtf(1.f + a.ss[n * a.C + c]);
        const float k1 = wc / wsum;
        const float k2 = (a.stats[n * (a.C + 1) + c] - a.stats[n * (a.C + 1) + a.C]) / (wsum * wc);
        const long long u = (long long)n * BT + bt;
        __syncwarp();
        for (int i = lane; i < 120; i += 32) s.od[i] = a.od[row * 120 + i];
        __syncwarp();
        // phase A: per-note empty-cell gradient
        for (int note = lane; note < 56; note += 32) {
            float ae[8], xp[8];
            me_note(k, s.od, note, ae, xp);
            const float4 *dp = reinterpret_cast<const float4 *>(a.dsum + (u * 56 + note) * 8);
            const float4 d0 = dp[0], d1 = dp[1];
            const float dsv[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
#pragma unroll
            for (int e = 0; e < 8; e++) {
                const float xe = leaky(xp[e]);
                const float v = (k1 * dsv[e] + 10.f * k2 * xe) * leaky_grad_from_out(xe);
                s.dxp[note * 8 + e] = v;
                s.dxa[note * 8 + e] = v;
            }
        }
        __syncwarp();
        // phase B: occupied cells swap their empty share for the true one (two cells of a row may share a note: atomics)
        for (uint32_t i = a.off[row] + lane; i < a.off[row + 1]; i += 32) {
            const CellRec r = a.recs[i];
            const int note = (int)(r.cell % 56);
            float ae[8], xp[8], bf[7], xt[8];
            me_note(k, s.od, note, ae, xp);
            me_occupied(k, r.p, xp, bf, xt);
            const float *d = a.dmelody + u * 4480 + r.cell * 8;
#pragma unroll
            for (int e = 0; e < 8; e++) {
                const float xe = leaky(xp[e]), x = leaky(xt[e]);
                const float de = (k1 * d[e] + k2 * xe) * leaky_grad_from_out(xe);
                const float dxt = (k1 * d[e] + k2 * x) * leaky_grad_from_out(x);
                atomicAdd(&s.dxp[note * 8 + e], -de);
                atomicAdd(&s.dxa[note * 8 + e], dxt - de);
            }
        }
        __syncwarp();
        // phase C: outer products with the per-note input; the outer add's gradient per note overwrites dxp
        for (int note = lane; note < 56; note += 32) {
            float ae[8], xp[8], da[8];
            me_note(k, s.od, note, ae, xp);
#pragma unroll
            for (int e = 0; e < 8; e++) {
                da[e] = s.dxa[note * 8 + e];
                Sp[e] += s.dxp[note * 8 + e]; Sa[e] += da[e];
#pragma unroll
                for (int q = 0; q < 8; q++) gWa[e * 8 + q] = fmaf(da[e], ae[q], gWa[e * 8 + q]);
            }
#pragma unroll
            for (int q = 0; q < 8; q++) {
                float v = 0.f;
#pragma unroll
                for (int e = 0; e < 8; e++) v = fmaf(c_me[kMeWl + e * 15 + q], da[e], v);
                s.dxp[note * 8 + q] = v * leaky_grad_from_out(ae[q]);
            }
        }
        __syncwarp();
        // octave rows sum over the 7 degrees, degree rows over the 8 octaves (note = o * 7 + d)
        for (int i = lane; i < 120; i += 32) {
            float v = 0.f;
            if (i < 64) {
                const int o = i >> 3, q = i & 7;
#pragma unroll
                for (int d = 0; d < 7; d++) v += s.dxp[(o * 7 + d) * 8 + q];
            } else {
                const int d = (i - 64) >> 3, q = i & 7;
#pragma unroll
                for (int o = 0; o < 8; o++) v += s.dxp[(o * 7 + d) * 8 + q];
            }
            a.dod[row * 120 + i] = v * leaky_grad_from_out(s.od[i]);
        }
    }
    // ---- flush: warp-reduce the register accumulators, add the empty-cell share of the bf branch
    __syncwarp();
#pragma unroll
    for (int e = 0; e < 8; e++) { Sp[e] = warp_sum(Sp[e]); Sa[e] = warp_sum(Sa[e]); }
#pragma unroll
    for (int i = 0; i < 64; i++) {
        const float v = warp_sum(gWa[i]);
        if (lane == 0 && v != 0.f) atomicAdd(a.gWl + (i >> 3) * 15 + (i & 7), v);
    }
    if (lane == 0) {
#pragma unroll
        for (int e = 0; e < 8; e++) {
            if (Sa[e] != 0.f) atomicAdd(a.gbl + e, Sa[e]);
#pragma unroll
            for (int q = 0; q < 7; q++) {
                const float v = Sp[e] * k.bfe[q];
                if (v != 0.f) atomicAdd(a.gWl + e * 15 + 8 + q, v);
            }
        }
#pragma unroll
        for (int q = 0; q < 7; q++) {
            float dbf = 0.f;
#pragma unroll
            for (int e = 0; e < 8; e++) dbf = fmaf(c_me[kMeWl + e * 15 + 8 + q], Sp[e], dbf);
            const float v = dbf * leaky_grad_from_out(k.bfe[q]);
            if (v != 0.f) atomicAdd(a.gbc + q, v);
        }
    }
}

// warp per beat row, lane per occupied cell: gradients of linear.weight[:, 8:] (56), channels_linear bias (7) and
// weight (35) in per-lane registers, reduced once per CTA
__global__ void __launch_bounds__(256)
k_me_bwd_cells(MeArgs2 a) {
    __shared__ float od_s[8][120];
    __shared__ float red[98];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const MeConsts k = me_consts();
    const long long BT = (long long)a.B * a.T, rows = (long long)a.N * a.C * BT;
    float g2[98];                       // [0,56) linear.weight[e][8 + q] at e * 7 + q, [56,63) bc, [63,98) Wc[q][p]
#pragma unroll
    for (int i = 0; i < 98; i++) g2[i] = 0.f;
    if (threadIdx.x < 98) red[threadIdx.x] = 0.f;
    for (long long row = (long long)blockIdx.x * 8 + warp; row < rows; row += (long long)gridDim.x * 8) {
        const uint32_t lo = a.off[row], hi = a.off[row + 1];
        if (lo == hi) continue;
        const long long nc = row / BT, bt = row - nc * BT;
        const int n = (int)(nc / a.C), c = (int)(nc - (long long)n * a.C);
        float wsum = 0.f;
        for (int j = 0; j < a.C; j++) wsum += sqrtf(1.f + a.ss[n * a.C + j]);
        const float wc = sqrtf(1.f + a.ss[n * a.C + c]);
        const float k1 = wc / wsum;
        const float k2 = (a.stats[n * (a.C + 1) + c] - a.stats[n * (a.C + 1) + a.C]) / (wsum * wc);
        const long long u = (long long)n * BT + bt;
        __syncwarp();
        for (int i = lane; i < 120; i += 32) od_s[warp][i] = a.od[row * 120 + i];
        __syncwarp();
        for (uint32_t i = lo + lane; i < hi; i += 32) {
            const CellRec r = a.recs[i];
            const int note = (int)(r.cell % 56);
            float ae[8], xp[8], bf[7], xt[8], dxt[8];
            me_note(k, od_s[warp], note, ae, xp);
            me_occupied(k, r.p, xp, bf, xt);
            const float *d = a.dmelody + u * 4480 + r.cell * 8;
#pragma unroll
            for (int e = 0; e < 8; e++) {
                const float x = leaky(xt[e]);
                dxt[e] = (k1 * d[e] + k2 * x) * leaky_grad_from_out(x);
            }
#pragma unroll
            for (int q = 0; q < 7; q++) {
                float dbf = 0.f;
#pragma unroll
                for (int e = 0; e < 8; e++) {
                    g2[e * 7 + q] = fmaf(dxt[e], bf[q], g2[e * 7 + q]);
                    dbf = fmaf(c_me[kMeWl + e * 15 + 8 + q], dxt[e], dbf);
                }
                dbf *= leaky_grad_from_out(bf[q]);
                g2[56 + q] += dbf;
#pragma unroll
                for (int p = 0; p < 5; p++) g2[63 + q * 5 + p] = fmaf(dbf, r.p[p], g2[63 + q * 5 + p]);
            }
        }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 98; i++) {
        const float v = warp_sum(g2[i]);
        if (lane == 0 && v != 0.f) atomicAdd(&red[i], v);
    }
    __syncthreads();
    if (threadIdx.x < 98) {
        const float v = red[threadIdx.x];
        const int i = threadIdx.x;
        if (v != 0.f) {
            if (i < 56) atomicAdd(a.gWl + (i / 7) * 15 + 8 + (i % 7), v);
            else if (i < 63) atomicAdd(a.gbc + (i - 56), v);
            else atomicAdd(a.gWc + (i - 63), v);
        }
    }
}

}  // namespace mstm
