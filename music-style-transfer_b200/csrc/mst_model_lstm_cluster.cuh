// mst_model_lstm_cluster.cuh — LSTM recurrence for a wide hidden state and few long sequences (StyleEncoder.bars_lstm,
// model.py:156: H = 192, one sequence of n_bars steps per sample).  W_hh (4H x H fp32 = 590 KB) fits no single SM, and
// streaming it from L2 every step pins the one-CTA-per-sequence kernel at the per-SM L2 bandwidth.  Here a thread-block
// CLUSTER of 4 CTAs shares the sequence: CTA r keeps the W_hh slice of its H/4 hidden units (all four gates, 147 KB) in
// shared memory for the whole kernel, computes those units' gates from the full h, updates its c / h slice and writes
// the new h slice into every CTA's shared memory (distributed shared memory); one cluster barrier per time step.
// Saved tensors, argument struct and results are those of k_lstm_fwd / k_lstm_bwd.
#pragma once
#include <cooperative_groups.h>
#include "mst_model_kernels2.cuh"

namespace mstm {
namespace cg = cooperative_groups;

constexpr int kClSize = 4;       // CTAs per cluster
constexpr int kClG = 2;          // sequences per cluster
constexpr int kClFwdParts = 4;   // K split of the forward matvec (threads = 4 * HQ * parts)
constexpr int kClBwdParts = 8;   // row split of the backward matvec (threads = HQ * parts)

__host__ __device__ inline size_t lstm_cl_fwd_smem(int H) {
    const int HQ = H / kClSize;
    return sizeof(float) * ((size_t)H * 4 * HQ + 2 * kClG * H + (size_t)kClFwdParts * kClG * 4 * HQ + kClG * 4 * HQ + kClG * HQ);
}
__host__ __device__ inline size_t lstm_cl_bwd_smem(int H) {
    const int HQ = H / kClSize;
    return sizeof(float) * ((size_t)4 * H * HQ + 2 * kClG * 4 * H + (size_t)kClBwdParts * kClG * HQ + 2 * kClG * HQ);
}

// KR > 0 (H = 192: KR = 48 forward, 96 backward): the thread's weights live in KR REGISTERS instead of the shared-memory
// slice (the shared-memory pipe, not the cluster barrier, was the larger part of a step) and the state / gate-gradient
// vectors are read with 16-byte broadcasts; same partial sums in the same order.  KR = 0 keeps the slice in shared memory.
template <int KR>
__global__ void __cluster_dims__(kClSize, 1, 1) __launch_bounds__(KR ? 768 : 1024)
k_lstm_fwd_cl(LstmArgs a) {
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    extern __shared__ __align__(16) float sm[];
    const int H = a.H, H4 = 4 * H, HQ = H / kClSize, LC = 4 * HQ;       // LC local gate columns
    float *w_s = sm;                          // [H][LC]
    float *h_full = w_s + (size_t)H * LC;     // [2][G][H]
    float *part_s = h_full + 2 * kClG * H;    // [parts][G][LC]
    float *g_loc = part_s + kClFwdParts * kClG * LC;   // [G][LC]
    float *c_loc = g_loc + kClG * LC;         // [G][HQ]
    const int j = threadIdx.x;
    const int s0 = (blockIdx.x / kClSize) * kClG;
    const int ng = min(kClG, a.S - s0);
    // local column -> global gate column
    auto gcol = [&](int col) { const int gate = col / HQ, u = col - gate * HQ; return gate * H + rank * HQ + u; };
    if (KR == 0) for (int i = j; i < H * LC; i += blockDim.x) { const int k = i / LC, col = i - k * LC; w_s[i] = a.w[(long long)k * H4 + gcol(col)]; }
    for (int i = j; i < 2 * kClG * H; i += blockDim.x) h_full[i] = 0.f;
    for (int i = j; i < kClG * HQ; i += blockDim.x) c_loc[i] = 0.f;
    cluster.sync();
    const int col = j % LC, part = j / LC;
    const int k_lo = H * part / kClFwdParts, k_hi = H * (part + 1) / kClFwdParts;
    float wr[KR ? KR : 1];
    if (KR) {
        const int gc = gcol(col);
#pragma unroll
        for (int i = 0; i < KR; i++) wr[i] = a.w[(long long)(k_lo + i) * H4 + gc];
    }
    // thread j < G * LC also owns the activation of (g, col): its input projection is fetched one step ahead
    const int ag = j / LC, acol = j - ag * LC;
    const bool act_thread = j < kClG * LC && ag < ng;
    const int a_gc = gcol(acol);
    float xnext = 0.f;
    if (act_thread) xnext = a.xproj[((long long)(s0 + ag) * a.T + (a.reverse ? a.T - 1 : 0)) * H4 + a_gc];
    int cur = 0;
    for (int step = 0; step < a.T; step++) {
        const int t = a.reverse ? a.T - 1 - step : step;
        const float xcur = xnext;
        if (act_thread && step + 1 < a.T) xnext = a.xproj[((long long)(s0 + ag) * a.T + (a.reverse ? t - 1 : t + 1)) * H4 + a_gc];
        {
            const float *h0 = h_full + (cur * kClG + 0) * H, *h1 = h_full + (cur * kClG + 1) * H;
            float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
            if (KR && ng == 1) {
                // one sequence in the cluster (N = 1, the reference's batch size): half the loads and FMAs per step
#pragma unroll
                for (int i = 0; i < KR; i += 4) {
                    const float4 x0 = *reinterpret_cast<const float4 *>(h0 + k_lo + i);
                    a0 = fmaf(wr[i], x0.x, a0); a1 = fmaf(wr[i + 1], x0.y, a1);
                    a0 = fmaf(wr[i + 2], x0.z, a0); a1 = fmaf(wr[i + 3], x0.w, a1);
                }
            } else if (KR) {
#pragma unroll
                for (int i = 0; i < KR; i += 4) {
                    const float4 x0 = *reinterpret_cast<const float4 *>(h0 + k_lo + i), x1 = *reinterpret_cast<const float4 *>(h1 + k_lo + i);
                    a0 = fmaf(wr[i], x0.x, a0); b0 = fmaf(wr[i], x1.x, b0);
                    a1 = fmaf(wr[i + 1], x0.y, a1); b1 = fmaf(wr[i + 1], x1.y, b1);
                    a0 = fmaf(wr[i + 2], x0.z, a0); b0 = fmaf(wr[i + 2], x1.z, b0);
                    a1 = fmaf(wr[i + 3], x0.w, a1); b1 = fmaf(wr[i + 3], x1.w, b1);
                }
            } else {
            int k = k_lo;
            for (; k + 2 <= k_hi; k += 2) {
                const float w0 = w_s[k * LC + col], w1 = w_s[(k + 1) * LC + col];
                a0 = fmaf(w0, h0[k], a0); b0 = fmaf(w0, h1[k], b0);
                a1 = fmaf(w1, h0[k + 1], a1); b1 = fmaf(w1, h1[k + 1], b1);
            }
            for (; k < k_hi; k++) { const float w0 = w_s[k * LC + col]; a0 = fmaf(w0, h0[k], a0); b0 = fmaf(w0, h1[k], b0); }
            }
            part_s[(part * kClG + 0) * LC + col] = a0 + a1;
            part_s[(part * kClG + 1) * LC + col] = b0 + b1;
        }
        __syncthreads();
        float act_v = 0.f;
        if (act_thread) {
            float v = xcur;
#pragma unroll
            for (int p = 0; p < kClFwdParts; p++) v += part_s[(p * kClG + ag) * LC + acol];
            const int gate = acol / HQ;
            v = gate == 2 ? tanhf(v) : sigmoidf(v);
            g_loc[ag * LC + acol] = v;
            act_v = v;
        }
        __syncthreads();
        // the new h goes to every CTA of the cluster first and the barrier is ARRIVED at right after it; the saved tensors are
        // written to global memory between arrive and wait, so the barrier's release does not have to wait for those stores
        const bool cell_thread = j < kClG * HQ && j / HQ < ng;
        float sv_hprev = 0.f, sv_c = 0.f, sv_h = 0.f;
        if (cell_thread) {
            const int g = j / HQ, u = j - g * HQ;
            const int unit = rank * HQ + u;
            const float *gs = g_loc + g * LC;
            sv_hprev = h_full[(cur * kClG + g) * H + unit];
            const float c = gs[HQ + u] * c_loc[j] + gs[u] * gs[2 * HQ + u];
            const float h = gs[3 * HQ + u] * tanhf(c);
            c_loc[j] = c;
            sv_c = c; sv_h = h;
#pragma unroll
            for (int rr = 0; rr < kClSize; rr++) {
                float *remote = cluster.map_shared_rank(h_full, rr);
                remote[((cur ^ 1) * kClG + g) * H + unit] = h;
            }
        }
        cluster.barrier_arrive();
        if (act_thread) a.gates[((long long)(s0 + ag) * a.T + t) * H4 + a_gc] = act_v;
        if (cell_thread) {
            const int g = j / HQ, u = j - g * HQ;
            const int unit = rank * HQ + u;
            const long long row = (long long)(s0 + g) * a.T + t;
            a.hprev[row * H + unit] = sv_hprev;
            a.cells[row * H + unit] = sv_c;
            a.out[row * a.out_ld + a.out_off + unit] = sv_h;
        }
        cluster.barrier_wait();
        cur ^= 1;
    }
}

template <int KR>
__global__ void __cluster_dims__(kClSize, 1, 1) __launch_bounds__(KR ? 384 : 1024)
k_lstm_bwd_cl(LstmArgs a) {
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    extern __shared__ __align__(16) float sm[];
    const int H = a.H, H4 = 4 * H, HQ = H / kClSize;
    float *w_s = sm;                               // [4H][HQ]: W_hh[:, own units]
    float *dg_full = w_s + (size_t)H4 * HQ;        // [2][G][4H]
    float *part_s = dg_full + 2 * kClG * H4;       // [parts][G][HQ]
    float *dh_loc = part_s + kClBwdParts * kClG * HQ;   // [G][HQ]
    float *dc_loc = dh_loc + kClG * HQ;            // [G][HQ]
    const int j = threadIdx.x;
    const int s0 = (blockIdx.x / kClSize) * kClG;
    const int ng = min(kClG, a.S - s0);
    if (KR == 0) for (int i = j; i < H4 * HQ; i += blockDim.x) { const int r = i / HQ, u = i - r * HQ; w_s[i] = a.w[(long long)r * H + rank * HQ + u]; }
    for (int i = j; i < 2 * kClG * H4; i += blockDim.x) dg_full[i] = 0.f;
    for (int i = j; i < kClG * HQ; i += blockDim.x) { dh_loc[i] = 0.f; dc_loc[i] = 0.f; }
    cluster.sync();
    const int ku = j % HQ, part = j / HQ;
    const int r_lo = H4 * part / kClBwdParts, r_hi = H4 * (part + 1) / kClBwdParts;
    float wr[KR ? KR : 1];
    if (KR) {
#pragma unroll
        for (int i = 0; i < KR; i++) wr[i] = a.w[(long long)(r_lo + i) * H + rank * HQ + ku];
    }
    int cur = 0;
    // cell threads: the saved tensors of a step (4 gates, cell, previous cell, upstream gradient) do not depend on the
    // recurrence - they are loaded one step ahead, so that their L2 latency is off the per-step critical path
    const bool cell_thread = j < kClG * HQ && j / HQ < ng;
    const int cg_ = j / HQ, cu_ = j - cg_ * HQ, cunit = rank * HQ + cu_;
    float n_gi = 0.f, n_gf = 0.f, n_gg = 0.f, n_go = 0.f, n_c = 0.f, n_cprev = 0.f, n_dout = 0.f;
    auto fetch = [&](int step) {
        const int t = a.reverse ? a.T - 1 - step : step;
        const int t_prev = a.reverse ? t + 1 : t - 1;
        const long long row = (long long)(s0 + cg_) * a.T + t;
        const float *gs = a.gates + row * H4;
        n_gi = gs[cunit]; n_gf = gs[H + cunit]; n_gg = gs[2 * H + cunit]; n_go = gs[3 * H + cunit];
        n_c = a.cells[row * H + cunit];
        n_cprev = step > 0 ? a.cells[((long long)(s0 + cg_) * a.T + t_prev) * H + cunit] : 0.f;
        n_dout = a.dout ? a.dout[row * a.out_ld + a.out_off + cunit] : 0.f;
    };
    if (cell_thread && a.T > 0) fetch(a.T - 1);
    for (int step = a.T - 1; step >= 0; step--) {
        const int t = a.reverse ? a.T - 1 - step : step;
        float sv_i = 0.f, sv_f = 0.f, sv_g = 0.f, sv_o = 0.f;
        if (j < kClG * HQ) {
            const int g = j / HQ, u = j - g * HQ;
            if (g < ng) {
                const int unit = rank * HQ + u;
                const float gi = n_gi, gf = n_gf, gg = n_gg, go = n_go, c = n_c, c_prev = n_cprev, dout_v = n_dout;
                const float tc = tanhf(c);
                const float dh = dh_loc[j] + dout_v;
                const float dc = dc_loc[j] + dh * go * (1.f - tc * tc);
                const float d_i = dc * gg * gi * (1.f - gi);
                const float d_f = dc * c_prev * gf * (1.f - gf);
                const float d_g = dc * gi * (1.f - gg * gg);
                const float d_o = dh * tc * go * (1.f - go);
                dc_loc[j] = dc * gf;
#pragma unroll
                for (int rr = 0; rr < kClSize; rr++) {
                    float *remote = cluster.map_shared_rank(dg_full, rr) + (cur * kClG + g) * H4;
                    remote[unit] = d_i; remote[H + unit] = d_f; remote[2 * H + unit] = d_g; remote[3 * H + unit] = d_o;
                }
                sv_i = d_i; sv_f = d_f; sv_g = d_g; sv_o = d_o;
            }
        }
        // arrive right after the distributed-shared-memory stores; the global stores of the gate gradients and the loads of the
        // next step's saved tensors go between arrive and wait (the barrier's release need not wait for them)
        cluster.barrier_arrive();
        if (cell_thread) {
            float *dgo = a.dgates + ((long long)(s0 + cg_) * a.T + t) * H4;
            dgo[cunit] = sv_i; dgo[H + cunit] = sv_f; dgo[2 * H + cunit] = sv_g; dgo[3 * H + cunit] = sv_o;
            if (step > 0) fetch(step - 1);                // consumed at the top of the next step
        }
        cluster.barrier_wait();
        {   // dh_{step-1}[own units] = W_hh[:, own units]^T . dgates
            const float *d0 = dg_full + (cur * kClG + 0) * H4, *d1 = dg_full + (cur * kClG + 1) * H4;
            float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
            if (KR && ng == 1) {
#pragma unroll
                for (int i = 0; i < KR; i += 4) {
                    const float4 x0 = *reinterpret_cast<const float4 *>(d0 + r_lo + i);
                    a0 = fmaf(wr[i], x0.x, a0); a1 = fmaf(wr[i + 1], x0.y, a1);
                    a0 = fmaf(wr[i + 2], x0.z, a0); a1 = fmaf(wr[i + 3], x0.w, a1);
                }
            } else if (KR) {
#pragma unroll
                for (int i = 0; i < KR; i += 4) {
                    const float4 x0 = *reinterpret_cast<const float4 *>(d0 + r_lo + i), x1 = *reinterpret_cast<const float4 *>(d1 + r_lo + i);
                    a0 = fmaf(wr[i], x0.x, a0); b0 = fmaf(wr[i], x1.x, b0);
                    a1 = fmaf(wr[i + 1], x0.y, a1); b1 = fmaf(wr[i + 1], x1.y, b1);
                    a0 = fmaf(wr[i + 2], x0.z, a0); b0 = fmaf(wr[i + 2], x1.z, b0);
                    a1 = fmaf(wr[i + 3], x0.w, a1); b1 = fmaf(wr[i + 3], x1.w, b1);
                }
            } else {
            int r = r_lo;
            for (; r + 2 <= r_hi; r += 2) {
                const float w0 = w_s[r * HQ + ku], w1 = w_s[(r + 1) * HQ + ku];
                a0 = fmaf(w0, d0[r], a0); b0 = fmaf(w0, d1[r], b0);
                a1 = fmaf(w1, d0[r + 1], a1); b1 = fmaf(w1, d1[r + 1], b1);
            }
            for (; r < r_hi; r++) { const float w0 = w_s[r * HQ + ku]; a0 = fmaf(w0, d0[r], a0); b0 = fmaf(w0, d1[r], b0); }
            }
            part_s[(part * kClG + 0) * HQ + ku] = a0 + a1;
            part_s[(part * kClG + 1) * HQ + ku] = b0 + b1;
        }
        __syncthreads();
        if (j < kClG * HQ) {
            const int g = j / HQ, u = j - g * HQ;
            float v = 0.f;
#pragma unroll
            for (int p = 0; p < kClBwdParts; p++) v += part_s[(p * kClG + g) * HQ + u];
            dh_loc[j] = v;
        }
        __syncthreads();
        cur ^= 1;
    }
}

}  // namespace mstm
