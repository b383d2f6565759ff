// mst_model_kernels2.cuh — LSTM recurrences and `combine` (forward + backward), fp32.
#pragma once
#include "mst_model_kernels.cuh"

namespace mstm {

// ------------------------------------------------------------------------------------------------
// LSTM (nn.LSTM, 1 layer, batch_first; style/model.py:62,69,113,120,156,462,468).  The input projection
// x·W_ih^T + b_ih + b_hh for all time steps is one GEMM (xproj); only h·W_hh^T runs serially here.
// One CTA carries up to kLstmG sequences through time; thread j owns gate row j (gate order i, f, g, o).
// W_hh^T lives in shared memory when it fits (H = 64: 64 KB), else it streams from L2 (H = 192).
// Saved for backward: post-activation gates, cell states, and h_{t-1} per step.
// ------------------------------------------------------------------------------------------------
constexpr int kLstmG = 8;

struct LstmArgs {
    const float *xproj;   // (S, T, 4H)
    const float *w;       // forward: W_hh^T (H, 4H); backward: W_hh (4H, H)
    float *out;           // forward: h written at out[(s*T + t) * out_ld + out_off + k]
    const float *dout;    // backward: gradient w.r.t. out, same addressing
    long long out_ld; int out_off;
    float *gates;         // (S, T, 4H) post-activation
    float *cells;         // (S, T, H)
    float *hprev;         // (S, T, H)  h_{t-1} as seen by step t
    float *dgates;        // backward out: (S, T, 4H) pre-activation gradients
    int S, T, H, G;
    int reverse;
    int w_smem;
};

constexpr int kLstmMaxJobs = 4;
struct LstmJobs { LstmArgs j[kLstmMaxJobs]; };   // independent recurrences of equal H launched together (blockIdx.y = job)

__global__ void k_lstm_fwd(LstmJobs jobs) {
    const LstmArgs &a = jobs.j[blockIdx.y];
    extern __shared__ float sm[];
    const int H = a.H, H4 = 4 * a.H, G = a.G;
    float *h_s = sm;                 // [G][H]
    float *c_s = h_s + G * H;        // [G][H]
    float *g_s = c_s + G * H;        // [G][4H]
    float *w_s = g_s + G * H4;       // [H][4H] if w_smem
    const int j = threadIdx.x;
    if (a.w_smem) for (int i = j; i < H * H4; i += blockDim.x) w_s[i] = a.w[i];
    const float *w = a.w_smem ? w_s : a.w;
    // persistent over the sequence groups: W_hh is staged once per CTA.  (Keeping the thread's W_hh row in registers, as
    // k_lstm_fwd_r64 and the backward kernel below do, measured 11 % SLOWER here: 126 registers, the eight sequences' packed
    // accumulators and the row compete for the register file.)
    for (int s0 = blockIdx.x * G; s0 < a.S; s0 += gridDim.x * G) {
    const int ng = min(G, a.S - s0);
    for (int i = j; i < G * H; i += blockDim.x) { h_s[i] = 0.f; c_s[i] = 0.f; }
    __syncthreads();
    float xnext[kLstmG];      // input projection of the next step, fetched while this step computes
    {
        const int t0 = a.reverse ? a.T - 1 : 0;
#pragma unroll
        for (int g = 0; g < kLstmG; g++) xnext[g] = (j < H4 && g < ng) ? a.xproj[((long long)(s0 + g) * a.T + t0) * H4 + j] : 0.f;
    }
    for (int step = 0; step < a.T; step++) {
        const int t = a.reverse ? a.T - 1 - step : step;
        if (j < H4) {
            float acc[kLstmG];
#pragma unroll
            for (int g = 0; g < kLstmG; g++) acc[g] = xnext[g];
            if (step + 1 < a.T) {
                const int tn = a.reverse ? t - 1 : t + 1;
#pragma unroll
                for (int g = 0; g < kLstmG; g++) if (g < ng) xnext[g] = a.xproj[((long long)(s0 + g) * a.T + tn) * H4 + j];
            }
            if (ng == 1) {
                // one sequence per CTA (few long sequences): W_hh streams from L2 every step, so keep 16 loads in flight
                float a0 = acc[0], a1 = 0.f, a2 = 0.f, a3 = 0.f;
                const float *wp = w + j;
                int k = 0;
                for (; k + 16 <= H; k += 16) {
                    float wv[16];
#pragma unroll
                    for (int u = 0; u < 16; u++) wv[u] = wp[(k + u) * H4];
#pragma unroll
                    for (int u = 0; u < 16; u += 4) {
                        a0 = fmaf(wv[u], h_s[k + u], a0);
                        a1 = fmaf(wv[u + 1], h_s[k + u + 1], a1);
                        a2 = fmaf(wv[u + 2], h_s[k + u + 2], a2);
                        a3 = fmaf(wv[u + 3], h_s[k + u + 3], a3);
                    }
                }
                for (; k < H; k++) a0 = fmaf(wp[k * H4], h_s[k], a0);
                acc[0] = (a0 + a1) + (a2 + a3);
            } else {
                // G == kLstmG sequences: h is kept k-major ([k][G]) so one step reads the 8 states with two 16-byte loads
                // (packed fp32: four FFMA2 per k for the eight sequences)
                f32x2 ac2[4] = {pk2(acc[0], acc[1]), pk2(acc[2], acc[3]), pk2(acc[4], acc[5]), pk2(acc[6], acc[7])};
#pragma unroll 4
                for (int k = 0; k < H; k++) {
                    const float wv = w[k * H4 + j];
                    const f32x2 ww = pk2(wv, wv);
                    const ulonglong2 h0 = *reinterpret_cast<const ulonglong2 *>(h_s + k * kLstmG);
                    const ulonglong2 h1 = *reinterpret_cast<const ulonglong2 *>(h_s + k * kLstmG + 4);
                    ac2[0] = fma2(ww, h0.x, ac2[0]); ac2[1] = fma2(ww, h0.y, ac2[1]);
                    ac2[2] = fma2(ww, h1.x, ac2[2]); ac2[3] = fma2(ww, h1.y, ac2[3]);
                }
#pragma unroll
                for (int g2 = 0; g2 < 4; g2++) upk2(ac2[g2], acc[g2 * 2], acc[g2 * 2 + 1]);
            }
            const bool is_tanh = (j >= 2 * H) && (j < 3 * H);
#pragma unroll
            for (int g = 0; g < kLstmG; g++) if (g < ng) {
                const float v = is_tanh ? tanhf(acc[g]) : sigmoid_fast(acc[g]);
                g_s[g * H4 + j] = v;
                a.gates[((long long)(s0 + g) * a.T + t) * H4 + j] = v;
            }
        }
        __syncthreads();
        for (int i = j; i < ng * H; i += blockDim.x) {
            const int g = i / H, k = i - g * H;
            const float *gs = g_s + g * H4;
            const long long row = (long long)(s0 + g) * a.T + t;
            const int hi = G == 1 ? k : k * kLstmG + g;          // k-major state when several sequences share the CTA
            a.hprev[row * H + k] = h_s[hi];
            const float c = gs[H + k] * c_s[i] + gs[k] * gs[2 * H + k];
            const float h = gs[3 * H + k] * tanhf(c);
            c_s[i] = c;
            h_s[hi] = h;
            a.cells[row * H + k] = c;
            a.out[row * a.out_ld + a.out_off + k] = h;
        }
        __syncthreads();
    }
    }
}

__global__ void k_lstm_bwd(LstmJobs jobs) {
    const LstmArgs &a = jobs.j[blockIdx.y];
    extern __shared__ float sm[];
    const int H = a.H, H4 = 4 * a.H, G = a.G;
    float *dh_s = sm;                 // [G][H]  dL/dh carried from the later step
    float *dc_s = dh_s + G * H;       // [G][H]
    float *dg_s = dc_s + G * H;       // [G][4H]
    float *part_s = dg_s + G * H4;    // [4][G][H] partial dh sums
    float *w_s = part_s + 4 * G * H;  // [4H][H] if w_smem
    const int j = threadIdx.x;
    if (a.w_smem) for (int i = j; i < H * H4; i += blockDim.x) w_s[i] = a.w[i];
    const float *w = a.w_smem ? w_s : a.w;
    float wr[64];                     // G > 1, H = 64, 256 threads: rows 64 (j / 64) .. + 63 of W_hh, column j % 64
    if (H == 64 && G > 1 && blockDim.x == 256) {
#pragma unroll
        for (int i = 0; i < 64; i++) wr[i] = a.w[((j >> 6) * 64 + i) * H + (j & 63)];
    }
    for (int s0 = blockIdx.x * G; s0 < a.S; s0 += gridDim.x * G) {          // persistent over the sequence groups
    const int ng = min(G, a.S - s0);
    for (int i = j; i < G * H; i += blockDim.x) { dh_s[i] = 0.f; dc_s[i] = 0.f; }
    __syncthreads();
    for (int step = a.T - 1; step >= 0; step--) {
        const int t = a.reverse ? a.T - 1 - step : step;          // time index of forward step `step`
        const int t_prev = a.reverse ? t + 1 : t - 1;             // where c_{step-1} lives
        for (int i = j; i < ng * H; i += blockDim.x) {
            const int g = i / H, k = i - g * H;
            const long long row = (long long)(s0 + g) * a.T + t;
            const float *gs = a.gates + row * H4;
            const float gi = gs[k], gf = gs[H + k], gg = gs[2 * H + k], go = gs[3 * H + k];
            const float c = a.cells[row * H + k];
            const float c_prev = step > 0 ? a.cells[((long long)(s0 + g) * a.T + t_prev) * H + k] : 0.f;
            const float tc = tanhf(c);
            const float dh = dh_s[i] + (a.dout ? a.dout[row * a.out_ld + a.out_off + k] : 0.f);
            const float dc = dc_s[i] + dh * go * (1.f - tc * tc);
            const float d_i = dc * gg * gi * (1.f - gi);
            const float d_f = dc * c_prev * gf * (1.f - gf);
            const float d_g = dc * gi * (1.f - gg * gg);
            const float d_o = dh * tc * go * (1.f - go);
            dc_s[i] = dc * gf;
            float *dgo = a.dgates + row * H4;
            dgo[k] = d_i; dgo[H + k] = d_f; dgo[2 * H + k] = d_g; dgo[3 * H + k] = d_o;
            if (G == 1) {
                float *ds = dg_s;
                ds[k] = d_i; ds[H + k] = d_f; ds[2 * H + k] = d_g; ds[3 * H + k] = d_o;
            } else {                                              // r-major: [4H][G]
                dg_s[k * kLstmG + g] = d_i; dg_s[(H + k) * kLstmG + g] = d_f;
                dg_s[(2 * H + k) * kLstmG + g] = d_g; dg_s[(3 * H + k) * kLstmG + g] = d_o;
            }
        }
        __syncthreads();
        // dh_{step-1} = W_hh^T · dgates.  With few sequences per CTA the 4H rows are split over `parts` thread groups
        // (partials through shared memory) and every thread keeps 16 loads in flight.
        const int items = ng * H;
        const int parts = min(4, (int)blockDim.x / items);
        if (G > 1) {
            // thread (part, k): the 8 sequences' dh[k] over a quarter of the 4H rows, 16-byte loads of the r-major dgates
            const int qparts = min(4, (int)blockDim.x / H);
            const int part = j / H, k = j - part * H;
            if (part < qparts) {
                const int r_lo = (int)((long long)H4 * part / qparts), r_hi = (int)((long long)H4 * (part + 1) / qparts);
                f32x2 ac2[4] = {pk2(0.f, 0.f), pk2(0.f, 0.f), pk2(0.f, 0.f), pk2(0.f, 0.f)};
                if (H == 64 && qparts == 4) {
#pragma unroll
                    for (int i = 0; i < 64; i++) {
                        const f32x2 ww = pk2(wr[i], wr[i]);
                        const ulonglong2 d0 = *reinterpret_cast<const ulonglong2 *>(dg_s + (r_lo + i) * kLstmG);
                        const ulonglong2 d1 = *reinterpret_cast<const ulonglong2 *>(dg_s + (r_lo + i) * kLstmG + 4);
                        ac2[0] = fma2(ww, d0.x, ac2[0]); ac2[1] = fma2(ww, d0.y, ac2[1]);
                        ac2[2] = fma2(ww, d1.x, ac2[2]); ac2[3] = fma2(ww, d1.y, ac2[3]);
                    }
                } else
#pragma unroll 4
                for (int r = r_lo; r < r_hi; r++) {
                    const float wv = w[r * H + k];
                    const f32x2 ww = pk2(wv, wv);
                    const ulonglong2 d0 = *reinterpret_cast<const ulonglong2 *>(dg_s + r * kLstmG);
                    const ulonglong2 d1 = *reinterpret_cast<const ulonglong2 *>(dg_s + r * kLstmG + 4);
                    ac2[0] = fma2(ww, d0.x, ac2[0]); ac2[1] = fma2(ww, d0.y, ac2[1]);
                    ac2[2] = fma2(ww, d1.x, ac2[2]); ac2[3] = fma2(ww, d1.y, ac2[3]);
                }
                float acc[kLstmG];
#pragma unroll
                for (int g2 = 0; g2 < 4; g2++) upk2(ac2[g2], acc[g2 * 2], acc[g2 * 2 + 1]);
#pragma unroll
                for (int g = 0; g < kLstmG; g++) part_s[(part * kLstmG + g) * H + k] = acc[g];
            }
            __syncthreads();
            for (int i = j; i < items; i += blockDim.x) {
                const int g = i / H, k = i - g * H;
                float v = 0.f;
                for (int q = 0; q < qparts; q++) v += part_s[(q * kLstmG + g) * H + k];
                dh_s[i] = v;
            }
        } else if (parts >= 2) {
            const int part = j / items, i = j - part * items;
            if (part < parts) {
                const int g = i / H, k = i - g * H;
                const float *ds = dg_s + g * H4;
                const int r_lo = (int)((long long)H4 * part / parts), r_hi = (int)((long long)H4 * (part + 1) / parts);
                float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
                const float *wp = w + k;
                int r = r_lo;
                for (; r + 16 <= r_hi; r += 16) {
                    float wv[16];
#pragma unroll
                    for (int u = 0; u < 16; u++) wv[u] = wp[(r + u) * H];
#pragma unroll
                    for (int u = 0; u < 16; u += 4) {
                        a0 = fmaf(wv[u], ds[r + u], a0);
                        a1 = fmaf(wv[u + 1], ds[r + u + 1], a1);
                        a2 = fmaf(wv[u + 2], ds[r + u + 2], a2);
                        a3 = fmaf(wv[u + 3], ds[r + u + 3], a3);
                    }
                }
                for (; r < r_hi; r++) a0 = fmaf(wp[r * H], ds[r], a0);
                part_s[part * items + i] = (a0 + a1) + (a2 + a3);
            }
            __syncthreads();
            if (j < items) {
                float v = part_s[j];
                for (int q = 1; q < parts; q++) v += part_s[q * items + j];
                dh_s[j] = v;
            }
        } else {
            for (int i = j; i < items; i += blockDim.x) {
                const int g = i / H, k = i - g * H;
                const float *ds = dg_s + g * H4;
                float acc = 0.f;
                for (int r = 0; r < H4; r++) acc = fmaf(w[r * H + k], ds[r], acc);
                dh_s[i] = acc;
            }
        }
        __syncthreads();
    }
    }
}

// ------------------------------------------------------------------------------------------------
// Bars-axis recurrences (style/model.py:69,120: one long sequence per sample, H = 64): one sequence per CTA and the
// step latency is everything.  Thread j keeps its row of W_hh in 64 REGISTERS (k_lstm_fwd streams it from shared memory:
// 64 loads per thread and step, the shared-memory pipe was the bound), the state is read with 16-byte broadcasts.
// Same partial-sum assignment and order as the one-sequence branch of k_lstm_fwd / k_lstm_bwd: bit-identical results.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_lstm_fwd_r64(LstmJobs jobs) {
    const LstmArgs &a = jobs.j[blockIdx.y];
    constexpr int H = 64, H4 = 256;
    __shared__ __align__(16) float h_s[H];
    __shared__ float c_s[H];
    __shared__ float g_s[H4];
    const int j = threadIdx.x;
    float wr[H];
#pragma unroll
    for (int k = 0; k < H; k++) wr[k] = a.w[k * H4 + j];            // W_hh^T (H, 4H): column j
    const bool is_tanh = (j >= 2 * H) && (j < 3 * H);
    for (int s = blockIdx.x; s < a.S; s += gridDim.x) {
        if (j < H) { h_s[j] = 0.f; c_s[j] = 0.f; }
        __syncthreads();
        // the input projection is fetched FOUR steps ahead (a step is shorter than the L2 latency): four registers, the
        // step loop unrolled by four so that each is addressed statically
        auto x_at = [&](int step) { return a.xproj[((long long)s * a.T + (a.reverse ? a.T - 1 - step : step)) * H4 + j]; };
        float xr[4];
#pragma unroll
        for (int d = 0; d < 4; d++) xr[d] = d < a.T ? x_at(d) : 0.f;
        for (int step0 = 0; step0 < a.T; step0 += 4) {
#pragma unroll
            for (int d = 0; d < 4; d++) {
                const int step = step0 + d;
                if (step >= a.T) break;
                const int t = a.reverse ? a.T - 1 - step : step;
                float a0 = xr[d], a1 = 0.f, a2 = 0.f, a3 = 0.f;
                if (step + 4 < a.T) xr[d] = x_at(step + 4);
#pragma unroll
                for (int k = 0; k < H; k += 4) {
                    const float4 hv = *reinterpret_cast<const float4 *>(h_s + k);
                    a0 = fmaf(wr[k], hv.x, a0); a1 = fmaf(wr[k + 1], hv.y, a1);
                    a2 = fmaf(wr[k + 2], hv.z, a2); a3 = fmaf(wr[k + 3], hv.w, a3);
                }
                const float acc = (a0 + a1) + (a2 + a3);
                const float v = is_tanh ? tanhf(acc) : sigmoid_fast(acc);
                const long long row = (long long)s * a.T + t;
                g_s[j] = v;
                a.gates[row * H4 + j] = v;
                __syncthreads();
                if (j < H) {
                    a.hprev[row * H + j] = h_s[j];
                    const float c = g_s[H + j] * c_s[j] + g_s[j] * g_s[2 * H + j];
                    const float h = g_s[3 * H + j] * tanhf(c);
                    c_s[j] = c;
                    h_s[j] = h;
                    a.cells[row * H + j] = c;
                    a.out[row * a.out_ld + a.out_off + j] = h;
                }
                __syncthreads();
            }
        }
    }
}

__global__ void __launch_bounds__(256)
k_lstm_bwd_r64(LstmJobs jobs) {
    const LstmArgs &a = jobs.j[blockIdx.y];
    constexpr int H = 64, H4 = 256;
    __shared__ float dh_s[H], dc_s[H];
    __shared__ __align__(16) float dg_s[H4];
    __shared__ float part_s[4 * H];
    const int j = threadIdx.x;
    const int part = j / H, k = j - part * H;
    float wr[H];                                                     // W_hh (4H, H): rows part * 64 .., column k
#pragma unroll
    for (int r = 0; r < H; r++) wr[r] = a.w[(part * H + r) * H + k];
    for (int s = blockIdx.x; s < a.S; s += gridDim.x) {
        if (j < H) { dh_s[j] = 0.f; dc_s[j] = 0.f; }
        __syncthreads();
        // the saved tensors of a step (gates, cell, previous cell, upstream gradient) do not depend on the recurrence: they are
        // loaded one step ahead, during the matrix-vector product, so their L2 latency is off the per-step critical path
        float n_gi = 0.f, n_gf = 0.f, n_gg = 0.f, n_go = 0.f, n_c = 0.f, n_cprev = 0.f, n_dout = 0.f;
        auto fetch = [&](int step) {
            const int t = a.reverse ? a.T - 1 - step : step;
            const int t_prev = a.reverse ? t + 1 : t - 1;
            const long long row = (long long)s * a.T + t;
            const float *gs = a.gates + row * H4;
            n_gi = gs[j]; n_gf = gs[H + j]; n_gg = gs[2 * H + j]; n_go = gs[3 * H + j];
            n_c = a.cells[row * H + j];
            n_cprev = step > 0 ? a.cells[((long long)s * a.T + t_prev) * H + j] : 0.f;
            n_dout = a.dout ? a.dout[row * a.out_ld + a.out_off + j] : 0.f;
        };
        if (j < H && a.T > 0) fetch(a.T - 1);
        for (int step = a.T - 1; step >= 0; step--) {
            const int t = a.reverse ? a.T - 1 - step : step;
            if (j < H) {
                const long long row = (long long)s * a.T + t;
                const float gi = n_gi, gf = n_gf, gg = n_gg, go = n_go, c = n_c, c_prev = n_cprev;
                const float tc = tanhf(c);
                const float dh = dh_s[j] + n_dout;
                const float dc = dc_s[j] + dh * go * (1.f - tc * tc);
                const float d_i = dc * gg * gi * (1.f - gi);
                const float d_f = dc * c_prev * gf * (1.f - gf);
                const float d_g = dc * gi * (1.f - gg * gg);
                const float d_o = dh * tc * go * (1.f - go);
                dc_s[j] = dc * gf;
                float *dgo = a.dgates + row * H4;
                dgo[j] = d_i; dgo[H + j] = d_f; dgo[2 * H + j] = d_g; dgo[3 * H + j] = d_o;
                dg_s[j] = d_i; dg_s[H + j] = d_f; dg_s[2 * H + j] = d_g; dg_s[3 * H + j] = d_o;
            }
            __syncthreads();
            if (j < H && step > 0) fetch(step - 1);
            {   // dh_{step-1}[k] = sum_r W_hh[r][k] dgates[r]: a quarter of the rows per thread, partials through shared memory
                const float *ds = dg_s + part * H;
                float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
                for (int r = 0; r < H; r += 4) {
                    const float4 dv = *reinterpret_cast<const float4 *>(ds + r);
                    a0 = fmaf(wr[r], dv.x, a0); a1 = fmaf(wr[r + 1], dv.y, a1);
                    a2 = fmaf(wr[r + 2], dv.z, a2); a3 = fmaf(wr[r + 3], dv.w, a3);
                }
                part_s[part * H + k] = (a0 + a1) + (a2 + a3);
            }
            __syncthreads();
            if (j < H) {
                float v = part_s[j];
                for (int q = 1; q < 4; q++) v += part_s[q * H + j];
                dh_s[j] = v;
            }
            __syncthreads();
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Narrow recurrences (SongInfoModel.bars_lstm, style/model.py:468: H = 8, one sequence of n_bars steps per sample):
// ONE WARP per sequence, lane j = gate row j (4H = 32), weights in 8 registers, the state and the gates move by
// shuffles: no shared memory and no barrier inside a step; the next step's global operands are fetched one step ahead.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32)
k_lstm_fwd_w8(LstmArgs a) {
    constexpr int H = 8, H4 = 32;
    const int j = threadIdx.x, u = j & 7, gate = j >> 3;
    float wr[H];
#pragma unroll
    for (int k = 0; k < H; k++) wr[k] = a.w[k * H4 + j];            // W_hh^T (H, 4H): column j
    for (int s = blockIdx.x; s < a.S; s += gridDim.x) {
        float h = 0.f, c = 0.f;                                      // lanes 0..7: unit u
        auto x_at = [&](int step) { return a.xproj[((long long)s * a.T + (a.reverse ? a.T - 1 - step : step)) * H4 + j]; };
        float xr[4];                                                 // input projections, fetched four steps ahead
#pragma unroll
        for (int d = 0; d < 4; d++) xr[d] = d < a.T ? x_at(d) : 0.f;
        for (int step0 = 0; step0 < a.T; step0 += 4) {
#pragma unroll
            for (int d = 0; d < 4; d++) {
                const int step = step0 + d;
                if (step >= a.T) break;
                const int t = a.reverse ? a.T - 1 - step : step;
                float acc = xr[d];
                if (step + 4 < a.T) xr[d] = x_at(step + 4);
#pragma unroll
                for (int k = 0; k < H; k++) acc = fmaf(wr[k], __shfl_sync(0xffffffffu, h, k), acc);
                const float v = gate == 2 ? tanhf(acc) : sigmoid_fast(acc);
                const long long row = (long long)s * a.T + t;
                a.gates[row * H4 + j] = v;
                const float gi = __shfl_sync(0xffffffffu, v, u), gf = __shfl_sync(0xffffffffu, v, 8 + u);
                const float gg = __shfl_sync(0xffffffffu, v, 16 + u), go = __shfl_sync(0xffffffffu, v, 24 + u);
                if (j < H) {
                    a.hprev[row * H + j] = h;
                    c = gf * c + gi * gg;
                    h = go * tanhf(c);
                    a.cells[row * H + j] = c;
                    a.out[row * a.out_ld + a.out_off + j] = h;
                }
            }
        }
    }
}

__global__ void __launch_bounds__(32)
k_lstm_bwd_w8(LstmArgs a) {
    constexpr int H = 8, H4 = 32;
    const int j = threadIdx.x, u = j & 7, part = j >> 3;
    float wr[H];                                                     // W_hh (4H, H): rows part * 8 .., column u
#pragma unroll
    for (int i = 0; i < H; i++) wr[i] = a.w[(part * H + i) * H + u];
    for (int s = blockIdx.x; s < a.S; s += gridDim.x) {
        float dh_c = 0.f, dc_c = 0.f;                                // carried gradients, lanes 0..7
        // operands of the step, lanes 0..7 (unit u); fetched one step ahead
        auto fetch = [&](int step, float (&g4)[4], float &cc, float &cp, float &dout) {
            const int t = a.reverse ? a.T - 1 - step : step;
            const int t_prev = a.reverse ? t + 1 : t - 1;
            const long long row = (long long)s * a.T + t;
            const float *gs = a.gates + row * H4;
            g4[0] = gs[u]; g4[1] = gs[H + u]; g4[2] = gs[2 * H + u]; g4[3] = gs[3 * H + u];
            cc = a.cells[row * H + u];
            cp = step > 0 ? a.cells[((long long)s * a.T + t_prev) * H + u] : 0.f;
            dout = a.dout ? a.dout[row * a.out_ld + a.out_off + u] : 0.f;
        };
        float g4[4] = {0.f, 0.f, 0.f, 0.f}, cc = 0.f, cp = 0.f, dout = 0.f;
        if (j < H) fetch(a.T - 1, g4, cc, cp, dout);
        for (int step = a.T - 1; step >= 0; step--) {
            const int t = a.reverse ? a.T - 1 - step : step;
            const long long row = (long long)s * a.T + t;
            const float gi = g4[0], gf = g4[1], gg = g4[2], go = g4[3], c = cc, c_prev = cp, dout_t = dout;
            if (j < H && step > 0) fetch(step - 1, g4, cc, cp, dout);
            float d4[4] = {0.f, 0.f, 0.f, 0.f};
            if (j < H) {
                const float tc = tanhf(c);
                const float dh = dh_c + dout_t;
                const float dc = dc_c + dh * go * (1.f - tc * tc);
                d4[0] = dc * gg * gi * (1.f - gi);
                d4[1] = dc * c_prev * gf * (1.f - gf);
                d4[2] = dc * gi * (1.f - gg * gg);
                d4[3] = dh * tc * go * (1.f - go);
                dc_c = dc * gf;
            }
            // lane r = gate * 8 + unit takes dgates[r] from lane `unit`
            const float e0 = __shfl_sync(0xffffffffu, d4[0], u), e1 = __shfl_sync(0xffffffffu, d4[1], u);
            const float e2 = __shfl_sync(0xffffffffu, d4[2], u), e3 = __shfl_sync(0xffffffffu, d4[3], u);
            const float dg = part == 0 ? e0 : (part == 1 ? e1 : (part == 2 ? e2 : e3));
            a.dgates[row * H4 + j] = dg;
            // dh_{step-1}[u] = sum_r W_hh[r][u] dgates[r]: lane (part, u) sums rows part * 8 .., then the four parts
            float acc = 0.f;
#pragma unroll
            for (int i = 0; i < H; i++) acc = fmaf(wr[i], __shfl_sync(0xffffffffu, dg, part * H + i), acc);
            acc += __shfl_xor_sync(0xffffffffu, acc, 8);
            acc += __shfl_xor_sync(0xffffffffu, acc, 16);
            dh_c = acc;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// combine (style/model.py:796-815): out[n, m] = sum_k x[n,k,m] w[n,k] / sum_k w[n,k],  w = sqrt(1 + sum_m x^2).
// x is addressed as x[n * n_stride + k * k_stride + m].
// ------------------------------------------------------------------------------------------------
struct CombineArgs {
    const float *x; long long n_stride, k_stride;
    int N, K; long long M;
    float *ss;          // [N, K] sums of squares (zeroed by the caller before k_combine_sumsq)
    float *out;         // [N, M]
    const float *dout;  // [N, M]
    float *stats;       // backward: [N, K + 1]: A_k = sum dout*x_k, then B = sum dout*out (zeroed by the caller)
    float *dx;          // same addressing as x; accumulate flag below
    int accumulate;
};

__device__ __forceinline__ float block_sum(float v, float *tmp) {
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) tmp[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = 0.f;
    if (threadIdx.x < 32) {
        r = threadIdx.x < (blockDim.x >> 5) ? tmp[threadIdx.x] : 0.f;
        for (int o = 16; o; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
    }
    __syncthreads();
    return r;   // valid on thread 0
}

// grid: (chunks, N*K)
__global__ void __launch_bounds__(256)
k_combine_sumsq(CombineArgs a) {
    __shared__ float tmp[8];
    const int nk = blockIdx.y, n = nk / a.K, k = nk - n * a.K;
    const float *p = a.x + n * a.n_stride + k * a.k_stride;
    float s = 0.f;
    for (long long m = (long long)blockIdx.x * blockDim.x + threadIdx.x; m < a.M; m += (long long)gridDim.x * blockDim.x) s = fmaf(p[m], p[m], s);
    s = block_sum(s, tmp);
    if (threadIdx.x == 0) atomicAdd(a.ss + nk, s);
}

// grid: (chunks, N)
__global__ void __launch_bounds__(256)
k_combine_fwd(CombineArgs a) {
    const int n = blockIdx.y;
    float wsum = 0.f;
    for (int k = 0; k < a.K; k++) wsum += sqrtf(1.f + a.ss[n * a.K + k]);
    for (long long m = (long long)blockIdx.x * blockDim.x + threadIdx.x; m < a.M; m += (long long)gridDim.x * blockDim.x) {
        float v = 0.f;
        for (int k = 0; k < a.K; k++) v = fmaf(a.x[n * a.n_stride + k * a.k_stride + m], sqrtf(1.f + a.ss[n * a.K + k]), v);
        a.out[(long long)n * a.M + m] = v / wsum;
    }
}

// grid: (chunks, N): stats[n, k] += sum_m dout*x_k ; stats[n, K] += sum_m dout*out
__global__ void __launch_bounds__(256)
k_combine_bwd_stats(CombineArgs a) {
    __shared__ float tmp[8];
    const int n = blockIdx.y;
    for (int k = 0; k <= a.K; k++) {
        float s = 0.f;
        for (long long m = (long long)blockIdx.x * blockDim.x + threadIdx.x; m < a.M; m += (long long)gridDim.x * blockDim.x) {
            const float d = a.dout[(long long)n * a.M + m];
            const float v = k < a.K ? a.x[n * a.n_stride + k * a.k_stride + m] : a.out[(long long)n * a.M + m];
            s = fmaf(d, v, s);
        }
        s = block_sum(s, tmp);
        if (threadIdx.x == 0) atomicAdd(a.stats + n * (a.K + 1) + k, s);
    }
}

// dx[n,k,m] (+)= dout[n,m] * w_k / S + (A_k - B) / (S * w_k) * x[n,k,m]
__global__ void __launch_bounds__(256)
k_combine_bwd(CombineArgs a) {
    const int n = blockIdx.y;
    float wsum = 0.f;
    for (int k = 0; k < a.K; k++) wsum += sqrtf(1.f + a.ss[n * a.K + k]);
    const float Bs = a.stats[n * (a.K + 1) + a.K];
    for (long long m = (long long)blockIdx.x * blockDim.x + threadIdx.x; m < a.M; m += (long long)gridDim.x * blockDim.x) {
        const float d = a.dout[(long long)n * a.M + m];
        for (int k = 0; k < a.K; k++) {
            const float w = sqrtf(1.f + a.ss[n * a.K + k]);
            const long long idx = n * a.n_stride + k * a.k_stride + m;
            const float g = d * w / wsum + (a.stats[n * (a.K + 1) + k] - Bs) / (wsum * w) * a.x[idx];
            if (a.accumulate) a.dx[idx] += g; else a.dx[idx] = g;
        }
    }
}

}  // namespace mstm
