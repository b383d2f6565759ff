// mst_model_tails_bwd.cuh — backward of the fused per-note kernels (mst_model_tails.cuh).  The per-note
// intermediates (note-rows x {15, 30, 50}) are RECOMPUTED from the same small per-row vectors the forward used,
// so the backward also never moves a note-row-sized activation through HBM: it reads the target grid (and, for
// the melody encoder, the melody gradient) once and writes only per-row gradients.
// The loss derivative w.r.t. each predicted cell (mst_model_loss.cuh: loss_cell_grad) is evaluated in place.
#pragma once
#include "mst_model_loss.cuh"
#include "mst_model_tails.cuh"

namespace mstm {

// ------------------------------------------------------------------------------------------------
// PitchedStyleApplier tail backward.  CTA = one (n, b, t, f) unit, channels processed in groups of kPsaCG.
// Outputs: dZr[u][450] (sum over channels of the od pre-activation gradient), dZi[n,c][450] (atomic, flushed when the
// CTA moves to another sample), dmelody[u][56][8], and atomics into the gradients of melody_linear / linear.
// ------------------------------------------------------------------------------------------------
constexpr int kPsaCG = 4;
constexpr int kPsaItems = kPsaCG * 56;     // 224 (channel, note) items per pass
constexpr int kPsaThreads = 256;

struct PsaBwdArgs {
    PsaArgs f;                       // forward arguments (out unused)
    const float *target;             // P (N,C,B,T,10,56,5)
    const float *coef;               // (N, kLossCoefs)
    float *dZr, *dZi;                // (N,B,T,10,450) written; (N,C,450) atomically accumulated (zeroed by the caller)
    float *dmelody;                  // (N,B,T,10,56,8) written
    float *gWm, *gbm, *gWl, *gbl;    // parameter gradients (atomic accumulate)
};

struct PsaBwdSmem {
    PsaSmemW w;                      // 435 floats
    float gW[436];                   // [0,160) Wm, [160,180) bm, [180,430) Wl, [430,435) bl
    float mel[56 * 8];
    float dmf[56 * 20];              // sum over channels of the melody-branch pre-activation gradient
    float dzr[450];
    float coef[kLossCoefs];
    float od[kPsaCG * 450];
    float dod[kPsaCG * 450];
    float s_d[kPsaItems * 5];
    float s_in[kPsaItems * 51];      // padded rows (51) to spread banks
};

static inline size_t psa_bwd_smem_bytes(int C) { return sizeof(PsaBwdSmem) + (size_t)C * 450 * sizeof(float); }

__global__ void __launch_bounds__(kPsaThreads)
k_psa_bwd(PsaBwdArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    PsaBwdSmem &s = *reinterpret_cast<PsaBwdSmem *>(smem_raw);
    float *dzi_s = reinterpret_cast<float *>(smem_raw + sizeof(PsaBwdSmem));   // [C][450], lives across units of one sample
    const PsaArgs &f = a.f;
    const int tid = threadIdx.x, C = f.C;
    psa_load_w(s.w, f.w);
    for (int i = tid; i < 436; i += blockDim.x) s.gW[i] = 0.f;
    for (int i = tid; i < C * 450; i += blockDim.x) dzi_s[i] = 0.f;
    const long long per_n = (long long)f.B * f.T * 10;
    const long long units = (long long)f.N * per_n;
    // contiguous unit ranges per CTA so that dZi is flushed at most twice per CTA and sample
    const long long per_cta = (units + gridDim.x - 1) / gridDim.x;
    const long long u_lo = (long long)blockIdx.x * per_cta, u_hi = min(units, u_lo + per_cta);
    int cur_n = -1;
    for (long long u = u_lo; u < u_hi; u++) {
        const int n = (int)(u / per_n);
        const long long btf = u - (long long)n * per_n;
        __syncthreads();
        if (n != cur_n) {
            if (cur_n >= 0)
                for (int i = tid; i < C * 450; i += blockDim.x) {
                    if (dzi_s[i] != 0.f) atomicAdd(a.dZi + (long long)cur_n * C * 450 + i, dzi_s[i]);
                    dzi_s[i] = 0.f;
                }
            cur_n = n;
            if (tid < kLossCoefs) s.coef[tid] = a.coef[n * kLossCoefs + tid];
        }
        for (int i = tid; i < 448; i += blockDim.x) s.mel[i] = f.melody[u * 448 + i];
        for (int i = tid; i < 56 * 20; i += blockDim.x) s.dmf[i] = 0.f;
        for (int i = tid; i < 450; i += blockDim.x) s.dzr[i] = 0.f;
        const float *zr = f.Zr + u * 450;
        const float *zs = f.Zs + (long long)n * 450;
        for (int c0 = 0; c0 < C; c0 += kPsaCG) {
            const int cg = min(kPsaCG, C - c0);
            __syncthreads();
            for (int i = tid; i < cg * 450; i += blockDim.x) {
                const int c = i / 450, j = i - c * 450;
                s.od[i] = leaky(f.w.od_bias[j] + zr[j] + f.Zi[((long long)n * C + c0 + c) * 450 + j] + zs[j]);
                s.dod[i] = 0.f;
            }
            __syncthreads();
            if (tid < cg * 56) {
                const int c = tid / 56, note = tid - c * 56;
                const int o = note / 7, d = note - o * 7;
                float in50[50], y[5], dpre[5];
                psa_cell(s.w, s.od + c * 450, s.mel + note * 8, note, in50, y);
                const float *tg = a.target + (((((long long)n * C + c0 + c) * per_n) + btf) * 56 + note) * 5;
                float t5[5];
#pragma unroll
                for (int e = 0; e < 5; e++) t5[e] = tg[e];
                loss_cell_grad<5>(y, t5, s.coef, dpre);
#pragma unroll
                for (int e = 0; e < 5; e++) s.s_d[tid * 5 + e] = dpre[e];
#pragma unroll
                for (int q = 0; q < 50; q++) {
                    s.s_in[tid * 51 + q] = in50[q];
                    float din = 0.f;
#pragma unroll
                    for (int e = 0; e < 5; e++) din = fmaf(s.w.Wl[e * 50 + q], dpre[e], din);
                    din *= leaky_grad_from_out(in50[q]);
                    if (q < 30) {
                        atomicAdd(&s.dod[c * 450 + o * 30 + q], din);
                        atomicAdd(&s.dod[c * 450 + 240 + d * 30 + q], din);
                    } else {
                        atomicAdd(&s.dmf[note * 20 + (q - 30)], din);
                    }
                }
            }
            __syncthreads();
            // linear (5 x 50) + bias gradient: thread = (q, cell group of 5); 51st "input" = 1 gives the bias
            if (tid < 255) {
                const int q = tid % 51, grp = tid / 51;
                float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
                for (int it = grp; it < cg * 56; it += 5) {
                    const float x = q < 50 ? s.s_in[it * 51 + q] : 1.f;
#pragma unroll
                    for (int e = 0; e < 5; e++) acc[e] = fmaf(s.s_d[it * 5 + e], x, acc[e]);
                }
#pragma unroll
                for (int e = 0; e < 5; e++) {
                    if (q < 50) atomicAdd(&s.gW[180 + e * 50 + q], acc[e]);
                    else atomicAdd(&s.gW[430 + e], acc[e]);
                }
            }
            // od pre-activation gradient of this channel group
            for (int i = tid; i < cg * 450; i += blockDim.x) {
                const int c = i / 450, j = i - c * 450;
                const float g = s.dod[i] * leaky_grad_from_out(s.od[i]);
                dzi_s[(c0 + c) * 450 + j] += g;
                atomicAdd(&s.dzr[j], g);
            }
        }
        __syncthreads();
        for (int j = tid; j < 450; j += blockDim.x) a.dZr[u * 450 + j] = s.dzr[j];
        // melody branch: dmelody = Wm^T dmf ; dWm += dmf (x) melody ; dbm += dmf
        for (int i = tid; i < 448; i += blockDim.x) {
            const int note = i >> 3, p = i & 7;
            float v = 0.f;
#pragma unroll
            for (int j = 0; j < 20; j++) v = fmaf(s.w.Wm[j * 8 + p], s.dmf[note * 20 + j], v);
            a.dmelody[u * 448 + i] = v;
        }
        if (tid < 180) {
            float v = 0.f;
            if (tid < 160) {
                const int j = tid >> 3, p = tid & 7;
                for (int note = 0; note < 56; note++) v = fmaf(s.dmf[note * 20 + j], s.mel[note * 8 + p], v);
            } else {
                const int j = tid - 160;
                for (int note = 0; note < 56; note++) v += s.dmf[note * 20 + j];
            }
            s.gW[tid] += v;
        }
    }
    __syncthreads();
    if (cur_n >= 0)
        for (int i = tid; i < C * 450; i += blockDim.x)
            if (dzi_s[i] != 0.f) atomicAdd(a.dZi + (long long)cur_n * C * 450 + i, dzi_s[i]);
    for (int i = tid; i < 435; i += blockDim.x) {
        const float v = s.gW[i];
        if (v == 0.f) continue;
        if (i < 160) atomicAdd(a.gWm + i, v);
        else if (i < 180) atomicAdd(a.gbm + (i - 160), v);
        else if (i < 430) atomicAdd(a.gWl + (i - 180), v);
        else atomicAdd(a.gbl + (i - 430), v);
    }
}

// ------------------------------------------------------------------------------------------------
// UnpitchedStyleApplier tail backward.  Unit (n, b, t, f): dHpre[u][376] = (Wl^T dy) * leaky'(H); linear (2 x 8) grads.
// ------------------------------------------------------------------------------------------------
struct UsaBwdArgs {
    UsaArgs f;                 // H read, out unused
    const float *target;       // Ug (N,B,T,10,47,2)
    const float *coef;         // (N, kLossCoefs)
    float *dHpre;              // (N,B,T,10,376)
    float *gWl, *gbl;          // (2,8), (2)
};

__global__ void __launch_bounds__(64)
k_usa_bwd(UsaBwdArgs a) {
    __shared__ float g_s[18];
    const UsaArgs &f = a.f;
    if (threadIdx.x < 18) g_s[threadIdx.x] = 0.f;
    __syncthreads();
    const long long per_n = (long long)f.B * f.T * 10;
    const long long units = (long long)f.N * per_n;
    float gw[18];
#pragma unroll
    for (int i = 0; i < 18; i++) gw[i] = 0.f;
    for (long long u = blockIdx.x; u < units; u += gridDim.x) {
        const int n = (int)(u / per_n);
        if (threadIdx.x < 47) {
            const int note = threadIdx.x;
            const float *h = f.H + u * 376 + note * 8;
            float hv[8], y[2], t2[2], dpre[2];
#pragma unroll
            for (int q = 0; q < 8; q++) hv[q] = h[q];
#pragma unroll
            for (int e = 0; e < 2; e++) {
                float v = f.bl[e];
#pragma unroll
                for (int q = 0; q < 8; q++) v = fmaf(f.Wl[e * 8 + q], hv[q], v);
                y[e] = sigmoidf(v) * (e == 0 ? 6.f : 1.f);
                t2[e] = a.target[(u * 47 + note) * 2 + e];
            }
            loss_cell_grad<2>(y, t2, a.coef + n * kLossCoefs, dpre);
#pragma unroll
            for (int q = 0; q < 8; q++) {
                const float dh = f.Wl[q] * dpre[0] + f.Wl[8 + q] * dpre[1];
                a.dHpre[u * 376 + note * 8 + q] = dh * leaky_grad_from_out(hv[q]);
                gw[q] = fmaf(dpre[0], hv[q], gw[q]);
                gw[8 + q] = fmaf(dpre[1], hv[q], gw[8 + q]);
            }
            gw[16] += dpre[0]; gw[17] += dpre[1];
        }
    }
#pragma unroll
    for (int i = 0; i < 18; i++) {
        float v = gw[i];
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) atomicAdd(&g_s[i], v);
    }
    __syncthreads();
    if (threadIdx.x < 16) atomicAdd(a.gWl + threadIdx.x, g_s[threadIdx.x]);
    else if (threadIdx.x < 18) atomicAdd(a.gbl + (threadIdx.x - 16), g_s[threadIdx.x]);
}

// ------------------------------------------------------------------------------------------------
// MelodyEncoder tail backward.  melody = combine_c(x_c):  dx_c = dmel * w_c / S + (A_c - Bs) / (S w_c) * x_c with
// A_c = sum dmel . x_c over the whole channel, Bs = sum dmel . melody over the sample (see k_combine_bwd).
//   k_me_bwd_stats: A (N, C) and Bs (N) -> stats[n * (C + 1) + {c, C}]   (zeroed by the caller)
//   k_me_bwd:       per beat row (n, c, b, t): d od_pre[row][120] and the channels_linear / linear gradients.
// ------------------------------------------------------------------------------------------------
struct MeBwdArgs {
    MeArgs f;                   // melody read
    const float *dmelody;       // (N,B,T,10,56,8)
    float *stats;               // (N, C + 1)
    float *dod;                 // (N,C,B,T,120) pre-activation gradient, written
    float *gWc, *gbc, *gWl, *gbl;
};

// grid: persistent over (n, b, t) units
__global__ void __launch_bounds__(256)
k_me_bwd_stats(MeBwdArgs a) {
    extern __shared__ float dyn[];        // [C][120] od rows, [C] partial A
    __shared__ MeSmemW sw;
    __shared__ float tmp[8];
    const MeArgs &f = a.f;
    me_load_w(sw, f.w);
    float *od_s = dyn;
    const long long units = (long long)f.N * f.B * f.T;
    for (long long u = blockIdx.x; u < units; u += gridDim.x) {
        const int n = (int)(u / ((long long)f.B * f.T));
        const long long bt = u - (long long)n * f.B * f.T;
        __syncthreads();
        for (int i = threadIdx.x; i < f.C * 120; i += blockDim.x) {
            const int c = i / 120;
            od_s[i] = f.od[(((long long)n * f.C + c) * f.B * f.T + bt) * 120 + (i - c * 120)];
        }
        __syncthreads();
        float bs = 0.f;
        for (int c = -1; c < f.C; c++) {
            float s = 0.f;
            for (int cell = threadIdx.x; cell < 560; cell += blockDim.x) {
                const float *dm = a.dmelody + (u * 560 + cell) * 8;
                if (c < 0) {
                    const float *m = f.melody + (u * 560 + cell) * 8;
#pragma unroll
                    for (int e = 0; e < 8; e++) s = fmaf(dm[e], m[e], s);
                } else {
                    const float *Pc = f.P + ((((long long)n * f.C + c) * f.B * f.T + bt) * 560 + cell) * 5;
                    float p5[5], in15[15], x[8];
#pragma unroll
                    for (int p = 0; p < 5; p++) p5[p] = Pc[p];
                    me_cell(sw, od_s + c * 120, p5, cell % 56, in15, x);
#pragma unroll
                    for (int e = 0; e < 8; e++) s = fmaf(dm[e], x[e], s);
                }
            }
            s = block_sum(s, tmp);
            if (threadIdx.x == 0) atomicAdd(a.stats + n * (f.C + 1) + (c < 0 ? f.C : c), s);
        }
        (void)bs;
    }
}

struct MeBwdSmem {
    MeSmemW w;
    float od[120], dod[120];
    float gW[176];              // [0,120) Wl, [120,128) bl, [128,163) Wc, [163,170) bc
    float s_in[560 * 17];       // in15 ++ 1, padded to 17
    float s_dx[560 * 8];
};

// grid: persistent over (n, c, b, t) beat rows
__global__ void __launch_bounds__(256)
k_me_bwd(MeBwdArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    MeBwdSmem &s = *reinterpret_cast<MeBwdSmem *>(smem_raw);
    const MeArgs &f = a.f;
    const int tid = threadIdx.x;
    me_load_w(s.w, f.w);
    for (int i = tid; i < 176; i += blockDim.x) s.gW[i] = 0.f;
    float gbc[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};   // channels_linear bias gradient, reduced once at the end
    const long long BT = (long long)f.B * f.T;
    const long long rows = (long long)f.N * f.C * BT;
    for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
        const long long nc = row / BT, bt = row - nc * BT;
        const int n = (int)(nc / f.C), c = (int)(nc - (long long)n * f.C);
        __syncthreads();
        for (int i = tid; i < 120; i += blockDim.x) { s.od[i] = f.od[row * 120 + i]; s.dod[i] = 0.f; }
        // combine weights of this sample
        float wsum = 0.f;
        for (int k = 0; k < f.C; k++) wsum += sqrtf(1.f + f.ss[n * f.C + k]);
        const float wc = sqrtf(1.f + f.ss[n * f.C + c]);
        const float k1 = wc / wsum;
        const float k2 = (a.stats[n * (f.C + 1) + c] - a.stats[n * (f.C + 1) + f.C]) / (wsum * wc);
        __syncthreads();
        const float *Prow = f.P + row * 2800;
        const float *dmrow = a.dmelody + ((long long)n * BT + bt) * 560 * 8;
        for (int cell = tid; cell < 560; cell += blockDim.x) {
            const int note = cell % 56, o = note / 7, d = note - o * 7;
            float p5[5], in15[15], x[8], dx[8];
#pragma unroll
            for (int p = 0; p < 5; p++) p5[p] = Prow[cell * 5 + p];
            me_cell(s.w, s.od, p5, note, in15, x);
#pragma unroll
            for (int e = 0; e < 8; e++) {
                dx[e] = (dmrow[cell * 8 + e] * k1 + k2 * x[e]) * leaky_grad_from_out(x[e]);
                s.s_dx[cell * 8 + e] = dx[e];
            }
#pragma unroll
            for (int q = 0; q < 15; q++) {
                s.s_in[cell * 17 + q] = in15[q];
                float din = 0.f;
#pragma unroll
                for (int e = 0; e < 8; e++) din = fmaf(s.w.Wl[e * 15 + q], dx[e], din);
                din *= leaky_grad_from_out(in15[q]);
                if (q < 8) {
                    atomicAdd(&s.dod[o * 8 + q], din);
                    atomicAdd(&s.dod[64 + d * 8 + q], din);
                } else {
                    // channels_linear (5 -> 7): the input is zero for almost every cell
                    const int qq = q - 8;
                    gbc[qq] += din;
#pragma unroll
                    for (int p = 0; p < 5; p++) if (p5[p] != 0.f) atomicAdd(&s.gW[128 + qq * 5 + p], din * p5[p]);
                }
            }
            s.s_in[cell * 17 + 15] = 1.f;
        }
        __syncthreads();
        // linear (8 x 15) + bias: thread = (e, q in 0..15, half); q == 15 is the bias column
        {
            const int half = tid >> 7, eq = tid & 127, e = eq >> 4, q = eq & 15;
            float acc = 0.f;
            for (int cell = half * 280; cell < (half + 1) * 280; cell++) acc = fmaf(s.s_dx[cell * 8 + e], s.s_in[cell * 17 + q], acc);
            if (q < 15) atomicAdd(&s.gW[e * 15 + q], acc); else atomicAdd(&s.gW[120 + e], acc);
        }
        for (int i = tid; i < 120; i += blockDim.x) a.dod[row * 120 + i] = s.dod[i] * leaky_grad_from_out(s.od[i]);
    }
#pragma unroll
    for (int q = 0; q < 7; q++) {
        float v = gbc[q];
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((tid & 31) == 0) atomicAdd(&s.gW[163 + q], v);
    }
    __syncthreads();
    for (int i = tid; i < 170; i += blockDim.x) {
        const float v = s.gW[i];
        if (v == 0.f) continue;
        if (i < 120) atomicAdd(a.gWl + i, v);
        else if (i < 128) atomicAdd(a.gbl + (i - 120), v);
        else if (i < 163) atomicAdd(a.gWc + (i - 128), v);
        else atomicAdd(a.gbc + (i - 163), v);
    }
}

}  // namespace mstm
