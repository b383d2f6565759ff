// mst_model_loss.cuh — get_total_loss (style/model.py:836-997) as two kernels per grid:
//   k_loss_sums   one pass over prediction + target producing the 7 masked sums of get_channels_losses (:909-921)
//                 per sample (TP/FP/FN of the smooth F-score :863-883, mask count, velocity / duration / accidental sums);
//   k_loss_tree   the scalar tree (safe_div, quadratic means with gradient-carrying weights :924-932, tanh
//                 normalisation :957-975, song-info losses :899-906) evaluated once per sample in f64, together with
//                 its reverse-mode derivative: one coefficient per masked sum, which the fused tail backward kernels
//                 turn into per-cell gradients, plus the gradients of the three song-info heads.
// Every reduction is segmented per sample: a batch is N independent batch-1 samples (the reference only runs N = 1).
#pragma once
#include "mst_model_kernels2.cuh"

namespace mstm {

constexpr int kLossSums = 8;     // tp, fp, fn, n_mask, sum_vel, sum_dur, sum_acc, (unused)
constexpr int kLossLeaves = 16;  // see enum below
constexpr int kLossCoefs = 8;    // c_tp, c_fp, c_fn, c_sv, c_sd, c_sa

enum LossLeaf {
    LL_TOTAL = 0, LL_CHANNELS, LL_P_TOTAL, LL_P_NOTES, LL_P_VELOCITY, LL_P_DURATION, LL_P_ACCIDENTALS,
    LL_U_TOTAL, LL_U_NOTES, LL_U_VELOCITY, LL_U_DURATION, LL_SONG_TOTAL, LL_INSTRUMENTS, LL_MODE, LL_BPM
};

__device__ __forceinline__ double block_sum_d(double v, double *tmp) {
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) tmp[threadIdx.x >> 5] = v;
    __syncthreads();
    double r = 0.0;
    if (threadIdx.x < 32) {
        r = threadIdx.x < (blockDim.x >> 5) ? tmp[threadIdx.x] : 0.0;
        for (int o = 16; o; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
    }
    __syncthreads();
    return r;
}

// F.binary_cross_entropy element (torch clamps both logs at -100)
__device__ __forceinline__ float bce_elem(float p, float t) {
    const float lp = fmaxf(logf(p), -100.f), l1p = fmaxf(log1pf(-p), -100.f);
    return (t - 1.f) * l1p - t * lp;
}

struct LossSumsArgs {
    const float *pred, *target;   // (N, cells, FEATS)
    long long cells;              // per sample
    double *sums;                 // (N, kLossSums), zeroed by the caller
};

// grid: (chunks, N)
template <int FEATS>
__global__ void __launch_bounds__(256)
k_loss_sums(LossSumsArgs a) {
    __shared__ double tmp[8];
    const int n = blockIdx.y;
    const float *pr = a.pred + (long long)n * a.cells * FEATS, *tg = a.target + (long long)n * a.cells * FEATS;
    double acc[7];
#pragma unroll
    for (int i = 0; i < 7; i++) acc[i] = 0.0;
    // float partials over short runs, folded into doubles
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long base = (long long)blockIdx.x * blockDim.x + threadIdx.x; base < a.cells; base += stride * 32) {
        float s[7];
#pragma unroll
        for (int i = 0; i < 7; i++) s[i] = 0.f;
        for (int it = 0; it < 32; it++) {
            const long long cell = base + it * stride;
            if (cell >= a.cells) break;
            const float *p = pr + cell * FEATS, *t = tg + cell * FEATS;
            const float v = p[1], tv = t[1];
            s[0] += fminf(v, tv);
            s[1] += fmaxf(v - tv, 0.f);
            s[2] += fmaxf(tv - v, 0.f);
            if (tv > 0.f) {
                s[3] += 1.f;
                const float dv = tv - v;
                s[4] = fmaf(dv, dv, s[4]);
                const float dd = (p[0] - fminf(t[0], 6.f)) / 6.f;
                s[5] = fmaf(dd, dd, s[5]);
                if (FEATS == 5) s[6] += bce_elem(p[2], t[2]) + bce_elem(p[3], t[3]) + bce_elem(p[4], t[4]);
            }
        }
#pragma unroll
        for (int i = 0; i < 7; i++) acc[i] += (double)s[i];
    }
#pragma unroll
    for (int i = 0; i < 7; i++) {
        const double r = block_sum_d(acc[i], tmp);
        if (threadIdx.x == 0 && r != 0.0) atomicAdd(a.sums + n * kLossSums + i, r);
    }
}

// ------------------------------------------------------------------------------------------------ scalar tree
struct SafeDiv { double val, d_num, d_den; };
__device__ __forceinline__ SafeDiv safe_div_d(double num, double den) {   // model.py:854-860
    if (fabs(den) < 1e-7) den = den < 0 ? den - 1e-7 : den + 1e-7;
    return SafeDiv{num / den, 1.0 / den, -num / (den * den)};
}

struct ChannelTree {
    double total, notes, velocity, duration, accidentals;    // leaves (accidentals after the optional tanh)
    double c_tp, c_fp, c_fn, c_sv, c_sd, c_sa;               // d total / d sums
};

// get_channels_losses + combine_losses for one grid of one sample
__device__ inline ChannelTree channel_tree(const double *s, bool pitched, bool normalize) {
    ChannelTree r{};
    const double tp = s[0], fp = s[1], fn = s[2], n = s[3];
    const SafeDiv P = safe_div_d(tp, tp + fp), R = safe_div_d(tp, tp + fn);
    const SafeDiv Fd = safe_div_d(P.val * R.val, P.val + R.val);
    r.notes = 1.0 - 2.0 * Fd.val;
    r.velocity = s[4] / n;
    r.duration = s[5] / n;
    const double acc_raw = pitched ? s[6] / (3.0 * n) : 0.0;
    r.accidentals = pitched ? (normalize ? tanh(acc_raw) : acc_raw) : 0.0;
    // notes2 = sqrt(notes * notes^2 + (1 - notes) * velocity^2)     (weights carry gradient)
    const double g = r.notes * r.notes * r.notes + (1.0 - r.notes) * r.velocity * r.velocity;
    const double notes2 = g == 0.0 ? 0.0 : sqrt(g);
    const double k = pitched ? 3.0 : 2.0;
    const double q = (r.duration * r.duration + r.accidentals * r.accidentals + g) / k;
    r.total = q == 0.0 ? 0.0 : sqrt(q);
    // reverse mode, seed d total = 1
    const double inv = r.total == 0.0 ? 0.0 : 1.0 / (k * r.total);
    const double d_dur = r.duration * inv, d_acc = r.accidentals * inv;
    const double d_notes2 = notes2 * inv;
    const double d_g = notes2 == 0.0 ? 0.0 : d_notes2 / (2.0 * notes2);
    const double d_notes = d_g * (3.0 * r.notes * r.notes - r.velocity * r.velocity);
    const double d_vel = d_g * 2.0 * (1.0 - r.notes) * r.velocity;
    const double d_f = -2.0 * d_notes;                      // notes = 1 - 2 * Fd
    const double d_pr = d_f * Fd.d_num, d_ps = d_f * Fd.d_den;
    const double d_P = d_pr * R.val + d_ps, d_R = d_pr * P.val + d_ps;
    r.c_tp = d_P * (P.d_num + P.d_den) + d_R * (R.d_num + R.d_den);
    r.c_fp = d_P * P.d_den;
    r.c_fn = d_R * R.d_den;
    r.c_sv = d_vel / n;
    r.c_sd = d_dur / n;
    const double d_acc_raw = pitched ? d_acc * (normalize ? 1.0 - r.accidentals * r.accidentals : 1.0) : 0.0;
    r.c_sa = pitched ? d_acc_raw / (3.0 * n) : 0.0;
    return r;
}

struct LossTreeArgs {
    const double *sums_p, *sums_u;          // (N, kLossSums); sums_u null without drums
    const float *instr, *instr_target;      // (N, 41) logits / multi-hot
    const float *mode, *mode_target;        // (N, 2) logits / one-hot
    const float *bpm, *bpm_target;          // (N) prediction after the sigmoid head / target
    float *leaves;                          // (N, kLossLeaves)
    float *coef_p, *coef_u;                 // (N, kLossCoefs)   may be null (loss only)
    float *d_instr, *d_mode, *d_bpm_pre;    // gradients w.r.t. logits / pre-sigmoid bpm; may be null
    int N, normalize;
    float seed;                             // d(objective) / d(total_n): 1 = sum over samples
};

__global__ void k_loss_tree(LossTreeArgs a) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= a.N) return;
    const bool norm = a.normalize != 0;
    const ChannelTree p = channel_tree(a.sums_p + n * kLossSums, true, norm);
    ChannelTree u{};
    double channels, d_p, d_u = 0.0;
    if (a.sums_u) {
        u = channel_tree(a.sums_u + n * kLossSums, false, norm);
        const double q = (p.total * p.total + u.total * u.total) / 2.0;
        channels = q == 0.0 ? 0.0 : sqrt(q);
        d_p = channels == 0.0 ? 0.0 : p.total / (2.0 * channels);
        d_u = channels == 0.0 ? 0.0 : u.total / (2.0 * channels);
    } else {
        channels = p.total; d_p = 1.0;
    }
    // song info (model.py:899-906); F.binary_cross_entropy_with_logits, mean over the 41 logits
    double instr = 0.0;
    for (int j = 0; j < 41; j++) {
        const double x = a.instr[n * 41 + j], t = a.instr_target[n * 41 + j];
        instr += fmax(x, 0.0) - x * t + log1p(exp(-fabs(x)));
    }
    instr /= 41.0;
    const double m0 = a.mode[n * 2], m1 = a.mode[n * 2 + 1];
    const int tgt = a.mode_target[n * 2 + 1] > a.mode_target[n * 2] ? 1 : 0;     // argmax, first index on ties
    const double mx = fmax(m0, m1), lse = mx + log(exp(m0 - mx) + exp(m1 - mx));
    double mode = lse - (tgt ? m1 : m0);
    const double bd = ((double)a.bpm[n] - (double)a.bpm_target[n]) / 150.0;
    const double bpm = bd * bd;
    const double instr_raw = instr, mode_raw = mode;
    if (norm) { instr = tanh(instr); mode = tanh(mode); }
    const double qs = (instr * instr + mode * mode + bpm * bpm) / 3.0;
    const double song = qs == 0.0 ? 0.0 : sqrt(qs);
    const double qt = (channels * channels + song * song) / 2.0;
    const double total = qt == 0.0 ? 0.0 : sqrt(qt);
    float *L = a.leaves + n * kLossLeaves;
    L[LL_TOTAL] = (float)total; L[LL_CHANNELS] = (float)channels;
    L[LL_P_TOTAL] = (float)p.total; L[LL_P_NOTES] = (float)p.notes; L[LL_P_VELOCITY] = (float)p.velocity;
    L[LL_P_DURATION] = (float)p.duration; L[LL_P_ACCIDENTALS] = (float)p.accidentals;
    L[LL_U_TOTAL] = (float)u.total; L[LL_U_NOTES] = (float)u.notes; L[LL_U_VELOCITY] = (float)u.velocity;
    L[LL_U_DURATION] = (float)u.duration; L[LL_SONG_TOTAL] = (float)song; L[LL_INSTRUMENTS] = (float)instr;
    L[LL_MODE] = (float)mode; L[LL_BPM] = (float)bpm; L[15] = 0.f;
    (void)instr_raw; (void)mode_raw;
    if (!a.coef_p) return;
    const double seed = a.seed;
    const double d_channels = total == 0.0 ? 0.0 : seed * channels / (2.0 * total);
    const double d_song = total == 0.0 ? 0.0 : seed * song / (2.0 * total);
    const double gp = d_channels * d_p, gu = d_channels * d_u;
    float *cp = a.coef_p + n * kLossCoefs;
    cp[0] = (float)(gp * p.c_tp); cp[1] = (float)(gp * p.c_fp); cp[2] = (float)(gp * p.c_fn);
    cp[3] = (float)(gp * p.c_sv); cp[4] = (float)(gp * p.c_sd); cp[5] = (float)(gp * p.c_sa); cp[6] = cp[7] = 0.f;
    if (a.coef_u) {
        float *cu = a.coef_u + n * kLossCoefs;
        cu[0] = (float)(gu * u.c_tp); cu[1] = (float)(gu * u.c_fp); cu[2] = (float)(gu * u.c_fn);
        cu[3] = (float)(gu * u.c_sv); cu[4] = (float)(gu * u.c_sd); cu[5] = 0.f; cu[6] = cu[7] = 0.f;
    }
    const double inv_s = song == 0.0 ? 0.0 : d_song / (3.0 * song);
    const double d_instr = instr * inv_s * (norm ? 1.0 - instr * instr : 1.0);
    const double d_mode = mode * inv_s * (norm ? 1.0 - mode * mode : 1.0);
    const double d_bpm = bpm * inv_s;
    for (int j = 0; j < 41; j++) {
        const double x = a.instr[n * 41 + j], t = a.instr_target[n * 41 + j];
        a.d_instr[n * 41 + j] = (float)(d_instr * (1.0 / (1.0 + exp(-x)) - t) / 41.0);
    }
    const double s0 = exp(m0 - lse), s1 = exp(m1 - lse);
    a.d_mode[n * 2] = (float)(d_mode * (s0 - (tgt == 0 ? 1.0 : 0.0)));
    a.d_mode[n * 2 + 1] = (float)(d_mode * (s1 - (tgt == 1 ? 1.0 : 0.0)));
    // bpm = sigmoid(pre) * 150 + 50 (model.py:552-554)
    const double sg = ((double)a.bpm[n] - 50.0) / 150.0;
    a.d_bpm_pre[n] = (float)(d_bpm * 2.0 * bd / 150.0 * 150.0 * sg * (1.0 - sg));
}

// per-cell gradient w.r.t. the PRE-sigmoid outputs of a style applier, from post-activation y, target t and the
// per-sample coefficients (torch: minimum splits ties 0.5/0.5, relu' (0) = 0, BCE backward divides by max(p(1-p), 1e-12))
template <int FEATS>
__device__ __forceinline__ void loss_cell_grad(const float *y, const float *t, const float *cf, float *dpre) {
    const float v = y[1], tv = t[1];
    const bool mask = tv > 0.f;
    float dv = cf[0] * (v < tv ? 1.f : (v == tv ? 0.5f : 0.f)) + (v > tv ? cf[1] : 0.f) - (tv > v ? cf[2] : 0.f);
    if (mask) dv += cf[3] * 2.f * (v - tv);
    dpre[1] = dv * v * (1.f - v);
    if (mask) {
        const float y0 = y[0];
        dpre[0] = cf[4] * 2.f * (y0 - fminf(t[0], 6.f)) / 36.f * y0 * (1.f - y0 / 6.f);
        if (FEATS == 5) {
#pragma unroll
            for (int e = 2; e < 5; e++) {
                const float p = y[e], pq = p * (1.f - p);
                dpre[e] = cf[5] * (p - t[e]) / fmaxf(pq, 1e-12f) * pq;
            }
        }
    } else {
#pragma unroll
        for (int e = 0; e < FEATS; e++) if (e != 1) dpre[e] = 0.f;
    }
}

}  // namespace mstm
