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
// PitchedStyleApplier tail backward (pairs with k_psa_fwd2): one warp per (n, b, t, f) unit, lane = (channel, octave),
// loop over the 7 degrees.  Everything is recomputed from Zr / Zci / melody; the loss derivative is evaluated per cell.
//   * per-lane registers: the octave row (30), its gradient (30), the tensor-core accumulators (24);
//   * weight gradients are real contractions over cells and go to the tensor cores (mma.sync m16n8k8, 3xTF32 split so
//     the result is fp32-accurate):  gWl[:, :30] += a^T.dpre per 32-cell step, gWl[:, 30:] (+ bias) += mf^T.dps and
//     gWm += dmf^T.mel once per unit; operands are staged in shared memory in conflict-free layouts;
//   * the degree-row gradient (sum over the 8 octave lanes) is a shared-memory transposition; od is overwritten in
//     place by its own pre-activation gradient, then reduced to dZr (per unit) and dZi (CTA-level shared accumulators,
//     one global flush per CTA: a CTA never spans samples).
// ------------------------------------------------------------------------------------------------
constexpr int kPsaBwdWarps = 8;      // 8 x 24.7 KB of per-warp staging + C x 1.8 KB of dZi accumulators (<= 227 KB up to C = 14)
struct PsaBwdWarpSmem {
    float od[4 * 450];
    float snote[56 * 5];
    float mf[56 * 24];       // mf (col 20 = 1 for the bias row), later the pre-activation gradient dmf
    float mel[56 * 8];
    float dps[56 * 8];       // sum over channels of dpre per note; cols 5..7 stay 0
    float sa[32 * 36];       // [q][lane] staging (rows 30, 31 stay 0)
    float sd[32 * 8];        // dpre per lane; cols 5..7 stay 0
    float dzr[452];          // 450 used; keeps the per-warp stride a multiple of 16 bytes
};
static_assert(sizeof(PsaBwdWarpSmem) % 16 == 0, "per-warp shared block must stay 16-byte aligned");

__device__ __forceinline__ uint32_t to_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}
// D (16x8, fp32) += A (16x8) . B (8x8) with fp32 inputs split into tf32 hi + lo parts (3 MMAs: lo.hi + hi.lo + hi.hi)
__device__ __forceinline__ void mma_tf32x3(float *d, const float *a, const float *b) {
    uint32_t ah[4], al[4], bh[2], bl[2];
#pragma unroll
    for (int i = 0; i < 4; i++) { ah[i] = to_tf32(a[i]); al[i] = to_tf32(a[i] - __uint_as_float(ah[i])); }
#pragma unroll
    for (int i = 0; i < 2; i++) { bh[i] = to_tf32(b[i]); bl[i] = to_tf32(b[i] - __uint_as_float(bh[i])); }
#define MST_MMA(A, B)                                                                                              \
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};" \
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])                                                  \
                 : "r"(A[0]), "r"(A[1]), "r"(A[2]), "r"(A[3]), "r"(B[0]), "r"(B[1]))
    MST_MMA(al, bh);
    MST_MMA(ah, bl);
    MST_MMA(ah, bh);
#undef MST_MMA
}

struct PsaBwdArgs2 {
    const float *Zr, *Zci, *melody;   // as k_psa_fwd2
    const float *target;              // P (N,C,B,T,10,56,5)
    const float *coef;                // (N, kLossCoefs)
    float *dZr, *dZi;                 // (N,B,T,10,450) written; (N,C,450) atomically accumulated (zeroed by the caller)
    float *dmelody;                   // (N,B,T,10,56,8) written
    float *gWm, *gbm, *gWl, *gbl;     // parameter gradients (atomic accumulate)
    int N, C, B, T, ctas_per_n, warps;
};

__global__ void __launch_bounds__(kPsaBwdWarps * 32)
k_psa_bwd2(PsaBwdArgs2 a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    PsaBwdWarpSmem &s = reinterpret_cast<PsaBwdWarpSmem *>(smem_raw)[warp];
    float *dzi_s = reinterpret_cast<float *>(smem_raw + sizeof(PsaBwdWarpSmem) * a.warps);   // [C][450]
    const int C = a.C;
    const int n = blockIdx.x / a.ctas_per_n, kcta = blockIdx.x - n * a.ctas_per_n;
    const long long per_n = (long long)a.B * a.T * 10;
    const long long chunk = (per_n + a.ctas_per_n - 1) / a.ctas_per_n;
    const long long b_lo = (long long)kcta * chunk, b_hi = min(per_n, b_lo + chunk);
    for (int i = threadIdx.x; i < C * 450; i += blockDim.x) dzi_s[i] = 0.f;
    for (int i = lane; i < 32 * 36; i += 32) s.sa[i] = 0.f;
    for (int i = lane; i < 32 * 8; i += 32) s.sd[i] = 0.f;
    for (int i = lane; i < 56 * 8; i += 32) s.dps[i] = 0.f;
    for (int i = lane; i < 56 * 24; i += 32) s.mf[i] = 0.f;
    __syncthreads();
    float cf[6];
#pragma unroll
    for (int i = 0; i < 6; i++) cf[i] = a.coef[n * kLossCoefs + i];
    const int c_l = lane >> 3, o = lane & 7, g = lane >> 2, t = lane & 3;
    float accW[2][4], accM[2][4], accWm[2][4], gbm[20];
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) { accW[i][j] = 0.f; accM[i][j] = 0.f; accWm[i][j] = 0.f; }
#pragma unroll
    for (int j = 0; j < 20; j++) gbm[j] = 0.f;

    for (long long btf = b_lo + warp; btf < b_hi; btf += a.warps) {
        const long long u = (long long)n * per_n + btf;
        __syncwarp();
        // ---- phase 1: per-note melody branch (mf kept for the backward), zero the per-unit accumulators
        for (int note = lane; note < 56; note += 32) {
            const float4 *mp = reinterpret_cast<const float4 *>(a.melody + (u * 56 + note) * 8);
            const float4 m0 = mp[0], m1 = mp[1];
            const float mel8[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
            float mf[20], sn[5];
            psa_note(mel8, mf, sn);
#pragma unroll
            for (int e = 0; e < 5; e++) { s.snote[note * 5 + e] = sn[e]; s.dps[note * 8 + e] = 0.f; }
#pragma unroll
            for (int p = 0; p < 8; p++) s.mel[note * 8 + p] = mel8[p];
#pragma unroll
            for (int j = 0; j < 20; j++) s.mf[note * 24 + j] = mf[j];
            s.mf[note * 24 + 20] = 1.f;
        }
        for (int j = lane; j < 450; j += 32) s.dzr[j] = 0.f;
        const float *zr = a.Zr + u * 450;
        for (int c0 = 0; c0 < C; c0 += 4) {
            const int cg = min(4, C - c0);
            __syncwarp();
            for (int j = lane; j < 450; j += 32) {
                const float z = zr[j];
                for (int c = 0; c < cg; c++) s.od[c * 450 + j] = leaky(z + a.Zci[((long long)n * C + c0 + c) * 450 + j]);
            }
            const bool active = c_l < cg;
            if (!active) {
#pragma unroll
                for (int q = 0; q < 30; q++) s.sa[q * 36 + lane] = 0.f;
#pragma unroll
                for (int e = 0; e < 5; e++) s.sd[lane * 8 + e] = 0.f;
            }
            __syncwarp();
            float oct[30], doct[30];
#pragma unroll
            for (int q = 0; q < 30; q++) { oct[q] = active ? s.od[c_l * 450 + o * 30 + q] : 0.f; doct[q] = 0.f; }
            const float *deg = s.od + c_l * 450 + 240;
            const float *tg = a.target + ((((long long)n * C + c0 + c_l) * per_n) + btf) * 280;
            for (int d = 0; d < 7; d++) {
                const int note = o * 7 + d;
                float dpre[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
                if (active) {
                    float y[5], t5[5];
#pragma unroll
                    for (int e = 0; e < 5; e++) { y[e] = s.snote[note * 5 + e]; t5[e] = tg[note * 5 + e]; }
#pragma unroll
                    for (int q = 0; q < 30; q++) {
                        const float v = leaky(oct[q] + deg[d * 30 + q]);
                        s.sa[q * 36 + lane] = v;
#pragma unroll
                        for (int e = 0; e < 5; e++) y[e] = fmaf(c_psa[kPsaWl + e * 50 + q], v, y[e]);
                    }
#pragma unroll
                    for (int e = 0; e < 5; e++) y[e] = sigmoidf(y[e]) * (e == 0 ? 6.f : 1.f);
                    loss_cell_grad<5>(y, t5, cf, dpre);
#pragma unroll
                    for (int e = 0; e < 5; e++) s.sd[lane * 8 + e] = dpre[e];
                }
                // dps[note] += sum over the channel lanes (lane = c * 8 + o: xor 8, 16)
#pragma unroll
                for (int e = 0; e < 5; e++) {
                    float v = dpre[e];
                    v += __shfl_xor_sync(0xffffffffu, v, 8);
                    v += __shfl_xor_sync(0xffffffffu, v, 16);
                    if (c_l == 0) s.dps[note * 8 + e] += v;
                }
                __syncwarp();
                // gWl[:, :30] += a^T . dpre over the 32 cells of this step (tensor cores)
#pragma unroll
                for (int ks = 0; ks < 4; ks++) {
                    const float bf[2] = {s.sd[(ks * 8 + t) * 8 + g], s.sd[(ks * 8 + t + 4) * 8 + g]};
#pragma unroll
                    for (int mt = 0; mt < 2; mt++) {
                        const float *r0 = s.sa + (mt * 16 + g) * 36 + ks * 8 + t;
                        const float af[4] = {r0[0], r0[8 * 36], r0[4], r0[8 * 36 + 4]};
                        mma_tf32x3(accW[mt], af, bf);
                    }
                }
                __syncwarp();
                // input gradient of the outer add: din[q] = (Wl[:, q] . dpre) * leaky'(a[q])
                if (active) {
#pragma unroll
                    for (int q = 0; q < 30; q++) {
                        float din = 0.f;
#pragma unroll
                        for (int e = 0; e < 5; e++) din = fmaf(c_psa[kPsaWl + e * 50 + q], dpre[e], din);
                        din *= leaky_grad_from_out(s.sa[q * 36 + lane]);
                        doct[q] += din;
                        s.sa[q * 36 + lane] = din;
                    }
                }
                __syncwarp();
                // degree-row gradient: sum over the 8 octave lanes of each channel, written over od in place
                for (int i = lane; i < cg * 30; i += 32) {
                    const int c = i / 30, q = i - c * 30;
                    const float4 v0 = *reinterpret_cast<const float4 *>(s.sa + q * 36 + c * 8);
                    const float4 v1 = *reinterpret_cast<const float4 *>(s.sa + q * 36 + c * 8 + 4);
                    const float sum = ((v0.x + v0.y) + (v0.z + v0.w)) + ((v1.x + v1.y) + (v1.z + v1.w));
                    float *p = s.od + c * 450 + 240 + d * 30 + q;
                    *p = sum * leaky_grad_from_out(*p);
                }
                __syncwarp();
            }
            if (active) {
#pragma unroll
                for (int q = 0; q < 30; q++) s.od[c_l * 450 + o * 30 + q] = doct[q] * leaky_grad_from_out(oct[q]);
            }
            __syncwarp();
            for (int j = lane; j < 450; j += 32) {
                float sum = 0.f;
                for (int c = 0; c < cg; c++) {
                    const float v = s.od[c * 450 + j];
                    sum += v;
                    atomicAdd(&dzi_s[(c0 + c) * 450 + j], v);
                }
                s.dzr[j] += sum;
            }
        }
        __syncwarp();
        for (int j = lane; j < 450; j += 32) a.dZr[u * 450 + j] = s.dzr[j];
        // ---- melody branch.  gWl[:, 30:50] (+ bias through the all-ones column 20) += mf^T . dps
#pragma unroll
        for (int ks = 0; ks < 7; ks++) {
            const float bf[2] = {s.dps[(ks * 8 + t) * 8 + g], s.dps[(ks * 8 + t + 4) * 8 + g]};
            const float *r0 = s.mf + (ks * 8 + t) * 24 + g;
            const float a0[4] = {r0[0], r0[8], r0[4 * 24], r0[4 * 24 + 8]};
            const float a1[4] = {r0[16], 0.f, r0[4 * 24 + 16], 0.f};
            mma_tf32x3(accM[0], a0, bf);
            mma_tf32x3(accM[1], a1, bf);
        }
        __syncwarp();
        for (int note = lane; note < 56; note += 32) {
            float dmf[20], dp[5];
#pragma unroll
            for (int e = 0; e < 5; e++) dp[e] = s.dps[note * 8 + e];
#pragma unroll
            for (int j = 0; j < 20; j++) {
                float v = 0.f;
#pragma unroll
                for (int e = 0; e < 5; e++) v = fmaf(c_psa[kPsaWl + e * 50 + 30 + j], dp[e], v);
                v *= leaky_grad_from_out(s.mf[note * 24 + j]);
                dmf[j] = v;
                gbm[j] += v;
                s.mf[note * 24 + j] = v;
            }
            float dm[8];
#pragma unroll
            for (int p = 0; p < 8; p++) {
                float v = 0.f;
#pragma unroll
                for (int j = 0; j < 20; j++) v = fmaf(c_psa[kPsaWm + j * 8 + p], dmf[j], v);
                dm[p] = v;
            }
            float4 *dst = reinterpret_cast<float4 *>(a.dmelody + (u * 56 + note) * 8);
            dst[0] = make_float4(dm[0], dm[1], dm[2], dm[3]);
            dst[1] = make_float4(dm[4], dm[5], dm[6], dm[7]);
        }
        __syncwarp();
        // gWm += dmf^T . mel
#pragma unroll
        for (int ks = 0; ks < 7; ks++) {
            const float bf[2] = {s.mel[(ks * 8 + t) * 8 + g], s.mel[(ks * 8 + t + 4) * 8 + g]};
            const float *r0 = s.mf + (ks * 8 + t) * 24 + g;
            const float a0[4] = {r0[0], r0[8], r0[4 * 24], r0[4 * 24 + 8]};
            const float a1[4] = {r0[16], 0.f, r0[4 * 24 + 16], 0.f};
            mma_tf32x3(accWm[0], a0, bf);
            mma_tf32x3(accWm[1], a1, bf);
        }
    }
    // ---- flush the per-lane accumulators (fragment layout: rows g / g + 8, columns 2t / 2t + 1)
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int row = mt * 16 + g + (i >> 1) * 8, col = 2 * t + (i & 1);
            if (col < 5) {
                if (row < 30 && accW[mt][i] != 0.f) atomicAdd(a.gWl + col * 50 + row, accW[mt][i]);
                if (row < 20 && accM[mt][i] != 0.f) atomicAdd(a.gWl + col * 50 + 30 + row, accM[mt][i]);
                if (row == 20 && accM[mt][i] != 0.f) atomicAdd(a.gbl + col, accM[mt][i]);
            }
            if (row < 20 && accWm[mt][i] != 0.f) atomicAdd(a.gWm + row * 8 + col, accWm[mt][i]);
        }
#pragma unroll
    for (int j = 0; j < 20; j++) {
        float v = gbm[j];
        for (int sft = 16; sft; sft >>= 1) v += __shfl_xor_sync(0xffffffffu, v, sft);
        if (lane == 0 && v != 0.f) atomicAdd(a.gbm + j, v);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < C * 450; i += blockDim.x)
        if (dzi_s[i] != 0.f) atomicAdd(a.dZi + (long long)n * C * 450 + i, dzi_s[i]);
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

}  // namespace mstm
