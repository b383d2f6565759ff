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
// PitchedStyleApplier tail backward (pairs with k_psa_fwd2), two kernels.
//
// k_psa_bwd3 — one warp per (n, b, t, f) unit, lane = (channel, octave), loop over the 7 degrees; everything is
// recomputed from Zr / Zci / melody.  The loss derivative (mst_model_loss.cuh: loss_cell_grad) reaches the duration and
// accidental outputs only where the TARGET holds a note (velocity > 0: 0.1-0.6 % of the cells); everywhere else only
// the velocity output carries gradient.  So:
//   * dense part, every cell: velocity row of the final linear only (30 FFMA forward, 30 backward), its weight
//     gradient gWl[1, :30] in 30 per-lane registers;
//   * cells with a target note (found by ballot, handled one at a time by the WHOLE warp, lane = column q of the
//     30-wide row): the other four outputs, their loss derivative, their share of the input gradient and
//     gWl[{0,2,3,4}, q] in 4 per-lane registers.  Exact: nothing is dropped, only zeros are skipped;
//   * the degree-row gradient (sum over the 8 octave lanes) is a shared-memory transposition; od is overwritten in
//     place by its own pre-activation gradient, then reduced to dZr (per unit) and dZi (CTA-level shared accumulators,
//     one global flush per CTA: a CTA never spans samples);
//   * dps[note][e] = sum over channels of the pre-sigmoid gradient goes to global for k_psa_mel_bwd.
// k_psa_mel_bwd — the melody branch (melody_linear 8 -> 20, its 20 columns of the final linear): weight gradients on the
// tensor cores (mma.sync m16n8k8, 3xTF32 split = fp32-accurate), dmelody written per note.
// ------------------------------------------------------------------------------------------------
constexpr int kPsaBwdWarps = 12;
struct PsaBwdWarpSmem {
    float od[4 * 450];
    float snote[56 * 5];
    float dps[56 * 5];
    float sa[32 * 34];       // [lane][q] input-gradient staging of the current degree step (row stride 34: conflict-free 8-byte stores)
    float exq[32];           // a target-note cell's extra input gradient, handed back to its owner lane
    float mel[56 * 8];       // this unit's melody rows / Zr row, fetched with cp.async during the previous unit
    float zr[452];
};
static_assert(sizeof(PsaBwdWarpSmem) % 16 == 0, "per-warp shared block must stay 16-byte aligned");

__device__ __forceinline__ uint32_t to_tf32(float x) { return __float_as_uint(tf32_rna(x)); }
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

struct PsaBwdArgs3 {
    const float *Zr, *Zci, *melody;   // as k_psa_fwd2
    const float *target;              // P (N,C,B,T,10,56,5)
    const float *coef;                // (N, kLossCoefs)
    const float *Wl;                  // final linear (5, 50), global copy for per-lane column access
    float *dZr, *dZi;                 // (N,B,T,10,450) written; (N,C,450) atomically accumulated (zeroed by the caller)
    float *dps;                       // per unit (56, 5) at stride dps_stride: sum over channels of the pre-sigmoid gradient
    long long dps_stride;
    float *gWl;                       // (5, 50): columns 0..29 accumulated here
    int N, C, B, T, ctas_per_n, warps;
};

__global__ void __launch_bounds__(kPsaBwdWarps * 32, 1)
k_psa_bwd3(PsaBwdArgs3 a) {
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
    for (int i = lane; i < 32 * 34; i += 32) s.sa[i] = 0.f;
    __syncthreads();
    float cf[6];
#pragma unroll
    for (int i = 0; i < 6; i++) cf[i] = a.coef[n * kLossCoefs + i];
    const int c_l = lane >> 3, o = lane & 7;
    // lane = column q of the final linear for the target-note cells: its four non-velocity weights and their gradients
    const int qx = min(lane, 29);
    const float wx0 = a.Wl[0 * 50 + qx], wx2 = a.Wl[2 * 50 + qx], wx3 = a.Wl[3 * 50 + qx], wx4 = a.Wl[4 * 50 + qx];
    float gwx0 = 0.f, gwx2 = 0.f, gwx3 = 0.f, gwx4 = 0.f;
    f32x2 gw1[15];                                    // velocity-row weight gradient, columns (2 qp, 2 qp + 1)
#pragma unroll
    for (int qp = 0; qp < 15; qp++) gw1[qp] = pk2(0.f, 0.f);

    auto prefetch = [&](long long uu) {           // Zr rows are 1800 bytes: 8-byte aligned
        const float *mg = a.melody + uu * 448, *zg = a.Zr + uu * 450;
        for (int i = lane; i < 112; i += 32) cp_async16(s.mel + i * 4, mg + i * 4);
        for (int i = lane; i < 225; i += 32) cp_async8(s.zr + i * 2, zg + i * 2);
        asm volatile("cp.async.commit_group;");
    };
    if (b_lo + warp < b_hi) prefetch((long long)n * per_n + b_lo + warp);
    for (long long btf = b_lo + warp; btf < b_hi; btf += a.warps) {
        const long long u = (long long)n * per_n + btf;
        asm volatile("cp.async.wait_group 0;");
        __syncwarp();
        // ---- per-note melody branch of the forward (bl + Wl[:, 30:].mf), zero the per-unit accumulator
        for (int note = lane; note < 56; note += 32) {
            const float4 *mp = reinterpret_cast<const float4 *>(s.mel + note * 8);
            const float4 m0 = mp[0], m1 = mp[1];
            const f32x2 mel[4] = {pk2(m0.x, m0.y), pk2(m0.z, m0.w), pk2(m1.x, m1.y), pk2(m1.z, m1.w)};
            f32x2 mf2[10];
            float sn[5];
            psa_note2(mel, mf2, sn);
#pragma unroll
            for (int e = 0; e < 5; e++) { s.snote[note * 5 + e] = sn[e]; s.dps[note * 5 + e] = 0.f; }
        }
        float *dzr = a.dZr + u * 450;
        for (int c0 = 0; c0 < C; c0 += 4) {
            const int cg = min(4, C - c0);
            __syncwarp();
            psa_build_od(s.od, s.zr, a.Zci + ((long long)n * C + c0) * 450, cg, lane);
            if (c0 + 4 >= C && btf + a.warps < b_hi) {       // both staging rows are consumed: fetch the next unit's
                __syncwarp();
                prefetch(u + a.warps);
            }
            const bool active = c_l < cg;
            if (!active) {
#pragma unroll
                for (int qp = 0; qp < 15; qp++) *reinterpret_cast<f32x2 *>(s.sa + lane * 34 + qp * 2) = pk2(0.f, 0.f);
            }
            __syncwarp();
            f32x2 oct[15], doct[15];
#pragma unroll
            for (int qp = 0; qp < 15; qp++) {
                oct[qp] = active ? ld2(s.od + c_l * 450 + o * 30 + qp * 2) : pk2(0.f, 0.f);
                doct[qp] = pk2(0.f, 0.f);
            }
            const float *deg = s.od + c_l * 450 + 240;
            const float *tg = a.target + ((((long long)n * C + c0 + c_l) * per_n) + btf) * 280;
            float tvel[7];                                   // target velocities of this lane's 7 cells, fetched up front
#pragma unroll
            for (int d = 0; d < 7; d++) tvel[d] = active ? tg[(o * 7 + d) * 5 + 1] : 0.f;
#pragma unroll 1
            for (int d = 0; d < 7; d++) {
                const int note = o * 7 + d;
                float dpre1 = 0.f;
                bool has_note = false;
                if (active) {
                    // dense part: velocity output only (three partial sums keep the FMA pipe busy)
                    f32x2 av[15];
                    f32x2 ya = pk2(s.snote[note * 5 + 1], 0.f), yb = pk2(0.f, 0.f), yc = pk2(0.f, 0.f);
#pragma unroll
                    for (int qp = 0; qp < 15; qp += 3) {
                        av[qp] = leaky2(add2(oct[qp], ld2(deg + d * 30 + qp * 2)));
                        av[qp + 1] = leaky2(add2(oct[qp + 1], ld2(deg + d * 30 + qp * 2 + 2)));
                        av[qp + 2] = leaky2(add2(oct[qp + 2], ld2(deg + d * 30 + qp * 2 + 4)));
                        ya = fma2(MST_CPSA2(kPsaWl + 50 + qp * 2), av[qp], ya);
                        yb = fma2(MST_CPSA2(kPsaWl + 50 + qp * 2 + 2), av[qp + 1], yb);
                        yc = fma2(MST_CPSA2(kPsaWl + 50 + qp * 2 + 4), av[qp + 2], yc);
                    }
                    const float y1 = sum2(add2(ya, add2(yb, yc)));
                    float tv = tvel[0];
#pragma unroll
                    for (int dd = 1; dd < 7; dd++) tv = d == dd ? tvel[dd] : tv;
                    const float p = sigmoid_fast(y1);
                    has_note = tv > 0.f;
                    float dv = cf[0] * (p < tv ? 1.f : (p == tv ? 0.5f : 0.f)) + (p > tv ? cf[1] : 0.f) - (tv > p ? cf[2] : 0.f);
                    if (has_note) dv += cf[3] * 2.f * (p - tv);
                    dpre1 = dv * p * (1.f - p);
                    const float dpre1s = kSlope * dpre1;
                    const f32x2 dp2 = pk2(dpre1, dpre1);
#pragma unroll
                    for (int qp = 0; qp < 15; qp++) {
                        float al, ah;
                        upk2(av[qp], al, ah);
                        const f32x2 din = mul2(MST_CPSA2(kPsaWl + 50 + qp * 2), pk2(al > 0.f ? dpre1 : dpre1s, ah > 0.f ? dpre1 : dpre1s));
                        doct[qp] = add2(doct[qp], din);
                        gw1[qp] = fma2(av[qp], dp2, gw1[qp]);
                        *reinterpret_cast<f32x2 *>(s.sa + lane * 34 + qp * 2) = din;
                    }
                }
                {   // dps[note][1] += sum over the channel lanes (lane = c * 8 + o: xor 8, 16)
                    float v = dpre1;
                    v += __shfl_xor_sync(0xffffffffu, v, 8);
                    v += __shfl_xor_sync(0xffffffffu, v, 16);
                    if (c_l == 0) s.dps[note * 5 + 1] += v;
                }
                unsigned todo = __ballot_sync(0xffffffffu, has_note);
                __syncwarp();
                // cells with a target note: the four other outputs, cooperatively (lane = column q)
                while (todo) {
                    const int m = __ffs(todo) - 1;
                    todo &= todo - 1;
                    const int cm = m >> 3, om = m & 7, note_m = om * 7 + d;
                    const float am = lane < 30 ? leaky(s.od[cm * 450 + om * 30 + lane] + s.od[cm * 450 + 240 + d * 30 + lane]) : 0.f;
                    float y0 = wx0 * am, y2 = wx2 * am, y3 = wx3 * am, y4 = wx4 * am;
#pragma unroll
                    for (int sft = 16; sft; sft >>= 1) {
                        y0 += __shfl_xor_sync(0xffffffffu, y0, sft);
                        y2 += __shfl_xor_sync(0xffffffffu, y2, sft);
                        y3 += __shfl_xor_sync(0xffffffffu, y3, sft);
                        y4 += __shfl_xor_sync(0xffffffffu, y4, sft);
                    }
                    const float *sn = s.snote + note_m * 5;
                    const float *tm = a.target + ((((long long)n * C + c0 + cm) * per_n) + btf) * 280 + note_m * 5;
                    y0 = sigmoidf(y0 + sn[0]) * 6.f;
                    y2 = sigmoidf(y2 + sn[2]); y3 = sigmoidf(y3 + sn[3]); y4 = sigmoidf(y4 + sn[4]);
                    const float d0 = cf[4] * 2.f * (y0 - fminf(tm[0], 6.f)) / 36.f * y0 * (1.f - y0 / 6.f);
                    const float q2 = y2 * (1.f - y2), q3 = y3 * (1.f - y3), q4 = y4 * (1.f - y4);
                    const float d2 = cf[5] * (y2 - tm[2]) / fmaxf(q2, 1e-12f) * q2;
                    const float d3 = cf[5] * (y3 - tm[3]) / fmaxf(q3, 1e-12f) * q3;
                    const float d4 = cf[5] * (y4 - tm[4]) / fmaxf(q4, 1e-12f) * q4;
                    if (lane < 30) {
                        const float ex = (wx0 * d0 + wx2 * d2 + wx3 * d3 + wx4 * d4) * leaky_grad_from_out(am);
                        s.sa[m * 34 + lane] += ex;
                        s.exq[lane] = ex;
                        gwx0 = fmaf(am, d0, gwx0); gwx2 = fmaf(am, d2, gwx2); gwx3 = fmaf(am, d3, gwx3); gwx4 = fmaf(am, d4, gwx4);
                    }
                    if (lane == 0) {
                        float *dp = s.dps + note_m * 5;
                        dp[0] += d0; dp[2] += d2; dp[3] += d3; dp[4] += d4;
                    }
                    __syncwarp();
                    if (lane == m) {
#pragma unroll
                        for (int qp = 0; qp < 15; qp++) doct[qp] = add2(doct[qp], ld2(s.exq + qp * 2));
                    }
                    __syncwarp();
                }
                // degree-row gradient: sum over the 8 octave lanes of each channel (lane = column q), written over od in place
                if (lane < 30) {
                    for (int c = 0; c < cg; c++) {
                        const float *col = s.sa + c * 8 * 34 + lane;
                        const float sum = ((col[0] + col[34]) + (col[68] + col[102])) + ((col[136] + col[170]) + (col[204] + col[238]));
                        float *p = s.od + c * 450 + 240 + d * 30 + lane;
                        *p = sum * leaky_grad_from_out(*p);
                    }
                }
                __syncwarp();
            }
            if (active) {
#pragma unroll
                for (int qp = 0; qp < 15; qp++) {
                    float ol, oh;
                    upk2(oct[qp], ol, oh);
                    *reinterpret_cast<f32x2 *>(s.od + c_l * 450 + o * 30 + qp * 2) =
                        mul2(doct[qp], pk2(leaky_grad_from_out(ol), leaky_grad_from_out(oh)));
                }
            }
            __syncwarp();
            for (int jp = lane; jp < 225; jp += 32) {
                f32x2 sum = pk2(0.f, 0.f);
                for (int c = 0; c < cg; c++) {
                    const f32x2 v = ld2(s.od + c * 450 + jp * 2);
                    float vl, vh;
                    upk2(v, vl, vh);
                    sum = add2(sum, v);
                    atomicAdd(&dzi_s[(c0 + c) * 450 + jp * 2], vl);
                    atomicAdd(&dzi_s[(c0 + c) * 450 + jp * 2 + 1], vh);
                }
                f32x2 *dst = reinterpret_cast<f32x2 *>(dzr + jp * 2);
                *dst = c0 == 0 ? sum : add2(*dst, sum);
            }
        }
        __syncwarp();
        float *dps = a.dps + u * a.dps_stride;
        for (int i = lane; i < 280; i += 32) dps[i] = s.dps[i];
    }
    // ---- flush the per-lane weight-gradient accumulators
#pragma unroll
    for (int q = 0; q < 30; q++) {
        float lo, hi;
        upk2(gw1[q >> 1], lo, hi);
        float v = (q & 1) ? hi : lo;
        for (int sft = 16; sft; sft >>= 1) v += __shfl_xor_sync(0xffffffffu, v, sft);
        if (lane == 0 && v != 0.f) atomicAdd(a.gWl + 50 + q, v);
    }
    if (lane < 30) {
        if (gwx0 != 0.f) atomicAdd(a.gWl + 0 * 50 + lane, gwx0);
        if (gwx2 != 0.f) atomicAdd(a.gWl + 2 * 50 + lane, gwx2);
        if (gwx3 != 0.f) atomicAdd(a.gWl + 3 * 50 + lane, gwx3);
        if (gwx4 != 0.f) atomicAdd(a.gWl + 4 * 50 + lane, gwx4);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < C * 450; i += blockDim.x)
        if (dzi_s[i] != 0.f) atomicAdd(a.dZi + (long long)n * C * 450 + i, dzi_s[i]);
}

// melody branch of the PitchedStyleApplier backward: one warp per unit, operands staged for the tensor cores; the next
// unit's melody rows and dps rows are fetched with cp.async while the current unit is processed
constexpr int kPsaMelWarps = 16;
struct PsaMelWarpSmem {
    float mf[56 * 24];       // mf (col 20 = 1 for the bias row, 21..23 = 0), later the pre-activation gradient dmf
    float mel[2][56 * 8];    // double-buffered global rows, used in place as MMA operands
    float dps[2][56 * 5];
};
static_assert(sizeof(PsaMelWarpSmem) % 16 == 0, "per-warp shared block must stay 16-byte aligned");

struct PsaMelBwdArgs {
    const float *melody;              // (units, 56, 8)
    const float *dps;                 // per unit (56, 5) at stride dps_stride (may alias dmelody: read before written)
    long long dps_stride;             // multiple of 4 floats (16-byte copies)
    float *dmelody;                   // (units, 56, 8) written
    float *gWm, *gbm, *gWl, *gbl;     // melody_linear (20, 8), (20); final linear columns 30..49 of (5, 50), (5)
    long long units;
};

__global__ void __launch_bounds__(kPsaMelWarps * 32, 1)
k_psa_mel_bwd(PsaMelBwdArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    PsaMelWarpSmem &s = reinterpret_cast<PsaMelWarpSmem *>(smem_raw)[warp];
    for (int i = lane; i < 56 * 24; i += 32) s.mf[i] = 0.f;
    const int g = lane >> 2, t = lane & 3;
    float accM[2][4], accWm[2][4], gbm = 0.f;       // gbm: lane j < 20 owns melody_linear.bias[j]
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) { accM[i][j] = 0.f; accWm[i][j] = 0.f; }
    auto prefetch = [&](long long u, int buf) {
        const float *mg = a.melody + u * 448, *dg = a.dps + u * a.dps_stride;
        for (int i = lane; i < 112; i += 32) cp_async16(s.mel[buf] + i * 4, mg + i * 4);
        for (int i = lane; i < 70; i += 32) cp_async16(s.dps[buf] + i * 4, dg + i * 4);
        asm volatile("cp.async.commit_group;");
    };
    const long long u0 = (long long)blockIdx.x * kPsaMelWarps + warp, ustep = (long long)gridDim.x * kPsaMelWarps;
    if (u0 < a.units) prefetch(u0, 0);
    int buf = 0;
    for (long long u = u0; u < a.units; u += ustep, buf ^= 1) {
        __syncwarp();
        if (u + ustep < a.units) { prefetch(u + ustep, buf ^ 1); asm volatile("cp.async.wait_group 1;"); }
        else asm volatile("cp.async.wait_group 0;");
        __syncwarp();
        const float *mel = s.mel[buf], *dps = s.dps[buf];
#pragma unroll 1
        for (int note = lane; note < 56; note += 32) {
            const float4 m0 = *reinterpret_cast<const float4 *>(mel + note * 8), m1 = *reinterpret_cast<const float4 *>(mel + note * 8 + 4);
            const float mel8[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
            float mf[20], sn[5];
            psa_note(mel8, mf, sn);
#pragma unroll
            for (int j = 0; j < 20; j++) s.mf[note * 24 + j] = mf[j];
            s.mf[note * 24 + 20] = 1.f;
        }
        __syncwarp();
        // gWl[:, 30:50] (+ bias through the all-ones column 20) += mf^T . dps
#pragma unroll 1
        for (int ks = 0; ks < 7; ks++) {
            const float bf[2] = {g < 5 ? dps[(ks * 8 + t) * 5 + g] : 0.f, g < 5 ? dps[(ks * 8 + t + 4) * 5 + g] : 0.f};
            const float *r0 = s.mf + (ks * 8 + t) * 24 + g;
            const float a0[4] = {r0[0], r0[8], r0[4 * 24], r0[4 * 24 + 8]};
            const float a1[4] = {r0[16], 0.f, r0[4 * 24 + 16], 0.f};
            mma_tf32x3(accM[0], a0, bf);
            mma_tf32x3(accM[1], a1, bf);
        }
        __syncwarp();
#pragma unroll 1
        for (int note = lane; note < 56; note += 32) {
            float dmf[20], dp[5];
#pragma unroll
            for (int e = 0; e < 5; e++) dp[e] = dps[note * 5 + e];
#pragma unroll
            for (int j = 0; j < 20; j++) {
                float v = 0.f;
#pragma unroll
                for (int e = 0; e < 5; e++) v = fmaf(c_psa[kPsaWl + e * 50 + 30 + j], dp[e], v);
                v *= leaky_grad_from_out(s.mf[note * 24 + j]);
                dmf[j] = v;
                s.mf[note * 24 + j] = v;
            }
            s.mf[note * 24 + 20] = 0.f;
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
        if (lane < 20) {
            float v = 0.f;
#pragma unroll 8
            for (int note = 0; note < 56; note++) v += s.mf[note * 24 + lane];
            gbm += v;
        }
        // gWm += dmf^T . mel
#pragma unroll 1
        for (int ks = 0; ks < 7; ks++) {
            const float bf[2] = {mel[(ks * 8 + t) * 8 + g], mel[(ks * 8 + t + 4) * 8 + g]};
            const float *r0 = s.mf + (ks * 8 + t) * 24 + g;
            const float a0[4] = {r0[0], r0[8], r0[4 * 24], r0[4 * 24 + 8]};
            const float a1[4] = {r0[16], 0.f, r0[4 * 24 + 16], 0.f};
            mma_tf32x3(accWm[0], a0, bf);
            mma_tf32x3(accWm[1], a1, bf);
        }
    }
    // ---- flush (fragment layout: rows g / g + 8, columns 2t / 2t + 1)
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int row = mt * 16 + g + (i >> 1) * 8, col = 2 * t + (i & 1);
            if (col < 5) {
                if (row < 20 && accM[mt][i] != 0.f) atomicAdd(a.gWl + col * 50 + 30 + row, accM[mt][i]);
                if (row == 20 && accM[mt][i] != 0.f) atomicAdd(a.gbl + col, accM[mt][i]);
            }
            if (row < 20 && accWm[mt][i] != 0.f) atomicAdd(a.gWm + row * 8 + col, accWm[mt][i]);
        }
    if (lane < 20 && gbm != 0.f) atomicAdd(a.gbm + lane, gbm);
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

__global__ void __launch_bounds__(kUsaWarps * 32)
k_usa_bwd(UsaBwdArgs a) {
    __shared__ __align__(16) UsaWarpSmem sm[kUsaWarps];
    __shared__ float g_s[18];
    const UsaArgs &f = a.f;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    UsaWarpSmem &s = sm[warp];
    if (threadIdx.x < 18) g_s[threadIdx.x] = 0.f;
    __syncthreads();
    const long long per_n = (long long)f.B * f.T * 10;
    const long long units = (long long)f.N * per_n;
    float wl[16];
#pragma unroll
    for (int i = 0; i < 16; i++) wl[i] = f.Wl[i];
    const float bl0 = f.bl[0], bl1 = f.bl[1];
    float gw[18];
#pragma unroll
    for (int i = 0; i < 18; i++) gw[i] = 0.f;
    auto prefetch = [&](long long uu, int buf) {
        const float *hg = f.H + uu * 376;
        for (int i = lane; i < 94; i += 32) cp_async16(s.row[buf] + i * 4, hg + i * 4);
        asm volatile("cp.async.commit_group;");
    };
    const long long u0 = (long long)blockIdx.x * kUsaWarps + warp, ustep = (long long)gridDim.x * kUsaWarps;
    if (u0 < units) prefetch(u0, 0);
    int buf = 0;
    for (long long u = u0; u < units; u += ustep, buf ^= 1) {
        const int n = (int)(u / per_n);
        __syncwarp();
        if (u + ustep < units) { prefetch(u + ustep, buf ^ 1); asm volatile("cp.async.wait_group 1;"); }
        else asm volatile("cp.async.wait_group 0;");
        __syncwarp();
        const float *cf = a.coef + n * kLossCoefs;
        for (int note = lane; note < 47; note += 32) {
            const float4 h0 = *reinterpret_cast<const float4 *>(s.row[buf] + note * 8), h1 = *reinterpret_cast<const float4 *>(s.row[buf] + note * 8 + 4);
            const float hv[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
            float y[2], t2[2], dpre[2];
            float v0 = bl0, v1 = bl1;
#pragma unroll
            for (int q = 0; q < 8; q++) { v0 = fmaf(wl[q], hv[q], v0); v1 = fmaf(wl[8 + q], hv[q], v1); }
            y[0] = sigmoidf(v0) * 6.f; y[1] = sigmoidf(v1);
            const float2 tt = *reinterpret_cast<const float2 *>(a.target + (u * 47 + note) * 2);
            t2[0] = tt.x; t2[1] = tt.y;
            loss_cell_grad<2>(y, t2, cf, dpre);
            float dh[8];
#pragma unroll
            for (int q = 0; q < 8; q++) {
                dh[q] = (wl[q] * dpre[0] + wl[8 + q] * dpre[1]) * leaky_grad_from_out(hv[q]);
                gw[q] = fmaf(dpre[0], hv[q], gw[q]);
                gw[8 + q] = fmaf(dpre[1], hv[q], gw[8 + q]);
            }
            gw[16] += dpre[0]; gw[17] += dpre[1];
            float4 *dst = reinterpret_cast<float4 *>(a.dHpre + u * 376 + note * 8);
            dst[0] = make_float4(dh[0], dh[1], dh[2], dh[3]);
            dst[1] = make_float4(dh[4], dh[5], dh[6], dh[7]);
        }
    }
#pragma unroll
    for (int i = 0; i < 18; i++) {
        float v = gw[i];
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) atomicAdd(&g_s[i], v);
    }
    __syncthreads();
    if (threadIdx.x < 16) atomicAdd(a.gWl + threadIdx.x, g_s[threadIdx.x]);
    else if (threadIdx.x < 18) atomicAdd(a.gbl + (threadIdx.x - 16), g_s[threadIdx.x]);
}

}  // namespace mstm
