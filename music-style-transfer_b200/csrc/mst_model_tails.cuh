// mst_model_tails.cuh — fused per-note kernels (forward): the (rows x 56 notes) tensors of the reference's
// PitchedStyleApplier (:658-675) and UnpitchedStyleApplier (:713-724) (the MelodyEncoder tail lives in mst_model_me.cuh)
// are produced and consumed inside one kernel each; nothing of size note-rows x {15, 30, 50} reaches HBM.
#pragma once
#include "mst_model_kernels2.cuh"

namespace mstm {

// ------------------------------------------------------------------------------------------------
// PitchedStyleApplier tail (model.py:658-675).  Row (n, c, b, t, f): od[450] = leaky(bias + Zr[n,b,t,f] + Zi[n,c] + Zs[n])
// (240 octave values = 8 x 30, then 210 degree values = 7 x 30), then per note:
//   a[q] = leaky(oct[o][q] + deg[d][q]) (30), mf = leaky(bm + Wm.melody[n,b,t,f,note]) (20),
//   y = bl + Wl.[a; mf] (5) -> duration 6*sigmoid, the rest sigmoid (model.py:565-579, 668-673).
// ONE WARP per (n, b, t, f) unit, no block-level synchronisation.
//   * the small weights (melody_linear 20x8, linear 5x50 and their biases) sit in the constant bank, so every FFMA
//     takes its weight as an immediate c[bank][offset] operand: no load instruction per multiply;
//   * the melody branch (mf = leaky(bm + Wm.mel), its share of the final linear) is computed once per (unit, note)
//     and shared by all channels: y = snote + Wl[:, :30].a;
//   * lane = (channel-in-group c, octave o): the octave row of the outer add lives in 30 registers, the degree row
//     is a broadcast shared-memory read, the loop runs over the 7 degrees; 4 channels x 8 octaves fill the warp;
//   * the (C, 56, 5) output tile is staged in shared memory and written with 16-byte coalesced stores.
// Zci[n,c][450] = od_bias + Zi[n,c] + Zs[n] is pre-summed by k_psa_zci (tiny).
// ------------------------------------------------------------------------------------------------
constexpr int kPsaWm = 0, kPsaBm = 160, kPsaWl = 180, kPsaBl = 430, kPsaConst = 440;
__constant__ __align__(16) float c_psa[kPsaConst];

__global__ void __launch_bounds__(256)
k_psa_zci(const float *__restrict__ bias, const float *__restrict__ Zi, const float *__restrict__ Zs, float *__restrict__ Zci, int N, int C) {
    const int total = N * C * 450;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int j = i % 450, n = i / (450 * C);
        Zci[i] = bias[j] + Zi[i] + Zs[n * 450 + j];
    }
}

constexpr int kPsaWarps = 12;
struct PsaWarpSmem {
    float od[4 * 450];       // post-activation octave/degree rows of the current channel group; then the output staging
    float snote[56 * 5];     // bl + Wl[:, 30:].mf per note
    float mel[56 * 8];       // this unit's melody rows / Zr row, fetched with cp.async during the previous unit
    float zr[452];
};
static_assert(sizeof(PsaWarpSmem) % 16 == 0, "per-warp shared block must stay 16-byte aligned");

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc));
}
__device__ __forceinline__ void cp_async8(void *smem_dst, const void *gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc));
}

#define MST_CPSA2(off) (*reinterpret_cast<const f32x2 *>(&c_psa[off]))                                    // even offsets
// phase 1 for one note: mf[20] (as 10 pairs) and snote[5] from the melody vector (4 pairs)
__device__ __forceinline__ void psa_note2(const f32x2 *mel, f32x2 *mf2, float *sn) {
#pragma unroll
    for (int jp = 0; jp < 10; jp++) {
        float m[2];
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int j = jp * 2 + h;
            f32x2 v = pk2(c_psa[kPsaBm + j], 0.f);
#pragma unroll
            for (int pp = 0; pp < 4; pp++) v = fma2(MST_CPSA2(kPsaWm + j * 8 + pp * 2), mel[pp], v);
            m[h] = sum2(v);
        }
        mf2[jp] = leaky2(pk2(m[0], m[1]));
    }
#pragma unroll
    for (int e = 0; e < 5; e++) {
        f32x2 v = pk2(c_psa[kPsaBl + e], 0.f);
#pragma unroll
        for (int jp = 0; jp < 10; jp++) v = fma2(MST_CPSA2(kPsaWl + e * 50 + 30 + jp * 2), mf2[jp], v);
        sn[e] = sum2(v);
    }
}
// scalar form (the backward kernels use it)
__device__ __forceinline__ void psa_note(const float *mel8, float *mf, float *sn) {
    const f32x2 mel[4] = {pk2(mel8[0], mel8[1]), pk2(mel8[2], mel8[3]), pk2(mel8[4], mel8[5]), pk2(mel8[6], mel8[7])};
    f32x2 mf2[10];
    psa_note2(mel, mf2, sn);
#pragma unroll
    for (int jp = 0; jp < 10; jp++) upk2(mf2[jp], mf[jp * 2], mf[jp * 2 + 1]);
}

struct PsaArgs2 {
    const float *Zr, *Zci;       // (N,B,T,10,450), (N,C,450)
    const float *melody;         // (N,B,T,10,56,8)
    float *out;                  // (N,C,B,T,10,56,5)
    int N, C, B, T;
};

// od rows of one channel group: od[c][j] = leaky(zr[j] + Zci[n, c0 + c][j]), two columns per lane and step
__device__ __forceinline__ void psa_build_od(float *od, const float *zr, const float *zci, int cg, int lane) {
    for (int jp = lane; jp < 225; jp += 32) {
        const f32x2 z = ld2(zr + jp * 2);
        for (int c = 0; c < cg; c++)
            *reinterpret_cast<f32x2 *>(od + c * 450 + jp * 2) = leaky2(add2(z, ld2(zci + c * 450 + jp * 2)));
    }
}

__global__ void __launch_bounds__(kPsaWarps * 32, 1)
k_psa_fwd3(PsaArgs2 a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    PsaWarpSmem &s = reinterpret_cast<PsaWarpSmem *>(smem_raw)[warp];
    const long long per_n = (long long)a.B * a.T * 10;
    const long long units = (long long)a.N * per_n;
    const int c_l = lane >> 3, o = lane & 7;
    auto prefetch = [&](long long uu) {           // Zr rows are 1800 bytes: 8-byte aligned
        const float *mg = a.melody + uu * 448, *zg = a.Zr + uu * 450;
        for (int i = lane; i < 112; i += 32) cp_async16(s.mel + i * 4, mg + i * 4);
        for (int i = lane; i < 225; i += 32) cp_async8(s.zr + i * 2, zg + i * 2);
        asm volatile("cp.async.commit_group;");
    };
    const long long u0 = (long long)blockIdx.x * kPsaWarps + warp, ustep = (long long)gridDim.x * kPsaWarps;
    if (u0 < units) prefetch(u0);
    for (long long u = u0; u < units; u += ustep) {
        const int n = (int)(u / per_n);
        const long long btf = u - (long long)n * per_n;
        asm volatile("cp.async.wait_group 0;");
        __syncwarp();
        // phase 1: per-note melody branch
        for (int note = lane; note < 56; note += 32) {
            const float4 *mp = reinterpret_cast<const float4 *>(s.mel + note * 8);
            const float4 m0 = mp[0], m1 = mp[1];
            const f32x2 mel[4] = {pk2(m0.x, m0.y), pk2(m0.z, m0.w), pk2(m1.x, m1.y), pk2(m1.z, m1.w)};
            f32x2 mf2[10];
            float sn[5];
            psa_note2(mel, mf2, sn);
#pragma unroll
            for (int e = 0; e < 5; e++) s.snote[note * 5 + e] = sn[e];
        }
        for (int c0 = 0; c0 < a.C; c0 += 4) {
            const int cg = min(4, a.C - c0);
            __syncwarp();
            psa_build_od(s.od, s.zr, a.Zci + ((long long)n * a.C + c0) * 450, cg, lane);
            if (c0 + 4 >= a.C && u + ustep < units) {       // both staging rows are consumed: fetch the next unit's
                __syncwarp();
                prefetch(u + ustep);
            }
            __syncwarp();
            float y[7][5];
            if (c_l < cg) {
                // acc[d][e] = (sum over even q, sum over odd q) of Wl[e][q] . leaky(oct[q] + deg[d][q]); q pairs outermost
                // so that the five weight pairs of a step are fetched once for the seven degrees
                f32x2 acc[7][5];
#pragma unroll
                for (int d = 0; d < 7; d++)
#pragma unroll
                    for (int e = 0; e < 5; e++) acc[d][e] = pk2(s.snote[(o * 7 + d) * 5 + e], 0.f);
                const float *oct = s.od + c_l * 450 + o * 30, *deg = s.od + c_l * 450 + 240;
#pragma unroll
                for (int qp = 0; qp < 15; qp++) {
                    const f32x2 oc = ld2(oct + qp * 2);
                    f32x2 w[5];
#pragma unroll
                    for (int e = 0; e < 5; e++) w[e] = MST_CPSA2(kPsaWl + e * 50 + qp * 2);
#pragma unroll
                    for (int d = 0; d < 7; d++) {
                        const f32x2 v = leaky2(add2(oc, ld2(deg + d * 30 + qp * 2)));
#pragma unroll
                        for (int e = 0; e < 5; e++) acc[d][e] = fma2(w[e], v, acc[d][e]);
                    }
                }
#pragma unroll
                for (int d = 0; d < 7; d++)
#pragma unroll
                    for (int e = 0; e < 5; e++) y[d][e] = sigmoid_fast(sum2(acc[d][e])) * (e == 0 ? 6.f : 1.f);
            }
            __syncwarp();                                   // every lane is done with od: it becomes the output staging
            if (c_l < cg) {
#pragma unroll
                for (int d = 0; d < 7; d++)
#pragma unroll
                    for (int e = 0; e < 5; e++) s.od[c_l * 280 + (o * 7 + d) * 5 + e] = y[d][e];
            }
            __syncwarp();
            for (int c = 0; c < cg; c++) {
                float4 *dst = reinterpret_cast<float4 *>(a.out + ((((long long)n * a.C + c0 + c) * per_n) + btf) * 280);
                const float4 *src = reinterpret_cast<const float4 *>(s.od + c * 280);
                for (int i = lane; i < 70; i += 32) dst[i] = src[i];
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// UnpitchedStyleApplier tail.  Row (n, b, t, f): h[376] = leaky(bias + Zr[n,b,t,f] + Zs[n,f]) viewed (47, 8);
// y[note] = bl + Wl·h[note] (2) -> duration 6*sigmoid, velocity sigmoid.  H is kept for the backward pass.
// ------------------------------------------------------------------------------------------------
struct UsaArgs {
    const float *bias, *Wl, *bl;   // (376), (2,8), (2)
    const float *Zr, *Zs;          // (N,B,T,10,376), (N,10,376)
    float *H;                      // (N,B,T,10,376) post-leaky
    float *out;                    // (N,B,T,10,47,2)
    int N, B, T;
};

// warp per unit; the next unit's Zr row arrives by cp.async while the current one is processed
constexpr int kUsaWarps = 8;
struct UsaWarpSmem { float row[2][376]; float h[376]; };
static_assert(sizeof(UsaWarpSmem) % 16 == 0, "per-warp shared block must stay 16-byte aligned");

__global__ void __launch_bounds__(kUsaWarps * 32)
k_usa_fwd(UsaArgs a) {
    __shared__ __align__(16) UsaWarpSmem sm[kUsaWarps];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    UsaWarpSmem &s = sm[warp];
    const long long per_n = (long long)a.B * a.T * 10;
    const long long units = (long long)a.N * per_n;
    auto prefetch = [&](long long uu, int buf) {
        const float *zg = a.Zr + uu * 376;
        for (int i = lane; i < 94; i += 32) cp_async16(s.row[buf] + i * 4, zg + i * 4);
        asm volatile("cp.async.commit_group;");
    };
    const long long u0 = (long long)blockIdx.x * kUsaWarps + warp, ustep = (long long)gridDim.x * kUsaWarps;
    if (u0 < units) prefetch(u0, 0);
    int buf = 0;
    for (long long u = u0; u < units; u += ustep, buf ^= 1) {
        const int n = (int)(u / per_n);
        const int f = (int)((u - (long long)n * per_n) % 10);
        __syncwarp();
        if (u + ustep < units) { prefetch(u + ustep, buf ^ 1); asm volatile("cp.async.wait_group 1;"); }
        else asm volatile("cp.async.wait_group 0;");
        __syncwarp();
        const float *zs = a.Zs + ((long long)n * 10 + f) * 376;
        for (int j = lane; j < 376; j += 32) {
            const float v = leaky(a.bias[j] + s.row[buf][j] + zs[j]);
            s.h[j] = v;
            a.H[u * 376 + j] = v;
        }
        __syncwarp();
        for (int note = lane; note < 47; note += 32) {
            const float4 h0 = *reinterpret_cast<const float4 *>(s.h + note * 8), h1 = *reinterpret_cast<const float4 *>(s.h + note * 8 + 4);
            const float h[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
            float y[2];
#pragma unroll
            for (int e = 0; e < 2; e++) {
                float v = a.bl[e];
#pragma unroll
                for (int q = 0; q < 8; q++) v = fmaf(a.Wl[e * 8 + q], h[q], v);
                y[e] = sigmoidf(v) * (e == 0 ? 6.f : 1.f);
            }
            *reinterpret_cast<float2 *>(a.out + (u * 47 + note) * 2) = make_float2(y[0], y[1]);
        }
    }
}

}  // namespace mstm
