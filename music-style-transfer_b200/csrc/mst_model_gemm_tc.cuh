// mst_model_gemm_tc.cuh — tensor-core GEMM for the row-space linears (model.py: every nn.Linear / Distributed(Linear)
// over (n, c, b, t, f) rows, their input gradients dY.W and weight gradients dY^T.X).
//
// Same contract as k_gemm (GemmArgs: explicit strides, bias / leaky epilogue, store / += / atomic split-K), but the
// products run on the tensor cores: mma.sync.m16n8k8 TF32 with the 3-term split  a.b ~ a_hi.b_hi + a_hi.b_lo + a_lo.b_hi
// (a_hi = tf32(a), a_lo = tf32(a - a_hi)), accumulated in fp32.  The dropped a_lo.b_lo term is 2^-22 relative, so the
// result is fp32-accurate: the 1e-5 loss / 1e-4 gradient parity bars hold unchanged.
//
// These GEMMs are tall and skinny (M = 10^5..10^6 rows, N or K <= 64 on one side): the streamed operand is read once
// from HBM and the bound is HBM, not the tensor pipe.  The kernel is therefore built around the stream:
//   * 128 x BN x 16 tiles (BN = 64 or 32), 8 warps as 4 (M) x 2 (N), a 3-stage cp.async ring (two K-slabs in flight
//     while the third is multiplied; zero fill at the matrix / split-K edges), ONE __syncthreads per slab;
//   * each operand is staged in the shared layout that matches its contiguous direction in global memory
//     (template flags AK / BK: k-contiguous -> [row][16 + 4], otherwise [k][rows + 8]); both layouts give
//     conflict-free fragment loads (row stride = 20 or 8 mod 32 words) and coalesced global loads.
#pragma once
#include "mst_model_kernels.cuh"

namespace mstm {

constexpr int kTcBM = 128, kTcBK = 16, kTcThreads = 256, kTcStages = 3;

__device__ __forceinline__ void tc_split(float x, uint32_t &hi, uint32_t &lo) {
    const float h = tf32_rna(x);
    hi = __float_as_uint(h);
    lo = __float_as_uint(tf32_rna(x - h));
}

__device__ __forceinline__ void tc_mma(float *d, const uint32_t *a, const uint32_t *b) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// shared-memory footprint of one operand tile of R rows: k-contiguous [R][20] or row-contiguous [16][R + 8]
template <int R, bool KC> struct TcTile { static constexpr int kFloats = KC ? R * (kTcBK + 4) : kTcBK * (R + 8); };

template <int BN, bool AK, bool BK>
__global__ void __launch_bounds__(kTcThreads)
k_gemm_tc(GemmArgs g) {
    constexpr int BM = kTcBM, KT = kTcBK;
    constexpr int A_FL = TcTile<BM, AK>::kFloats, B_FL = TcTile<BN, BK>::kFloats;
    constexpr int A_PER = BM * KT / kTcThreads;          // 8 elements per thread
    constexpr int B_PER = BN * KT / kTcThreads;          // 4 or 2
    constexpr int WN = BN / 2, NT = WN / 8;              // warp tile 32 x WN: 2 m-tiles x NT n-tiles
    constexpr int STAGES = kTcStages, STAGE_FL = A_FL + B_FL;
    __shared__ __align__(16) float sm[STAGES * STAGE_FL];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int gq = lane >> 2, tq = lane & 3;
    const int wm = (warp & 3) * 32, wn = (warp >> 2) * WN;
    const int i0 = blockIdx.y * BM, j0 = blockIdx.x * BN;
    const int kz = gridDim.z;
    const int k_per = ((g.K + kz - 1) / kz + KT - 1) / KT * KT;
    const int k_lo = blockIdx.z * k_per, k_hi = min(g.K, k_lo + k_per);

    float acc[2][NT][4];
#pragma unroll
    for (int a = 0; a < 2; a++)
#pragma unroll
        for (int b = 0; b < NT; b++)
#pragma unroll
            for (int c = 0; c < 4; c++) acc[a][b][c] = 0.f;
    if (k_lo >= k_hi) return;

    // per-thread copy slots: global element pointer (advanced by one K-slab per issue), shared offset, k index, row validity
    const float *pa[A_PER], *pb[B_PER];
    int ka[A_PER], kb[B_PER];
    uint32_t sa[A_PER], sb[B_PER];
    const uint32_t sm_base = (uint32_t)__cvta_generic_to_shared(sm);
#pragma unroll
    for (int e = 0; e < A_PER; e++) {
        const int idx = tid + e * kTcThreads;
        int ii, kk;
        if (AK) { kk = idx & (KT - 1); ii = idx / KT; } else { ii = idx & (BM - 1); kk = idx / BM; }
        const int gi = i0 + ii;
        ka[e] = gi < g.M ? k_lo + kk : (1 << 30);                  // invalid rows never pass the k < k_hi test
        pa[e] = g.A + (long long)min(gi, g.M - 1) * g.a_si + (long long)(k_lo + kk) * g.a_sk;
        sa[e] = sm_base + 4u * (uint32_t)(AK ? ii * (KT + 4) + kk : kk * (BM + 8) + ii);
    }
#pragma unroll
    for (int e = 0; e < B_PER; e++) {
        const int idx = tid + e * kTcThreads;
        int jj, kk;
        if (BK) { kk = idx & (KT - 1); jj = idx / KT; } else { jj = idx & (BN - 1); kk = idx / BN; }
        const int gj = j0 + jj;
        kb[e] = gj < g.N ? k_lo + kk : (1 << 30);
        pb[e] = g.B + (long long)(k_lo + kk) * g.b_sk + (long long)min(gj, g.N - 1) * g.b_sj;
        sb[e] = sm_base + 4u * (uint32_t)(A_FL + (BK ? jj * (KT + 4) + kk : kk * (BN + 8) + jj));
    }
    const long long a_step = (long long)KT * g.a_sk, b_step = (long long)KT * g.b_sk;
    // one K-slab global -> shared with 4-byte cp.async (zero fill outside the matrix / the split-K range)
    auto issue = [&](int stage) {
        const uint32_t so = 4u * (uint32_t)(stage * STAGE_FL);
#pragma unroll
        for (int e = 0; e < A_PER; e++) {
            const bool ok = ka[e] < k_hi;
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(sa[e] + so), "l"(ok ? pa[e] : g.A), "r"(ok ? 4 : 0));
            pa[e] += a_step; ka[e] += KT;
        }
#pragma unroll
        for (int e = 0; e < B_PER; e++) {
            const bool ok = kb[e] < k_hi;
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(sb[e] + so), "l"(ok ? pb[e] : g.B), "r"(ok ? 4 : 0));
            pb[e] += b_step; kb[e] += KT;
        }
    };
    const int n_slabs = (k_hi - k_lo + KT - 1) / KT;
#pragma unroll
    for (int s = 0; s < STAGES - 1; s++) {
        if (s < n_slabs) issue(s);
        asm volatile("cp.async.commit_group;");
    }
    for (int it = 0; it < n_slabs; it++) {
        asm volatile("cp.async.wait_group %0;" ::"n"(STAGES - 2));
        __syncthreads();                         // slab `it` has landed for every thread; slab it - 1 is no longer read
        if (it + STAGES - 1 < n_slabs) issue((it + STAGES - 1) % STAGES);
        asm volatile("cp.async.commit_group;");
        const float *As = sm + (it % STAGES) * STAGE_FL, *Bs = As + A_FL;
        auto a_at = [&](int m, int k) -> float { return AK ? As[m * (KT + 4) + k] : As[k * (BM + 8) + m]; };
        auto b_at = [&](int n, int k) -> float { return BK ? Bs[n * (KT + 4) + k] : Bs[k * (BN + 8) + n]; };
#pragma unroll
        for (int ks = 0; ks < KT; ks += 8) {
            uint32_t bh[NT][2], bl[NT][2];
#pragma unroll
            for (int nt = 0; nt < NT; nt++) {
                tc_split(b_at(wn + nt * 8 + gq, ks + tq), bh[nt][0], bl[nt][0]);
                tc_split(b_at(wn + nt * 8 + gq, ks + tq + 4), bh[nt][1], bl[nt][1]);
            }
#pragma unroll
            for (int mt = 0; mt < 2; mt++) {
                uint32_t ah[4], al[4];
                const int m = wm + mt * 16 + gq;
                tc_split(a_at(m, ks + tq), ah[0], al[0]);
                tc_split(a_at(m + 8, ks + tq), ah[1], al[1]);
                tc_split(a_at(m, ks + tq + 4), ah[2], al[2]);
                tc_split(a_at(m + 8, ks + tq + 4), ah[3], al[3]);
#pragma unroll
                for (int nt = 0; nt < NT; nt++) {
                    tc_mma(acc[mt][nt], al, bh[nt]);
                    tc_mma(acc[mt][nt], ah, bl[nt]);
                    tc_mma(acc[mt][nt], ah, bh[nt]);
                }
            }
        }
    }
    asm volatile("cp.async.wait_group 0;");
    // epilogue: fragment (row gq / gq + 8, columns 2 tq, 2 tq + 1); 8-byte stores when the destination allows it
    const bool pair_ok = g.accumulate != 2 && (g.c_si & 1) == 0 && ((reinterpret_cast<uintptr_t>(g.C) & 7) == 0);
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
        for (int nt = 0; nt < NT; nt++)
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int gi = i0 + wm + mt * 16 + gq + h * 8;
                const int gj = j0 + wn + nt * 8 + 2 * tq;
                if (gi >= g.M || gj >= g.N) continue;
                float v0 = acc[mt][nt][h * 2], v1 = acc[mt][nt][h * 2 + 1];
                const bool has1 = gj + 1 < g.N;
                if (g.bias && blockIdx.z == 0) { v0 += g.bias[gj]; if (has1) v1 += g.bias[gj + 1]; }
                if (g.act == 1) { v0 = leaky(v0); v1 = leaky(v1); }
                float *dst = g.C + (long long)gi * g.c_si + gj;
                if (pair_ok && has1) {
                    float2 *d2 = reinterpret_cast<float2 *>(dst);
                    if (g.accumulate == 1) { const float2 o = *d2; v0 += o.x; v1 += o.y; }
                    *d2 = make_float2(v0, v1);
                } else if (g.accumulate == 0) { dst[0] = v0; if (has1) dst[1] = v1; }
                else if (g.accumulate == 1) { dst[0] += v0; if (has1) dst[1] += v1; }
                else { atomicAdd(dst, v0); if (has1) atomicAdd(dst + 1, v1); }
            }
}

// ------------------------------------------------------------------------------------------------
// k_gemm_smallk — C[M, N] (op)= act(A[M, K] . B + bias) for a SHORT contraction (K <= kSkMaxK) over a huge row count:
// the Zr projections of the style appliers (rhythm features, K = 10 / 17, to 450 / 376 columns).  These are pure
// output streams (0.9 GB written, ~20 MB read): the tensor-core kernels pay their staging and epilogue for nothing.
// Thread j owns output column j: its K weights sit in registers as (w, w) pairs for the whole (persistent) CTA; rows
// are taken four at a time from a transposed shared tile ([k][row], one broadcast 16-byte load per k) and multiplied as
// two packed FFMA2; a warp stores 128 contiguous bytes per row.  A rows must be k-contiguous (a_sk == 1).
// ------------------------------------------------------------------------------------------------
constexpr int kSkMaxK = 24, kSkRows = 64;

// KP = K rounded up to a multiple of 4 (compile-time trip count; the padding weights and tile rows are zero)
template <int KP>
__global__ void __launch_bounds__(512, 2)
k_gemm_smallk(GemmArgs g) {
    __shared__ __align__(16) float xs[2][KP][kSkRows];
    const int j = threadIdx.x, K = g.K;
    const bool col_ok = j < g.N;
    for (int i = threadIdx.x; i < 2 * (KP - K) * kSkRows; i += blockDim.x)
        xs[i / ((KP - K) * kSkRows)][K + (i / kSkRows) % (KP - K)][i % kSkRows] = 0.f;
    f32x2 w[KP];
#pragma unroll
    for (int k = 0; k < KP; k++) {
        const float v = (col_ok && k < K) ? g.B[(long long)k * g.b_sk + (long long)j * g.b_sj] : 0.f;
        w[k] = pk2(v, v);
    }
    const float bias = (col_ok && g.bias) ? g.bias[j] : 0.f;
    const f32x2 bias2 = pk2(bias, bias);
    const int n_blocks = (g.M + kSkRows - 1) / kSkRows;
    // x tile of block b -> xs[buf] transposed; rows past M are clamped (their outputs are never stored)
    auto stage = [&](int b, int buf) {
        const int r0 = b * kSkRows;
        for (int i = threadIdx.x; i < kSkRows * K; i += blockDim.x) {
            const int r = i / K, k = i - r * K;
            xs[buf][k][r] = g.A[(long long)min(r0 + r, g.M - 1) * g.a_si + k];
        }
    };
    int buf = 0;
    if ((int)blockIdx.x < n_blocks) stage(blockIdx.x, 0);
    __syncthreads();
    for (int b = blockIdx.x; b < n_blocks; b += gridDim.x, buf ^= 1) {
        if (b + (int)gridDim.x < n_blocks) stage(b + gridDim.x, buf ^ 1);
        if (col_ok) {
            const int r0 = b * kSkRows;
#pragma unroll 2
            for (int r = 0; r < kSkRows; r += 4) {
                f32x2 a0 = bias2, a1 = bias2;
#pragma unroll
                for (int k = 0; k < KP; k++) {
                    const ulonglong2 x = *reinterpret_cast<const ulonglong2 *>(&xs[buf][k][r]);
                    a0 = fma2(x.x, w[k], a0);
                    a1 = fma2(x.y, w[k], a1);
                }
                float v[4];
                upk2(a0, v[0], v[1]);
                upk2(a1, v[2], v[3]);
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    if (r0 + r + e < g.M) {
                        float o = g.act == 1 ? leaky(v[e]) : v[e];
                        float *dst = g.C + (long long)(r0 + r + e) * g.c_si + j;
                        *dst = g.accumulate == 1 ? *dst + o : o;
                    }
                }
            }
        }
        __syncthreads();
    }
}

}  // namespace mstm
