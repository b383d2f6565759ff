// mst_model_psa_tc.cuh — PitchedStyleApplier tail (model.py:658-675) on the 5th-generation tensor cores, bf16 dense mode.
//
// The tail is, per note row (n, c, b, t, f, note):  y[5] = bl + Wl[5 x 50] . [ leaky(oct[o] + deg[d]) (30) ; mf (20) ].
// The fp32 kernels (mst_model_tails.cuh) evaluate the 5 x 50 contraction as 250 FP32 FMAs per row and are issue-bound at
// half the FMA pipe.  Here the contraction is a tcgen05.mma tile:
//   M = 128 note rows (two channels of one (n, b, t, f) unit: 2 x 56 rows + 16 idle), N = 16 (5 outputs, zero padded),
//   K = 64 (30 outer-add columns | 2 zero | 20 melody columns | 12 zero), operands bf16, accumulator fp32 in TMEM.
//   * A is BUILT in shared memory by the thread that owns the row: packed bf16x2 add / leaky of the octave and degree
//     rows (8 + 8 sixteen-byte shared loads, 12 packed ALU operations, 4 sixteen-byte stores per row) and a copy of the
//     note's melody columns (computed once per unit, shared by all channels), in the canonical K-major no-swizzle UMMA
//     layout (the one k_gemm_b16 uses: 8-row x 16-byte core matrices, LBO 144 B, SBO 1152 B);
//   * one elected thread issues the four K = 16 MMAs of a tile and commits them to an mbarrier; two A tiles and two
//     accumulators alternate, so the MMAs of tile i overlap the epilogue of tile i - 1 and the build of tile i + 1;
//   * epilogue: tcgen05.ld (32 lanes x 32 bit x 8 columns: thread = row) -> bias, sigmoid as 0.5 tanh.approx(x / 2) + 0.5
//     (one SFU operation; duration x 6) -> per-warp
//     staging -> 16-byte coalesced stores of the (56, 5) output slabs.
// The next unit's Zr row and melody rows arrive by cp.async while the current unit is processed.
#pragma once
#include <cuda_bf16.h>

#include "mst_model_gemm_b16.cuh"
#include "mst_model_tails.cuh"

namespace mstm {

constexpr int kPtcThreads = 128;
constexpr uint32_t kPtcATile = 16 * kT5SBO;                 // 128 rows: 18432 B
constexpr uint32_t kPtcOffA = 0;                            // two A tiles
constexpr uint32_t kPtcOffW = 2 * kPtcATile;                // Wl tile: 16 rows (2 row groups)
constexpr uint32_t kPtcOffOd = kPtcOffW + 2 * kT5SBO;       // od[2 channels][15 rows][40 bf16] (row stride 80 B: conflict-free)
constexpr uint32_t kPtcOdCh = 15 * 80;
constexpr uint32_t kPtcOffM = kPtcOffOd + 2 * kPtcOdCh;     // mf[56 notes][24 bf16]
constexpr uint32_t kPtcOffZr = kPtcOffM + 56 * 48;          // 2 x zr[452 floats]
constexpr uint32_t kPtcOffMel = kPtcOffZr + 2 * 1808;       // 2 x melody[56][8 floats]
constexpr uint32_t kPtcOffOut = kPtcOffMel + 2 * 1792;      // 4 warps x 160 floats
constexpr uint32_t kPtcSmem = kPtcOffOut + 4 * 640;
static_assert(kPtcOffOd % 16 == 0 && kPtcOffM % 16 == 0 && kPtcOffZr % 16 == 0 && kPtcOffMel % 16 == 0 && kPtcOffOut % 16 == 0, "16-byte aligned regions");

// column of the K axis that holds melody column j (the outer-add columns are 0..29)
constexpr int kPtcMelK = 32;

__device__ __forceinline__ uint32_t bf2_as_u32(__nv_bfloat162 v) { return *reinterpret_cast<uint32_t *>(&v); }
__device__ __forceinline__ __nv_bfloat162 u32_as_bf2(uint32_t v) { return *reinterpret_cast<__nv_bfloat162 *>(&v); }
// leaky_relu(a + b) on two packed bf16 values
__device__ __forceinline__ uint32_t bf2_add_leaky(uint32_t a, uint32_t b) {
    const __nv_bfloat162 s = __hadd2(u32_as_bf2(a), u32_as_bf2(b));
    return bf2_as_u32(__hmax2(s, __hmul2(s, __float2bfloat162_rn(kSlope))));
}

// Wl (5 x 50, constant bank) as the B operand: N = 16 rows x K = 64, K-major canonical layout
__device__ __forceinline__ void ptc_build_w(unsigned char *w, int tid) {
    for (int i = tid; i < 16 * 64; i += kPtcThreads) {
        const int e = i >> 6, k = i & 63;
        float v = 0.f;
        if (e < 5) {
            if (k < 30) v = c_psa[kPsaWl + e * 50 + k];
            else if (k >= kPtcMelK && k < kPtcMelK + 20) v = c_psa[kPsaWl + e * 50 + 30 + (k - kPtcMelK)];
        }
        *reinterpret_cast<__nv_bfloat16 *>(w + (e >> 3) * kT5SBO + (k >> 3) * kT5LBO + (e & 7) * 16 + (k & 7) * 2) = __float2bfloat16_rn(v);
    }
}

__global__ void __launch_bounds__(kPtcThreads, 4)
k_psa_fwd_tc(PsaArgs2 a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) uint64_t bars[2];
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long per_n = (long long)a.B * a.T * 10;
    const long long units = (long long)a.N * per_n;
    const uint32_t bar0 = smem_u32(&bars[0]);

    // ---- one-time setup: barriers, TMEM, zeroed A tiles / pads, the weight tile
    if (tid == 0) {
        mbar_init(bar0, 1); mbar_init(bar0 + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "n"(32) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (uint32_t i = tid * 16; i < kPtcOffZr; i += kPtcThreads * 16) *reinterpret_cast<uint4 *>(smem + i) = make_uint4(0, 0, 0, 0);
    __syncthreads();
    ptc_build_w(smem + kPtcOffW, tid);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base_s;
    const uint32_t smem_base = smem_u32(smem);
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(16 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);

    auto prefetch = [&](long long uu, int sb) {
        if (uu < units) {
            const float *mg = a.melody + uu * 448, *zg = a.Zr + uu * 450;
            float *zs = reinterpret_cast<float *>(smem + kPtcOffZr + sb * 1808), *ms = reinterpret_cast<float *>(smem + kPtcOffMel + sb * 1792);
            if (tid < 112) cp_async16(ms + tid * 4, mg + tid * 4);
            for (int i = tid; i < 225; i += kPtcThreads) cp_async8(zs + i * 2, zg + i * 2);     // Zr rows are 1800 bytes: 8-byte aligned
        }
        asm volatile("cp.async.commit_group;");
    };

    // row owned by this thread inside a tile: two channels x 56 notes
    const int r_ch = tid / 56, r_note = tid - r_ch * 56, r_o = r_note / 7, r_d = r_note - r_o * 7;
    const bool r_valid = tid < 112;
    const uint32_t a_row_off = (uint32_t)(tid >> 3) * kT5SBO + (uint32_t)(tid & 7) * 16;
    const int m_note = (warp & 1) * 32 + lane;                  // this thread's note in the per-unit melody branch
    // (b) the od tasks of this thread: column pair q = tid + 128 i of the pair's 2 x 225 -> where it reads Zr and writes od
    int od_src[4], od_dst[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int q = tid + i * kPtcThreads, ch = q >= 225 ? 1 : 0, qq = q - 225 * ch, j = qq / 15, kk = (qq - j * 15) * 2;
        od_src[i] = 2 * qq;
        od_dst[i] = (int)kPtcOffOd + ch * (int)kPtcOdCh + j * 80 + kk * 2;
    }
    float bl[5];
#pragma unroll
    for (int e = 0; e < 5; e++) bl[e] = c_psa[kPsaBl + e];

    // the tile whose MMAs are in flight (its epilogue runs one iteration later)
    long long p_base0 = 0, p_base1 = 0;     // float offsets of the two channel slabs in `out`
    int p_nch = 0, p_ib = 0;
    uint32_t p_phase = 0;
    bool p_pending = false;
    float *stage = reinterpret_cast<float *>(smem + kPtcOffOut) + warp * 160;

    auto epilogue = [&]() {
        mbar_wait(bar0 + 8 * p_ib, p_phase);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        uint32_t v[8];
        const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(16 * p_ib);
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
#pragma unroll
        for (int e = 0; e < 5; e++) {
            float th;
            asm("tanh.approx.f32 %0, %1;" : "=f"(th) : "f"(0.5f * (__uint_as_float(v[e]) + bl[e])));
            stage[lane * 5 + e] = e == 0 ? fmaf(th, 3.f, 3.f) : fmaf(th, 0.5f, 0.5f);     // sigmoid = 0.5 tanh(x / 2) + 0.5; duration x 6
        }
        __syncwarp();
        // the warp's 32 rows = floats [160 warp, 160 warp + 160) of the two-slab buffer [ch0: 280][ch1: 280]
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int q = lane + 32 * h;                        // float4 index inside the warp's 160 floats
            if (q < 40) {
                const int f = 160 * warp + 4 * q;
                const int ch = f >= 280 ? 1 : 0;
                if (f < 560 && ch < p_nch) {
                    float4 *dst = reinterpret_cast<float4 *>(a.out + (ch ? p_base1 : p_base0) + (f - 280 * ch));
                    *dst = *reinterpret_cast<const float4 *>(stage + 4 * q);
                }
            }
        }
        __syncwarp();
    };

    // the (n, c) rows of Zci a tile adds to the unit's Zr row come from L2: fetched into registers ONE TILE AHEAD so that
    // their latency hides behind the current tile (task i of this thread = column pair q = tid + 128 i of the pair's 2 x 225)
    float2 zc[4];
    auto fetch_zci = [&](int nn, int c0n) {
        const int nchn = min(2, a.C - c0n);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int q = tid + i * kPtcThreads;
            if (q < 225 * nchn) {
                const int ch = q >= 225 ? 1 : 0, qq = q - 225 * ch;
                zc[i] = __ldg(reinterpret_cast<const float2 *>(a.Zci + ((long long)nn * a.C + c0n + ch) * 450 + 2 * qq));
            }
        }
    };
    long long it = 0;
    int sb = 0;
    prefetch(blockIdx.x, 0);
    if ((long long)blockIdx.x < units) fetch_zci((int)(blockIdx.x / per_n), 0);
    for (long long u = blockIdx.x; u < units; u += gridDim.x, sb ^= 1) {
        const int n = (int)(u / per_n);
        const long long btf = u - (long long)n * per_n;
        prefetch(u + gridDim.x, sb ^ 1);
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncthreads();                                        // this unit's staging rows are visible to every thread
        const float *zr = reinterpret_cast<const float *>(smem + kPtcOffZr + sb * 1808);
        const float *mel = reinterpret_cast<const float *>(smem + kPtcOffMel + sb * 1792);
        for (int c0 = 0; c0 < a.C; c0 += 2, it++) {
            const int nch = min(2, a.C - c0), ib = (int)(it & 1);
            // ---- S1: od rows of the pair (bf16), and once per unit the melody branch
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int q = tid + i * kPtcThreads;
                if (q < 225 * nch) {
                    const float2 z = *reinterpret_cast<const float2 *>(zr + od_src[i]);
                    *reinterpret_cast<uint32_t *>(smem + od_dst[i]) = pack_bf16(leaky(z.x + zc[i].x), leaky(z.y + zc[i].y));
                }
            }
            if (c0 + 2 < a.C) fetch_zci(n, c0 + 2);
            else if (u + gridDim.x < units) fetch_zci((int)((u + gridDim.x) / per_n), 0);
            if (c0 == 0 && m_note < 56) {
                // warps 0, 1 -> columns 0..9, warps 2, 3 -> columns 10..19 of the notes m_note = 32 (warp & 1) + lane: the
                // column index is uniform in a warp, so every weight is an immediate constant-bank operand of its FFMA
                const float4 m0 = *reinterpret_cast<const float4 *>(mel + m_note * 8), m1 = *reinterpret_cast<const float4 *>(mel + m_note * 8 + 4);
                const float mv[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
                float mf[10];
                if (warp < 2) {
#pragma unroll
                    for (int j = 0; j < 10; j++) {
                        float acc = c_psa[kPsaBm + j];
#pragma unroll
                        for (int p = 0; p < 8; p++) acc = fmaf(c_psa[kPsaWm + j * 8 + p], mv[p], acc);
                        mf[j] = leaky(acc);
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 10; j++) {
                        float acc = c_psa[kPsaBm + 10 + j];
#pragma unroll
                        for (int p = 0; p < 8; p++) acc = fmaf(c_psa[kPsaWm + (10 + j) * 8 + p], mv[p], acc);
                        mf[j] = leaky(acc);
                    }
                }
                uint32_t *dst = reinterpret_cast<uint32_t *>(smem + kPtcOffM + m_note * 48 + (warp >> 1) * 20);
#pragma unroll
                for (int jj = 0; jj < 5; jj++) dst[jj] = pack_bf16(mf[2 * jj], mf[2 * jj + 1]);
            }
            __syncthreads();
            // ---- S2: this thread's A row (the MMAs that read this A buffer two tiles ago were waited for in the epilogue)
            if (r_valid && r_ch < nch) {
                unsigned char *arow = smem + kPtcOffA + ib * kPtcATile + a_row_off;
                const uint4 *oct = reinterpret_cast<const uint4 *>(smem + kPtcOffOd + r_ch * kPtcOdCh + r_o * 80);
                const uint4 *deg = reinterpret_cast<const uint4 *>(smem + kPtcOffOd + r_ch * kPtcOdCh + (8 + r_d) * 80);
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const uint4 x = oct[j], y = deg[j];
                    *reinterpret_cast<uint4 *>(arow + j * kT5LBO) = make_uint4(bf2_add_leaky(x.x, y.x), bf2_add_leaky(x.y, y.y),
                                                                               bf2_add_leaky(x.z, y.z), bf2_add_leaky(x.w, y.w));
                }
                const uint4 *mrow = reinterpret_cast<const uint4 *>(smem + kPtcOffM + r_note * 48);
#pragma unroll
                for (int j = 0; j < 3; j++) *reinterpret_cast<uint4 *>(arow + (4 + j) * kT5LBO) = mrow[j];
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncthreads();
            // ---- S3: issue this tile's MMAs, then finish the previous tile while they run
            if (tid == 0) {
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t sa = smem_base + kPtcOffA + ib * kPtcATile, sw = smem_base + kPtcOffW;
#pragma unroll
                for (int ks = 0; ks < 4; ks++)
                    b16_mma(tmem + 16 * ib, t5_desc(sa + ks * 2 * kT5LBO, kT5LBO, kT5SBO), t5_desc(sw + ks * 2 * kT5LBO, kT5LBO, kT5SBO), idesc, ks != 0);
                t5_commit(bar0 + 8 * ib);
            }
            __syncwarp();
            if (p_pending) epilogue();
            p_base0 = (((long long)n * a.C + c0) * per_n + btf) * 280;
            p_base1 = (((long long)n * a.C + c0 + 1) * per_n + btf) * 280;
            p_nch = nch; p_ib = ib; p_phase = (uint32_t)((it >> 1) & 1); p_pending = true;
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    if (p_pending) epilogue();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(32) : "memory");
}

}  // namespace mstm

#include "mst_model_loss.cuh"

namespace mstm {

// ------------------------------------------------------------------------------------------------
// Backward of the tail on the tensor cores (replaces k_psa_bwd3 in the bf16 mode; the melody branch stays in
// k_psa_mel_bwd, fed through `dps`).  Per tile of 128 note rows (two channels of one unit):
//   G  [r][e]   = pre-sigmoid loss gradient of the row (loss_cell_grad on the stored prediction and the target), bf16
//   Az [r][k]   = leaky(oct[o] + deg[d]) recomputed as in the forward (k < 30)
//   MMA-1  dA[r][k]   = G . Wl[:, :30]                  (M 128, N 32, K 16)       -> dz = dA * leaky'(z), bf16 tile S
//   MMA-3  gWl^T[k][e] += Az^T . G                      (M 128 (32 used), N 16, K 128 rows; accumulates over ALL tiles of
//                                                        the CTA in TMEM, read once at the end)
//   MMA-2  dOD[(ch, o|d)][k] = Sel . S                   (M 128 (30 used), N 32, K 128 rows): the sums over the 7 degrees /
//                                                        8 octaves of a row group as a product with a constant 0/1 matrix
// All five tiles are the SAME 8 x 16-byte core-matrix images; which one is "K-major" and which "MN-major" is only a matter
// of the descriptor (cute/arch/mma_sm100_desc.hpp: K-major ((8,m),(T,2)):((1T,SBO),(1,LBO)); MN-major
// ((T,1,m),(8,k)):((1,T,SBO),(1T,LBO))), so the thread that owns row r writes its 16-byte chunks once and the tile is
// consumed as A[r][.] (contraction over the chunk axis) by one MMA and as A^T (contraction over r) by another.
// Rows that are not part of the tile (r >= 112, the second channel of an odd pair) hold ZEROS in G and S: a 0 x NaN in a
// contraction over r would poison every output.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t kPbSel = 0;                               // Sel: 4 row groups x 2048 B (groups 4..15 of the M = 128 operand read on into the tiles below: finite or not, they only reach unused output rows)
constexpr uint32_t kPbAz = kPbSel + 4 * 2048;                // Az: 16 row groups x 576 B (4 chunks of 144 B)
constexpr uint32_t kPbAzRG = 576;
constexpr uint32_t kPbG = kPbAz + 16 * kPbAzRG;              // G: [e 0..7] 16 x 128 B, then the all-zero [e 8..15] group
constexpr uint32_t kPbS = kPbG + 2 * 2048;                   // S: 4 k-groups x 2048 B
constexpr uint32_t kPbW = kPbS + 4 * 2048;                   // Wl^T: 4 k-groups x 256 B
constexpr uint32_t kPbOd = kPbW + 4 * 256;                   // od bf16 [2][15][40]
constexpr uint32_t kPbZr = kPbOd + 2 * kPtcOdCh;             // 2 x zr[452]
constexpr uint32_t kPbStage = kPbZr + 2 * 1808;              // 2 buffers x 4 warps x (pred[160] + target[160] floats)
constexpr uint32_t kPbDzr = kPbStage + 2 * 4 * 1280;             // dzr accumulator [450] (+2 pad)
constexpr uint32_t kPbDps = kPbDzr + 1808;                   // dps accumulator [280]
constexpr uint32_t kPbDzi = kPbDps + 1120;                   // dzi accumulators [C][450]
__host__ __device__ constexpr uint32_t ptc_bwd_smem(int C) { return kPbDzi + (uint32_t)C * 1800; }
static_assert(kPbAz % 16 == 0 && kPbG % 16 == 0 && kPbS % 16 == 0 && kPbW % 16 == 0 && kPbOd % 16 == 0 && kPbZr % 16 == 0 &&
              kPbStage % 16 == 0 && kPbDzr % 16 == 0 && kPbDps % 16 == 0 && kPbDzi % 16 == 0, "16-byte aligned regions");

struct PsaBwdTcArgs {
    const float *Zr, *Zci;            // as the forward
    const float *pred, *target;       // psa_out and P: (N,C,B,T,10,56,5)
    const float *coef;                // (N, kLossCoefs)
    float *dZr, *dZi;                 // (N,B,T,10,450) written; (N,C,450) atomically accumulated (zeroed by the caller)
    float *dps;                       // per unit (56, 5) at stride dps_stride: sum over channels of the pre-sigmoid gradient
    long long dps_stride;
    float *gWl;                       // (5, 50): columns 0..29 accumulated here
    int N, C, B, T, ctas_per_n;
};

// smem descriptor with explicit strides (version 1, no swizzle); the two strides mean (LBO, SBO) as the operand's major-ness defines them
__device__ __forceinline__ uint64_t ptc_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) { return t5_desc(saddr, lbo, sbo); }

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t *v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
                   "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
                   "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// dz pair = dA pair * leaky'(z pair): exact select between dA and slope * dA by the sign of z (bf16 bits)
__device__ __forceinline__ uint32_t bf2_leaky_bwd(float da0, float da1, uint32_t z2) {
    const __nv_bfloat162 d = __floats2bfloat162_rn(da0, da1);
    const uint32_t pos = __hgt2_mask(u32_as_bf2(z2), __float2bfloat162_rn(0.f));
    const uint32_t full = bf2_as_u32(d), small = bf2_as_u32(__floats2bfloat162_rn(da0 * kSlope, da1 * kSlope));
    return (full & pos) | (small & ~pos);
}

__global__ void __launch_bounds__(kPtcThreads, 3)
k_psa_bwd_tc(PsaBwdTcArgs a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) uint64_t bars[3];
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int C = a.C;
    const int n = blockIdx.x / a.ctas_per_n, kcta = blockIdx.x - n * a.ctas_per_n;
    const long long per_n = (long long)a.B * a.T * 10;
    const long long chunk = (per_n + a.ctas_per_n - 1) / a.ctas_per_n;
    const long long b_lo = (long long)kcta * chunk, b_hi = min(per_n, b_lo + chunk);
    const uint32_t bar0 = smem_u32(&bars[0]);
    const uint32_t bar1 = bar0, bar2 = bar0 + 8, bar3 = bar0 + 16;
    float *dzr_s = reinterpret_cast<float *>(smem + kPbDzr), *dps_s = reinterpret_cast<float *>(smem + kPbDps);
    float *dzi_s = reinterpret_cast<float *>(smem + kPbDzi);

    if (tid == 0) {
        mbar_init(bar1, 1); mbar_init(bar2, 1); mbar_init(bar3, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "n"(64) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    const uint32_t total = ptc_bwd_smem(C);
    for (uint32_t i = tid * 16; i < total; i += kPtcThreads * 16) *reinterpret_cast<uint4 *>(smem + i) = make_uint4(0, 0, 0, 0);
    __syncthreads();
    // Sel[row = ch * 15 + j][r = ch' * 56 + note] = 1 where ch' == ch and (j < 8 ? note / 7 == j : note % 7 == j - 8); K-major
    for (int i = tid; i < 30 * 112; i += kPtcThreads) {
        const int row = i / 112, r = i - row * 112;
        const int ch = row / 15, j = row - ch * 15, rc = r / 56, note = r - rc * 56;
        const bool one = rc == ch && (j < 8 ? note / 7 == j : note % 7 == j - 8);
        if (one) *reinterpret_cast<__nv_bfloat16 *>(smem + kPbSel + (row >> 3) * 2048 + (r >> 3) * 128 + (row & 7) * 16 + (r & 7) * 2) = __float2bfloat16_rn(1.f);
    }
    // Wl^T[k][e] (k < 30, e < 5), K-major over e
    for (int i = tid; i < 30 * 5; i += kPtcThreads) {
        const int k = i / 5, e = i - k * 5;
        *reinterpret_cast<__nv_bfloat16 *>(smem + kPbW + (k >> 3) * 256 + (k & 7) * 16 + e * 2) = __float2bfloat16_rn(c_psa[kPsaWl + e * 50 + k]);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base_s, smem_base = smem_u32(smem);
    constexpr uint32_t kBase = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(128 >> 4) << 24);
    constexpr uint32_t idesc1 = kBase | ((uint32_t)(32 >> 3) << 17);                             // G (K-major) . Wl^T (K-major)
    constexpr uint32_t idesc2 = kBase | ((uint32_t)(32 >> 3) << 17) | (1u << 16);                // Sel (K-major) . S (MN-major)
    constexpr uint32_t idesc3 = kBase | ((uint32_t)(16 >> 3) << 17) | (1u << 15) | (1u << 16);   // Az^T (MN-major) . G (MN-major)

    float cf[6];
#pragma unroll
    for (int i = 0; i < 6; i++) cf[i] = a.coef[n * kLossCoefs + i];
    const int r_ch = tid / 56, r_note = tid - r_ch * 56, r_o = r_note / 7, r_d = r_note - r_o * 7;
    const bool r_valid = tid < 112;
    unsigned char *az_row = smem + kPbAz + (uint32_t)(tid >> 3) * kPbAzRG + (uint32_t)(tid & 7) * 16;
    unsigned char *g_row = smem + kPbG + (uint32_t)(tid >> 3) * 128 + (uint32_t)(tid & 7) * 16;
    unsigned char *s_row = smem + kPbS + (uint32_t)(tid >> 3) * 128 + (uint32_t)(tid & 7) * 16;
    float *st_pred = reinterpret_cast<float *>(smem + kPbStage) + warp * 320, *st_tgt = st_pred + 160;      // buffer b: + b * 1280 floats

    // everything tile t + 1 needs from global memory travels in ONE cp.async group issued at the start of tile t: its
    // prediction / target rows (per-warp staging, 16-byte coalesced) and, when it opens a new unit, that unit's Zr row
    auto prefetch_tile = [&](long long btf, int c0, int buf, int zbuf) {
        if (btf < b_hi) {
            const int nch = min(2, C - c0);
            const long long base0 = (((long long)n * C + c0) * per_n + btf) * 280, base1 = (((long long)n * C + c0 + 1) * per_n + btf) * 280;
            float *sp = st_pred + buf * 1280, *stg = st_tgt + buf * 1280;
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int q = lane + 32 * h;
                if (q < 40) {
                    const int f = 160 * warp + 4 * q, ch = f >= 280 ? 1 : 0;
                    if (f < 560 && ch < nch) {
                        const long long g = (ch ? base1 : base0) + (f - 280 * ch);
                        cp_async16(sp + 4 * q, a.pred + g);
                        cp_async16(stg + 4 * q, a.target + g);
                    }
                }
            }
            if (c0 == 0) {
                const float *zg = a.Zr + ((long long)n * per_n + btf) * 450;
                float *zs = reinterpret_cast<float *>(smem + kPbZr + zbuf * 1808);
                for (int i = tid; i < 225; i += kPtcThreads) cp_async8(zs + i * 2, zg + i * 2);     // Zr rows are 1800 bytes: 8-byte aligned
            }
        }
        asm volatile("cp.async.commit_group;");
    };
    float2 zc[4];
    auto fetch_zci = [&](int c0n) {
        const int nchn = min(2, C - c0n);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int q = tid + i * kPtcThreads;
            if (q < 225 * nchn) {
                const int ch = q >= 225 ? 1 : 0, qq = q - 225 * ch;
                zc[i] = __ldg(reinterpret_cast<const float2 *>(a.Zci + ((long long)n * C + c0n + ch) * 450 + 2 * qq));
            }
        }
    };

    uint32_t tile = 0;                       // tiles issued so far by this CTA: barrier phase = tile & 1
    int sb = 0;
    prefetch_tile(b_lo, 0, 0, 0);
    fetch_zci(0);
    for (long long btf = b_lo; btf < b_hi; btf++, sb ^= 1) {
        const long long u = (long long)n * per_n + btf;
        const float *zr = reinterpret_cast<const float *>(smem + kPbZr + sb * 1808);
        for (int c0 = 0; c0 < C; c0 += 2, tile++) {
            const int nch = min(2, C - c0);
            const uint32_t ph = tile & 1;
            const int tb = (int)(tile & 1);
            if (c0 + 2 < C) prefetch_tile(btf, c0 + 2, tb ^ 1, sb); else prefetch_tile(btf + 1, 0, tb ^ 1, sb ^ 1);
            const float *sp = st_pred + tb * 1280, *stg = st_tgt + tb * 1280;
            asm volatile("cp.async.wait_group 1;" ::: "memory");    // everything but the group just issued: this tile's rows (and Zr row) landed
            __syncthreads();
            // ---- P1: od rows of the pair (bf16)
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int q = tid + i * kPtcThreads;
                if (q < 225 * nch) {
                    const int ch = q >= 225 ? 1 : 0, qq = q - 225 * ch;
                    const float2 z = *reinterpret_cast<const float2 *>(zr + 2 * qq);
                    const int j = qq / 15, kk = (qq - j * 15) * 2;
                    *reinterpret_cast<uint32_t *>(smem + kPbOd + ch * kPtcOdCh + j * 80 + kk * 2) = pack_bf16(leaky(z.x + zc[i].x), leaky(z.y + zc[i].y));
                }
            }
            if (c0 + 2 < C) fetch_zci(c0 + 2); else fetch_zci(0);
            if (tile > 0) mbar_wait(bar3, (tile - 1) & 1);       // the weight-gradient MMAs of the previous tile have read Az and G
            __syncthreads();
            // ---- P2: loss gradient of the row, its z row; G and Az chunks
            uint4 zreg[4];
            const bool active = r_valid && r_ch < nch;
            if (active) {
                float y[5], t[5], dpre[5];
#pragma unroll
                for (int e = 0; e < 5; e++) { y[e] = sp[lane * 5 + e]; t[e] = stg[lane * 5 + e]; }
                loss_cell_grad<5>(y, t, cf, dpre);
                *reinterpret_cast<uint4 *>(g_row) = make_uint4(pack_bf16(dpre[0], dpre[1]), pack_bf16(dpre[2], dpre[3]), pack_bf16(dpre[4], 0.f), 0u);
#pragma unroll
                for (int e = 0; e < 5; e++) if (dpre[e] != 0.f) atomicAdd(dps_s + r_note * 5 + e, dpre[e]);
                const uint4 *oct = reinterpret_cast<const uint4 *>(smem + kPbOd + r_ch * kPtcOdCh + r_o * 80);
                const uint4 *deg = reinterpret_cast<const uint4 *>(smem + kPbOd + r_ch * kPtcOdCh + (8 + r_d) * 80);
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const uint4 x = oct[j], w = deg[j];
                    zreg[j] = make_uint4(bf2_add_leaky(x.x, w.x), bf2_add_leaky(x.y, w.y), bf2_add_leaky(x.z, w.z), bf2_add_leaky(x.w, w.w));
                    *reinterpret_cast<uint4 *>(az_row + j * kT5LBO) = zreg[j];
                }
            } else if (r_valid) {
                *reinterpret_cast<uint4 *>(g_row) = make_uint4(0, 0, 0, 0);
#pragma unroll
                for (int j = 0; j < 4; j++) zreg[j] = make_uint4(0, 0, 0, 0);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncthreads();
            // ---- P3: dA = G . Wl^T, and the weight gradient
            if (tid == 0) {
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                b16_mma(tmem, ptc_desc(smem_base + kPbG, 2048, 128), ptc_desc(smem_base + kPbW, 128, 256), idesc1, 0);
                t5_commit(bar1);
#pragma unroll
                for (int ks = 0; ks < 8; ks++)
                    b16_mma(tmem + 32, ptc_desc(smem_base + kPbAz + ks * 2 * kPbAzRG, kPbAzRG, kT5LBO), ptc_desc(smem_base + kPbG + ks * 256, 128, 2048),
                            idesc3, (tile | ks) != 0);
                t5_commit(bar3);
            }
            __syncwarp();
            // ---- P4: dz = dA * leaky'(z) -> S
            mbar_wait(bar1, ph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            {
                uint32_t v[32];
                tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16), v);
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                if (r_valid) {
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const uint32_t zz[4] = {zreg[j].x, zreg[j].y, zreg[j].z, zreg[j].w};
                        uint32_t o4[4];
#pragma unroll
                        for (int h = 0; h < 4; h++)
                            o4[h] = active ? bf2_leaky_bwd(__uint_as_float(v[j * 8 + 2 * h]), __uint_as_float(v[j * 8 + 2 * h + 1]), zz[h]) : 0u;
                        *reinterpret_cast<uint4 *>(s_row + j * 2048) = make_uint4(o4[0], o4[1], o4[2], o4[3]);
                    }
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncthreads();
            // ---- P5: dOD = Sel . S
            if (tid == 0) {
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                for (int ks = 0; ks < 8; ks++)
                    b16_mma(tmem, ptc_desc(smem_base + kPbSel + ks * 256, 128, 2048), ptc_desc(smem_base + kPbS + ks * 256, 128, 2048), idesc2, ks != 0);
                t5_commit(bar2);
            }
            __syncwarp();
            // ---- P6 (warp 0 = output rows 0..29): through leaky'(od) into the dZr / dZi accumulators
            if (warp == 0) {
                mbar_wait(bar2, ph);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                uint32_t v[32];
                tmem_ld32(tmem, v);
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                const int ch = lane / 15, j = lane - ch * 15;
                const bool on = lane < 30 && ch < nch;
                const __nv_bfloat16 *odr = reinterpret_cast<const __nv_bfloat16 *>(smem + kPbOd + (on ? ch : 0) * kPtcOdCh + (on ? j : 0) * 80);
                float *zi = dzi_s + (c0 + (on ? ch : 0)) * 450 + (on ? j : 0) * 30;
                float *zrp = dzr_s + (on ? j : 0) * 30;
#pragma unroll
                for (int k = 0; k < 30; k++) {
                    const float g = on ? __uint_as_float(v[k]) * (__bfloat162float(odr[k]) > 0.f ? 1.f : kSlope) : 0.f;
                    v[k] = __float_as_uint(g);
                    if (on) zi[k] += g;
                }
                if (on && ch == 0) {
#pragma unroll
                    for (int k = 0; k < 30; k++) zrp[k] += __uint_as_float(v[k]);
                }
                __syncwarp();
                if (on && ch == 1) {
#pragma unroll
                    for (int k = 0; k < 30; k++) zrp[k] += __uint_as_float(v[k]);
                }
            }
            __syncthreads();
        }
        // ---- unit done: dZr row and the channel sums of the pre-sigmoid gradient go out, accumulators are cleared by their readers
        float *dzr = a.dZr + u * 450, *dps = a.dps + u * a.dps_stride;
        for (int i = tid; i < 225; i += kPtcThreads) {
            *reinterpret_cast<float2 *>(dzr + 2 * i) = *reinterpret_cast<const float2 *>(dzr_s + 2 * i);
            *reinterpret_cast<float2 *>(dzr_s + 2 * i) = make_float2(0.f, 0.f);
        }
        for (int i = tid; i < 280; i += kPtcThreads) { dps[i] = dps_s[i]; dps_s[i] = 0.f; }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    // ---- flush: weight gradient gWl[e][k] (TMEM rows = k, columns = e) and the dZi accumulators
    if (tile > 0 && warp == 0) {
        mbar_wait(bar3, (tile - 1) & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        uint32_t v[8];
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(tmem + 32));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        if (lane < 30) {
#pragma unroll
            for (int e = 0; e < 5; e++) {
                const float g = __uint_as_float(v[e]);
                if (g != 0.f) atomicAdd(a.gWl + e * 50 + lane, g);
            }
        }
    }
    __syncthreads();
    for (int i = tid; i < C * 450; i += kPtcThreads)
        if (dzi_s[i] != 0.f) atomicAdd(a.dZi + (long long)n * C * 450 + i, dzi_s[i]);
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(64) : "memory");
}

}  // namespace mstm
