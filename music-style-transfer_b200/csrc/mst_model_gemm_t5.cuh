// mst_model_gemm_t5.cuh — the row-space GEMMs on the 5th-generation tensor cores (tcgen05.mma, accumulator in TMEM).
//
// Same contract as k_gemm / k_gemm_tc (GemmArgs: explicit strides, bias / leaky epilogue, store / += / atomic split-K).
// fp32 accuracy is kept with the 3-term TF32 split  a.b ~ a_lo.b_hi + a_hi.b_lo + a_hi.b_hi  (a_hi = tf32(a), a_lo =
// tf32(a - a_hi)); the legacy mma.sync path tops out near 120 TFLOP/s of TF32 on this part (measured, tools/mma_probe.py),
// which with three products per term made k_gemm_tc MMA-bound at ~2 TB/s of operand traffic.  tcgen05 removes that
// bound: the kernel is a streaming producer (global -> split -> shared) in front of one MMA-issuing thread.
//
//   * CTA tile 128 x BN (BN = 16 / 32 / 64), K slab 32, a ring of NS shared-memory stages (NS = 2 is what the host
//     launches, two CTAs per SM); each stage holds A_hi, A_lo (128 x 32)
//     and B_hi, B_lo (BN x 32) in the canonical K-major no-swizzle UMMA layout: 8-row x 16-byte core matrices, the 8
//     K-chunks of a row group 144 B apart (LBO), row groups 1152 B apart (SBO);
//   * 8 producer warps keep the next TWO slabs in registers (any operand strides; the lane order follows the operand's
//     contiguous direction, 16 / 8-byte loads when aligned), split it, store 16-byte chunks (conflict-free for both
//     orders thanks to the 144-byte LBO), fence.proxy.async, and arrive on the stage's "full" mbarrier (no CTA barrier);
//     a ninth warp's elected thread waits on it, issues 4 k-steps x 3 products of tcgen05.mma.kind::tf32 (M = 128,
//     N = BN, K = 8) and commits to the stage's "empty" mbarrier, which gates the reuse of that stage two slabs later;
//   * epilogue: tcgen05.ld (32 lanes x 32 bit, warp w reads TMEM lanes 32 (w % 4) ...) -> shared tile -> coalesced
//     stores with bias / leaky / accumulate (16-byte where the destination allows).
#pragma once
#include "mst_model_kernels.cuh"

namespace mstm {

constexpr int kT5BM = 128, kT5BK = 32, kT5Threads = 256, kT5Block = kT5Threads + 32;   // 8 loader warps + the MMA warp

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0, spins = 0;
    while (!done) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (!done && ++spins > (1u << 26)) __trap();      // a lost arrival must fail loudly, not hang the device
    }
}
// K-major, no swizzle: start address, LBO (between the two 16-byte K chunks of one MMA), SBO (between 8-row groups)
__device__ __forceinline__ uint64_t t5_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) |
           (1ull << 46);                                   // descriptor version 1 (sm_100), layout type 0 = no swizzle
}
__device__ __forceinline__ void t5_mma(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void t5_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// x = hi + lo exactly: hi = tf32_rna(x) (mst_model_kernels.cuh), lo = x - hi.  The tensor core reads only the tf32 bits of lo, i.e. truncates it:
// |lo| <= 2^-11 |x|, so the dropped part is below 2^-21 |x|.
__device__ __forceinline__ void t5_split4(const float *v, float4 &hi, float4 &lo) {
    hi = make_float4(tf32_rna(v[0]), tf32_rna(v[1]), tf32_rna(v[2]), tf32_rna(v[3]));
    lo = make_float4(v[0] - hi.x, v[1] - hi.y, v[2] - hi.z, v[3] - hi.w);
}

// shared layout of one (rows x 32) operand tile: 16-byte chunk (row r, k-chunk kc) at (r / 8) * SBO + kc * LBO + (r % 8) * 16.
// LBO = 144 (not 128) so that BOTH loader mappings store without bank conflicts: 8 lanes on one row (kc = 0..7) hit
// kc * 144 = kc * 16 mod 128, 8 lanes on consecutive rows hit r * 16.
constexpr uint32_t kT5LBO = 144, kT5SBO = 8 * kT5LBO;
__host__ __device__ constexpr int t5_tile_bytes(int rows) { return rows / 8 * (int)kT5SBO; }

// operand access modes (chosen on the host from strides and alignment, compile-time in the kernel so that the slab loop
// holds exactly the loads it executes): 0 = k-contiguous, 16-byte aligned rows (one 16-byte load per chunk), 1 =
// k-contiguous, 8-byte aligned rows (two 8-byte loads), 2 = any strides (four 4-byte loads `sk` apart; coalesced over
// the lanes in whichever direction is contiguous)
template <int MODE>
__device__ __forceinline__ void t5_load_full(const float *p, long long sk, float *v) {
    if (MODE == 0) {
        const float4 x = __ldg(reinterpret_cast<const float4 *>(p));
        v[0] = x.x; v[1] = x.y; v[2] = x.z; v[3] = x.w;
    } else if (MODE == 1) {
        const float2 x = __ldg(reinterpret_cast<const float2 *>(p)), y = __ldg(reinterpret_cast<const float2 *>(p + 2));
        v[0] = x.x; v[1] = x.y; v[2] = y.x; v[3] = y.y;
    } else {
#pragma unroll
        for (int e = 0; e < 4; e++) v[e] = __ldg(p + e * sk);
    }
}
// the last, partial slab: zero fill at and beyond k_hi
__device__ __forceinline__ void t5_load_tail(const float *p, long long sk, int k, int k_hi, float *v) {
#pragma unroll
    for (int e = 0; e < 4; e++) v[e] = k + e < k_hi ? __ldg(p + e * sk) : 0.f;
}

template <int BN, int AMODE, int BMODE, int NS>
__global__ void __launch_bounds__(kT5Block, NS == 2 ? 2 : 1)
k_gemm_t5(GemmArgs g) {
    constexpr int BM = kT5BM, KT = kT5BK;
    constexpr int A_BYTES = t5_tile_bytes(BM), B_BYTES = t5_tile_bytes(BN);   // one hi or lo tile
    constexpr int STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;
    constexpr int A_CH = BM * 8 / kT5Threads;                                  // 16-byte chunks per thread: 4
    constexpr int B_CH = (BN * 8 + kT5Threads - 1) / kT5Threads;               // 2 / 1 / 1 (BN = 16: threads < 128)
    constexpr int TM_COLS = BN < 32 ? 32 : BN;
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) uint64_t bars[2 * NS + 1]; // full[NS] (8 loader warps), empty[NS] (tcgen05.commit), done
    __shared__ uint32_t tmem_base_s;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int i0 = blockIdx.y * BM, j0 = blockIdx.x * BN;
    const int kz = gridDim.z;
    const int k_per = ((g.K + kz - 1) / kz + KT - 1) / KT * KT;
    const int k_lo = blockIdx.z * k_per, k_hi = min(g.K, k_lo + k_per);
    if (k_lo >= k_hi) return;
    const int n_slabs = (k_hi - k_lo + KT - 1) / KT;

    const uint32_t bar0 = smem_u32(&bars[0]);
    const uint32_t bar_full = bar0, bar_empty = bar0 + 8 * NS, bar_done = bar0 + 16 * NS;
    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < NS; i++) { mbar_init(bar_full + 8 * i, 8); mbar_init(bar_empty + 8 * i, 1); }
        mbar_init(bar_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "n"(TM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base_s;

    // ---- loader roles.  A k-contiguous operand is walked chunk-fastest (8 lanes = one row's 128 bytes), a row-contiguous
    // one row-fastest (32 lanes = 32 consecutive rows of one k); either way thread-chunk i is (row, kc) below.
    // Per-thread chunk slots: global pointer of the chunk's first element in the current slab (rows clamped into the
    // matrix: rows / columns past the edge compute garbage that the epilogue never stores), its k offset inside a slab
    // and its shared-memory offset.
    const bool a_kc = g.a_sk == 1, b_kc = g.b_sk == 1;
    const float *pa[A_CH], *pb[B_CH];
    int a_k[A_CH], b_k[B_CH];
    uint32_t a_off[A_CH], b_off[B_CH];
#pragma unroll
    for (int i = 0; i < A_CH; i++) {
        const int c = tid + i * kT5Threads;
        int row, kc4;
        if (AMODE < 2 || a_kc) { kc4 = c & 7; row = c >> 3; } else { row = c & (BM - 1); kc4 = c / BM; }
        a_k[i] = kc4 * 4;
        a_off[i] = (uint32_t)(row >> 3) * kT5SBO + (uint32_t)kc4 * kT5LBO + (uint32_t)(row & 7) * 16;
        pa[i] = g.A + (long long)min(i0 + row, g.M - 1) * g.a_si + (long long)(k_lo + kc4 * 4) * g.a_sk;
    }
#pragma unroll
    for (int i = 0; i < B_CH; i++) {
        const int c = min(tid + i * kT5Threads, BN * 8 - 1);          // BN = 16: the upper threads repeat the last chunk
        int row, kc4;
        if (BMODE < 2 || b_kc) { kc4 = c & 7; row = c >> 3; } else { row = c % BN; kc4 = c / BN; }
        b_k[i] = kc4 * 4;
        b_off[i] = (uint32_t)(2 * A_BYTES) + (uint32_t)(row >> 3) * kT5SBO + (uint32_t)kc4 * kT5LBO + (uint32_t)(row & 7) * 16;
        pb[i] = g.B + (long long)(k_lo + kc4 * 4) * g.b_sk + (long long)min(j0 + row, g.N - 1) * g.b_sj;
    }
    const long long a_step = (long long)KT * g.a_sk, b_step = (long long)KT * g.b_sk;
    const int n_full = (k_hi - k_lo) / KT;                            // slabs that need no zero fill
    // two register sets: slab s lives in set s & 1 and is fetched TWO slabs ahead of its store, so that every loader
    // thread keeps two slabs of global loads in flight (one slab per thread is ~32 KB per SM, short of the
    // latency x bandwidth product of HBM)
    float ra[2][A_CH][4], rb[2][B_CH][4];
    auto load_slab = [&](int t, float (&xa)[A_CH][4], float (&xb)[B_CH][4]) {
        if (t < n_full) {
#pragma unroll
            for (int i = 0; i < A_CH; i++) { t5_load_full<AMODE>(pa[i], g.a_sk, xa[i]); pa[i] += a_step; }
#pragma unroll
            for (int i = 0; i < B_CH; i++) { t5_load_full<BMODE>(pb[i], g.b_sk, xb[i]); pb[i] += b_step; }
        } else {
            const int kb = k_lo + n_full * KT;
#pragma unroll
            for (int i = 0; i < A_CH; i++) t5_load_tail(pa[i], g.a_sk, kb + a_k[i], k_hi, xa[i]);
#pragma unroll
            for (int i = 0; i < B_CH; i++) t5_load_tail(pb[i], g.b_sk, kb + b_k[i], k_hi, xb[i]);
        }
    };
    auto store_slab = [&](int stage, float (&xa)[A_CH][4], float (&xb)[B_CH][4]) {
        unsigned char *st = smem + stage * STAGE_BYTES;
#pragma unroll
        for (int i = 0; i < A_CH; i++) {
            float4 hi, lo;
            t5_split4(xa[i], hi, lo);
            *reinterpret_cast<float4 *>(st + a_off[i]) = hi;
            *reinterpret_cast<float4 *>(st + A_BYTES + a_off[i]) = lo;
        }
#pragma unroll
        for (int i = 0; i < B_CH; i++) {
            if (BN * 8 >= kT5Threads * (i + 1) || tid + i * kT5Threads < BN * 8) {
                float4 hi, lo;
                t5_split4(xb[i], hi, lo);
                *reinterpret_cast<float4 *>(st + b_off[i]) = hi;
                *reinterpret_cast<float4 *>(st + B_BYTES + b_off[i]) = lo;
            }
        }
    };
    // instruction descriptor: D = f32, A = B = tf32, both K-major, N >> 3 at bit 17, M >> 4 at bit 24
    constexpr uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    const uint32_t smem_base = smem_u32(smem);

    if (warp < 8) {
        // ---- producers: split + store slab s, publish it on the stage's "full" barrier, then fetch slab s + 2
        auto produce = [&](int s, float (&xa)[A_CH][4], float (&xb)[B_CH][4]) {
            const int stage = s % NS;
            if (s >= NS) mbar_wait(bar_empty + 8 * stage, (uint32_t)((s / NS - 1) & 1));    // MMAs of slab s - NS have read it
            store_slab(stage, xa, xb);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_full + 8 * stage) : "memory");
            if (s + 2 < n_slabs) load_slab(s + 2, xa, xb);
        };
        load_slab(0, ra[0], rb[0]);
        if (n_slabs > 1) load_slab(1, ra[1], rb[1]);
        for (int s = 0; s < n_slabs; s += 2) {
            produce(s, ra[0], rb[0]);
            if (s + 1 < n_slabs) produce(s + 1, ra[1], rb[1]);
        }
    } else if (lane == 0) {
        // ---- consumer: one thread issues the MMAs of a slab as soon as all 8 producer warps have published it
        for (int s = 0; s < n_slabs; s++) {
            const int stage = s % NS;
            mbar_wait(bar_full + 8 * stage, (uint32_t)((s / NS) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t sa = smem_base + stage * STAGE_BYTES, sb = sa + 2 * A_BYTES;
#pragma unroll
            for (int ks = 0; ks < KT / 8; ks++) {
                const uint32_t ko = ks * 2 * kT5LBO;
                const uint64_t a_hi = t5_desc(sa + ko, kT5LBO, kT5SBO), a_lo = t5_desc(sa + A_BYTES + ko, kT5LBO, kT5SBO);
                const uint64_t b_hi = t5_desc(sb + ko, kT5LBO, kT5SBO), b_lo = t5_desc(sb + B_BYTES + ko, kT5LBO, kT5SBO);
                t5_mma(tmem, a_lo, b_hi, idesc, (s | ks) != 0);
                t5_mma(tmem, a_hi, b_lo, idesc, 1);
                t5_mma(tmem, a_hi, b_hi, idesc, 1);
            }
            t5_commit(bar_empty + 8 * stage);
            if (s + 1 == n_slabs) t5_commit(bar_done);
        }
    }
    __syncwarp();
    mbar_wait(bar_done, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    // ---- epilogue: TMEM -> registers (warp w reads lanes 32 (w % 4) .. + 31 = tile rows; warps 0-3 / 4-7 split the
    // columns) -> shared tile [128][BN + 4] -> coalesced global stores with bias / leaky / accumulate
    constexpr int LD = BN + 4;
    float *tile = reinterpret_cast<float *>(smem);            // the operand stages are idle now (all MMAs completed)
    constexpr int CW = BN >= 32 ? BN / 2 : BN;                // columns per warp (BN = 16: warps 4-7 idle)
    const int col0 = (BN >= 32 && warp >= 4) ? CW : 0;
    if (warp < 8 && (BN >= 32 || warp < 4)) {
        const int r = (warp & 3) * 32 + lane;
#pragma unroll
        for (int cb = 0; cb < CW; cb += 8) {
            uint32_t v[8];
            const uint32_t taddr = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(col0 + cb);
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            float4 *d = reinterpret_cast<float4 *>(tile + r * LD + col0 + cb);
            d[0] = make_float4(__uint_as_float(v[0]), __uint_as_float(v[1]), __uint_as_float(v[2]), __uint_as_float(v[3]));
            d[1] = make_float4(__uint_as_float(v[4]), __uint_as_float(v[5]), __uint_as_float(v[6]), __uint_as_float(v[7]));
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TM_COLS) : "memory");
    const bool vec_ok = g.accumulate != 2 && (g.c_si & 3) == 0 && ((reinterpret_cast<uintptr_t>(g.C) & 15) == 0);
    if (vec_ok) {
        for (int c = tid; c < BM * (BN / 4); c += kT5Block) {
            const int r = c / (BN / 4), c4 = (c - r * (BN / 4)) * 4;
            const int gi = i0 + r, gj = j0 + c4;
            if (gi >= g.M || gj >= g.N) continue;
            float4 x = *reinterpret_cast<const float4 *>(tile + r * LD + c4);
            float o[4] = {x.x, x.y, x.z, x.w};
            float *dst = g.C + (long long)gi * g.c_si + gj;
            if (gj + 3 < g.N) {
                if (g.bias && blockIdx.z == 0) { o[0] += g.bias[gj]; o[1] += g.bias[gj + 1]; o[2] += g.bias[gj + 2]; o[3] += g.bias[gj + 3]; }
                if (g.act == 1) { o[0] = leaky(o[0]); o[1] = leaky(o[1]); o[2] = leaky(o[2]); o[3] = leaky(o[3]); }
                float4 *d4 = reinterpret_cast<float4 *>(dst);
                if (g.accumulate == 1) { const float4 p = *d4; o[0] += p.x; o[1] += p.y; o[2] += p.z; o[3] += p.w; }
                *d4 = make_float4(o[0], o[1], o[2], o[3]);
            } else {
                for (int e = 0; e < 4 && gj + e < g.N; e++) {
                    float v = o[e];
                    if (g.bias && blockIdx.z == 0) v += g.bias[gj + e];
                    if (g.act == 1) v = leaky(v);
                    dst[e] = g.accumulate == 1 ? dst[e] + v : v;
                }
            }
        }
    } else {
        for (int c = tid; c < BM * BN; c += kT5Block) {
            const int r = c / BN, cc = c - r * BN;
            const int gi = i0 + r, gj = j0 + cc;
            if (gi >= g.M || gj >= g.N) continue;
            float v = tile[r * LD + cc];
            if (g.bias && blockIdx.z == 0) v += g.bias[gj];
            if (g.act == 1) v = leaky(v);
            float *dst = g.C + (long long)gi * g.c_si + gj;
            if (g.accumulate == 0) *dst = v;
            else if (g.accumulate == 1) *dst += v;
            else atomicAdd(dst, v);
        }
    }
}

template <int BN, int NS> constexpr size_t t5_smem_bytes() { return NS * (2 * t5_tile_bytes(kT5BM) + 2 * t5_tile_bytes(BN)); }

}  // namespace mstm
