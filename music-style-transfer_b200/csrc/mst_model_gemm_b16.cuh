// mst_model_gemm_b16.cuh — the row-space GEMMs in the bf16 dense mode (BASELINE.json configs[3]: "bf16 dense layers"):
// tcgen05.mma kind::f16 with bf16 operands and the fp32 accumulator in TMEM.
//
// Same contract and the same warp-specialised shape as k_gemm_t5 (mst_model_gemm_t5.cuh): 8 producer warps, one MMA-issuing
// thread, mbarrier ring, TMEM -> registers -> shared -> coalesced epilogue.  What the bf16 mode changes:
//   * the operands stay fp32 in HBM (every SIMT consumer of the activations reads fp32) and are rounded to bf16 by the
//     producers (cvt.rn.bf16x2.f32: one instruction per two elements instead of the five of the 3-term TF32 split), so
//     a stage holds ONE tile per operand instead of a hi and a lo tile, and one MMA per k-step replaces three;
//   * the K slab is 64 (64 bf16 = 128 bytes = the same eight 16-byte chunks per row as a 32-wide tf32 slab): the
//     canonical K-major no-swizzle layout, the 144-byte LBO and all shared-memory offsets carry over unchanged;
//   * one slab of global loads (48 registers) is in flight per producer thread - the same bytes as the two 32-wide slabs
//     of the fp32 kernel.
// Accuracy: bf16 operands (8-bit mantissa), fp32 accumulation: the tolerance of the mode is 2e-2 on gradients / 1e-3 on
// losses, tested in tests/test_gpu_train.py.
#pragma once
#include <cuda_bf16.h>

#include "mst_model_gemm_t5.cuh"

namespace mstm {

constexpr int kB16BK = 64;

__device__ __forceinline__ void b16_mma(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);      // .x (low half, lower address) = lo
    return *reinterpret_cast<const uint32_t *>(&v);
}
__device__ __forceinline__ uint4 pack_bf16x8(const float *v) {
    return make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
}

// eight consecutive k of one row: MODE 0 = 16-byte aligned k-contiguous rows, 1 = 8-byte aligned, 2 = any strides
template <int MODE>
__device__ __forceinline__ void b16_load_full(const float *p, long long sk, float *v) {
    if (MODE == 0) {
        const float4 x = __ldg(reinterpret_cast<const float4 *>(p)), y = __ldg(reinterpret_cast<const float4 *>(p + 4));
        v[0] = x.x; v[1] = x.y; v[2] = x.z; v[3] = x.w; v[4] = y.x; v[5] = y.y; v[6] = y.z; v[7] = y.w;
    } else if (MODE == 1) {
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const float2 x = __ldg(reinterpret_cast<const float2 *>(p + 2 * e));
            v[2 * e] = x.x; v[2 * e + 1] = x.y;
        }
    } else {
#pragma unroll
        for (int e = 0; e < 8; e++) v[e] = __ldg(p + e * sk);
    }
}
__device__ __forceinline__ void b16_load_tail(const float *p, long long sk, int k, int k_hi, float *v) {
#pragma unroll
    for (int e = 0; e < 8; e++) v[e] = k + e < k_hi ? __ldg(p + e * sk) : 0.f;
}

template <int BN, int AMODE, int BMODE>
__global__ void __launch_bounds__(kT5Block, 2)
k_gemm_b16(GemmArgs g) {
    constexpr int BM = kT5BM, KT = kB16BK, NS = 2;
    constexpr int A_BYTES = t5_tile_bytes(BM), B_BYTES = t5_tile_bytes(BN);
    constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    constexpr int A_CH = BM * 8 / kT5Threads;                                  // 16-byte (8-element) chunks per thread: 4
    constexpr int B_CH = (BN * 8 + kT5Threads - 1) / kT5Threads;               // 2 / 1 / 1
    constexpr int TM_COLS = BN < 32 ? 32 : BN;
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) uint64_t bars[2 * NS + 1];
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

    const bool a_kc = g.a_sk == 1, b_kc = g.b_sk == 1;
    const float *pa[A_CH], *pb[B_CH];
    int a_k[A_CH], b_k[B_CH];
    uint32_t a_off[A_CH], b_off[B_CH];
#pragma unroll
    for (int i = 0; i < A_CH; i++) {
        const int c = tid + i * kT5Threads;
        int row, kc8;
        if (AMODE < 2 || a_kc) { kc8 = c & 7; row = c >> 3; } else { row = c & (BM - 1); kc8 = c / BM; }
        a_k[i] = kc8 * 8;
        a_off[i] = (uint32_t)(row >> 3) * kT5SBO + (uint32_t)kc8 * kT5LBO + (uint32_t)(row & 7) * 16;
        pa[i] = g.A + (long long)min(i0 + row, g.M - 1) * g.a_si + (long long)(k_lo + kc8 * 8) * g.a_sk;
    }
#pragma unroll
    for (int i = 0; i < B_CH; i++) {
        const int c = min(tid + i * kT5Threads, BN * 8 - 1);
        int row, kc8;
        if (BMODE < 2 || b_kc) { kc8 = c & 7; row = c >> 3; } else { row = c % BN; kc8 = c / BN; }
        b_k[i] = kc8 * 8;
        b_off[i] = (uint32_t)A_BYTES + (uint32_t)(row >> 3) * kT5SBO + (uint32_t)kc8 * kT5LBO + (uint32_t)(row & 7) * 16;
        pb[i] = g.B + (long long)(k_lo + kc8 * 8) * g.b_sk + (long long)min(j0 + row, g.N - 1) * g.b_sj;
    }
    const long long a_step = (long long)KT * g.a_sk, b_step = (long long)KT * g.b_sk;
    const int n_full = (k_hi - k_lo) / KT;
    float ra[A_CH][8], rb[B_CH][8];
    auto load_slab = [&](int t) {
        if (t < n_full) {
#pragma unroll
            for (int i = 0; i < A_CH; i++) { b16_load_full<AMODE>(pa[i], g.a_sk, ra[i]); pa[i] += a_step; }
#pragma unroll
            for (int i = 0; i < B_CH; i++) { b16_load_full<BMODE>(pb[i], g.b_sk, rb[i]); pb[i] += b_step; }
        } else {
            const int kb = k_lo + n_full * KT;
#pragma unroll
            for (int i = 0; i < A_CH; i++) b16_load_tail(pa[i], g.a_sk, kb + a_k[i], k_hi, ra[i]);
#pragma unroll
            for (int i = 0; i < B_CH; i++) b16_load_tail(pb[i], g.b_sk, kb + b_k[i], k_hi, rb[i]);
        }
    };
    // instruction descriptor: D = f32 (bit 4), A = B = bf16 (1 at bits 7 and 10), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    const uint32_t smem_base = smem_u32(smem);

    if (warp < 8) {
        load_slab(0);
        for (int s = 0; s < n_slabs; s++) {
            const int stage = s % NS;
            if (s >= NS) mbar_wait(bar_empty + 8 * stage, (uint32_t)((s / NS - 1) & 1));
            unsigned char *st = smem + stage * STAGE_BYTES;
#pragma unroll
            for (int i = 0; i < A_CH; i++) *reinterpret_cast<uint4 *>(st + a_off[i]) = pack_bf16x8(ra[i]);
#pragma unroll
            for (int i = 0; i < B_CH; i++)
                if (BN * 8 >= kT5Threads * (i + 1) || tid + i * kT5Threads < BN * 8) *reinterpret_cast<uint4 *>(st + b_off[i]) = pack_bf16x8(rb[i]);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_full + 8 * stage) : "memory");
            if (s + 1 < n_slabs) load_slab(s + 1);
        }
    } else if (lane == 0) {
        for (int s = 0; s < n_slabs; s++) {
            const int stage = s % NS;
            mbar_wait(bar_full + 8 * stage, (uint32_t)((s / NS) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t sa = smem_base + stage * STAGE_BYTES, sb = sa + A_BYTES;
#pragma unroll
            for (int ks = 0; ks < KT / 16; ks++) {                // one MMA = K 16 = two 16-byte chunks, LBO apart
                const uint32_t ko = ks * 2 * kT5LBO;
                b16_mma(tmem, t5_desc(sa + ko, kT5LBO, kT5SBO), t5_desc(sb + ko, kT5LBO, kT5SBO), idesc, (s | ks) != 0);
            }
            t5_commit(bar_empty + 8 * stage);
            if (s + 1 == n_slabs) t5_commit(bar_done);
        }
    }
    __syncwarp();
    mbar_wait(bar_done, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    // ---- epilogue (as k_gemm_t5): TMEM -> registers -> shared tile [128][BN + 4] -> coalesced stores
    constexpr int LD = BN + 4;
    float *tile = reinterpret_cast<float *>(smem);
    constexpr int CW = BN >= 32 ? BN / 2 : BN;
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

template <int BN> constexpr size_t b16_smem_bytes() {
    // two operand stages, or the [128][BN + 4] fp32 epilogue tile that reuses them
    return 2 * (t5_tile_bytes(kT5BM) + t5_tile_bytes(BN)) > (size_t)kT5BM * (BN + 4) * 4 ? 2 * (t5_tile_bytes(kT5BM) + t5_tile_bytes(BN))
                                                                                        : (size_t)kT5BM * (BN + 4) * 4;
}

}  // namespace mstm
