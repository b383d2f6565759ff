// mst_model_gemm_tma.cuh — the canonical Blackwell GEMM for the one contraction of the model that is large enough to
// deserve it: PitchedChannelsEncoder.linear over the conv features, (N.C.B.T rows) x 456 -> 64 (model.py:57-60, 23 % of
// the step's FLOPs).  bf16 dense mode only: the producer of the activation (k_conv_sparse_fwd) writes a
// bf16 copy, the weight block is cast once per step, and the kernel is fed by tensor-map TMA:
//   * A = activation (M, K) bf16, K-contiguous; B = weight (N, K) bf16, K-contiguous; C = act(A.B^T + bias) fp32;
//   * cp.async.bulk.tensor.2d (SASS UTMALDG) moves 128 x 64 (A) and BN x 64 (B) boxes into a 4-stage shared-memory ring
//     with SWIZZLE_128B (a 64-element bf16 row = one 128-byte swizzle span; out-of-bounds rows / columns arrive as zeros,
//     so K = 456 needs no tail code), completing on the stage's "full" mbarrier by transaction bytes;
//   * one thread issues tcgen05.mma kind::f16 (M 128, N BN, K 16; four per stage, advancing the descriptor start address
//     by 32 bytes inside the swizzle atom) and frees the stage with tcgen05.commit on its "empty" mbarrier;
//   * no thread touches the operands: the only SIMT work is the epilogue (TMEM -> registers -> shared -> coalesced stores).
#pragma once
#include <cuda.h>

#include "mst_model_gemm_b16.cuh"

namespace mstm {

constexpr int kTmaBM = 128, kTmaBK = 64, kTmaStages = 4, kTmaThreads = 128;

struct GemmTmaArgs {
    float *C; const float *bias;
    long long c_si;
    int M, N, K, act, accumulate;
};

// K-major SWIZZLE_128B operand tile: rows of 128 bytes, 8-row groups 1024 bytes apart (SBO), LBO unused
__device__ __forceinline__ uint64_t tma_desc_sw128(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void tma_load_2d(uint32_t smem_dst, const CUtensorMap *map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1) : "memory");
}

template <int BN>
__global__ void __launch_bounds__(kTmaThreads, 2)
k_gemm_b16_tma(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, GemmTmaArgs g) {
    constexpr int BM = kTmaBM, KT = kTmaBK, NS = kTmaStages;
    constexpr uint32_t A_BYTES = BM * KT * 2, B_BYTES = BN * KT * 2, STAGE_BYTES = A_BYTES + B_BYTES;
    constexpr int TM_COLS = BN < 32 ? 32 : BN;
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) uint64_t bars[2 * NS + 1];
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int i0 = blockIdx.y * BM, j0 = blockIdx.x * BN;
    const int n_slabs = (g.K + KT - 1) / KT;
    const uint32_t bar0 = smem_u32(&bars[0]);
    const uint32_t bar_full = bar0, bar_empty = bar0 + 8 * NS, bar_done = bar0 + 16 * NS;
    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < NS; i++) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, 1); }
        mbar_init(bar_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "n"(TM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base_s;
    const uint32_t smem_base = smem_u32(smem);
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

    if (tid == 0) {
        // ---- TMA producer: one thread, NS slabs ahead of the MMAs
        for (int s = 0; s < n_slabs; s++) {
            const int stage = s % NS;
            if (s >= NS) mbar_wait(bar_empty + 8 * stage, (uint32_t)((s / NS - 1) & 1));
            const uint32_t sa = smem_base + stage * STAGE_BYTES, full = bar_full + 8 * stage;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full), "r"(STAGE_BYTES) : "memory");
            tma_load_2d(sa, &map_a, full, s * KT, i0);
            tma_load_2d(sa + A_BYTES, &map_b, full, s * KT, j0);
        }
    } else if (tid == 32) {
        // ---- MMA issuer
        for (int s = 0; s < n_slabs; s++) {
            const int stage = s % NS;
            mbar_wait(bar_full + 8 * stage, (uint32_t)((s / NS) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t sa = smem_base + stage * STAGE_BYTES, sb = sa + A_BYTES;
#pragma unroll
            for (int ks = 0; ks < KT / 16; ks++)
                b16_mma(tmem, tma_desc_sw128(sa + ks * 32), tma_desc_sw128(sb + ks * 32), idesc, (s | ks) != 0);
            t5_commit(bar_empty + 8 * stage);
            if (s + 1 == n_slabs) t5_commit(bar_done);
        }
    }
    __syncwarp();
    mbar_wait(bar_done, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    // ---- epilogue: warp w owns TMEM lanes 32 w .. (= tile rows); TMEM -> registers -> shared tile -> coalesced stores
    constexpr int LD = BN + 4;
    float *tile = reinterpret_cast<float *>(smem);            // the operand ring is idle: every MMA has completed
    {
        const int r = warp * 32 + lane;
#pragma unroll
        for (int cb = 0; cb < BN; cb += 8) {
            uint32_t v[8];
            const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)cb;
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            float4 *d = reinterpret_cast<float4 *>(tile + r * LD + cb);
            d[0] = make_float4(__uint_as_float(v[0]), __uint_as_float(v[1]), __uint_as_float(v[2]), __uint_as_float(v[3]));
            d[1] = make_float4(__uint_as_float(v[4]), __uint_as_float(v[5]), __uint_as_float(v[6]), __uint_as_float(v[7]));
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TM_COLS) : "memory");
    const bool vec_ok = (g.c_si & 3) == 0 && ((reinterpret_cast<uintptr_t>(g.C) & 15) == 0) && (g.N & 3) == 0;
    for (int c = tid; c < BM * (BN / 4); c += kTmaThreads) {
        const int r = c / (BN / 4), c4 = (c - r * (BN / 4)) * 4;
        const int gi = i0 + r, gj = j0 + c4;
        if (gi >= g.M || gj >= g.N) continue;
        const float4 x = *reinterpret_cast<const float4 *>(tile + r * LD + c4);
        float o[4] = {x.x, x.y, x.z, x.w};
        float *dst = g.C + (long long)gi * g.c_si + gj;
#pragma unroll
        for (int e = 0; e < 4; e++) {
            if (gj + e < g.N) {
                if (g.bias) o[e] += g.bias[gj + e];
                if (g.act == 1) o[e] = leaky(o[e]);
                if (g.accumulate == 1) o[e] += dst[e];
            }
        }
        if (vec_ok && gj + 3 < g.N) *reinterpret_cast<float4 *>(dst) = make_float4(o[0], o[1], o[2], o[3]);
        else for (int e = 0; e < 4 && gj + e < g.N; e++) dst[e] = o[e];
    }
}

template <int BN> constexpr size_t tma_gemm_smem_bytes() {
    return (size_t)kTmaStages * (kTmaBM * kTmaBK * 2 + BN * kTmaBK * 2) > (size_t)kTmaBM * (BN + 4) * 4
               ? (size_t)kTmaStages * (kTmaBM * kTmaBK * 2 + BN * kTmaBK * 2) : (size_t)kTmaBM * (BN + 4) * 4;
}

// fp32 (rows, cols) at row stride ld -> bf16 (rows, cols_pad) at row stride ld16 (columns >= cols zero-filled)
__global__ void __launch_bounds__(256)
k_cast_bf16(const float *__restrict__ src, long long ld, __nv_bfloat16 *__restrict__ dst, long long ld16, long long rows, int cols, int cols_pad) {
    const long long total = rows * cols_pad;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / cols_pad;
        const int c = (int)(i - r * cols_pad);
        dst[r * ld16 + c] = __float2bfloat16_rn(c < cols ? src[r * ld + c] : 0.f);
    }
}

// host: a 2-D bf16 tensor map (inner dimension K contiguous), 128-byte swizzle, box {64, box_rows}
typedef CUresult (*PFN_tensorMapEncodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                             const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline PFN_tensorMapEncodeTiled tensor_map_encoder() {
    static PFN_tensorMapEncodeTiled fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void *p = nullptr;
        cudaDriverEntryPointQueryResult st;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &st) == cudaSuccess && st == cudaDriverEntryPointSuccess)
            fn = (PFN_tensorMapEncodeTiled)p;
        else cudaGetLastError();
    }
    return fn;
}
inline bool make_bf16_map(CUtensorMap *map, const void *base, long long rows, long long cols, long long row_stride_elems, int box_rows) {
    PFN_tensorMapEncodeTiled enc = tensor_map_encoder();
    if (!enc) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)row_stride_elems * 2};
    const cuuint32_t box[2] = {(cuuint32_t)kTmaBK, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace mstm
