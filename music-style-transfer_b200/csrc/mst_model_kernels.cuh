// mst_model_kernels.cuh — device kernels of the style-transfer network's training step (fp32 path).
//
// The reference (style/model.py) is a stack of small nn.Linear / nn.LSTM / Conv1d layers glued by
// cat_with_broadcast (style/utils/pytorch.py:54-65) and `combine` (model.py:796-815).  Here every
// "Linear over a broadcast concat" is decomposed: W·cat(x1..xk) = sum_s W_s·x_s, each term computed at the
// source's own (small) row space by one GEMM and summed by a broadcast-add kernel, so the concatenated
// tensors (rows x 135, rows x 88, note-rows x 50 ...) never exist.  The backward pass mirrors it: the
// pre-activation gradient is reduced over the broadcast axes and each source sees a plain GEMM.
#pragma once
#include "mst_device.cuh"

namespace mstm {

constexpr float kSlope = 0.01f;   // F.leaky_relu default negative_slope, used everywhere in the reference

__device__ __forceinline__ float leaky(float x) { return x > 0.f ? x : kSlope * x; }
// derivative from the OUTPUT (sign preserved by a positive slope; torch uses slope at x == 0)
__device__ __forceinline__ float leaky_grad_from_out(float y) { return y > 0.f ? 1.f : kSlope; }
__device__ __forceinline__ float sigmoidf(float x) { return 1.f / (1.f + expf(-x)); }

// ---- packed fp32 pairs (sm_100 FFMA2 / FADD2 / FMUL2): one issue slot carries two fp32 operations.  These kernels are
// issue-bound, not FMA-pipe-bound (profiles/r1_ffma2_issue_probe.txt), so the pairs run over the contraction index and
// the two partial sums are added once at the end.  A pair is a 64-bit register: .x = low word, .y = high word.
// fp32 -> tf32, round to nearest with ties away (add half an ulp of the 10-bit mantissa to the bit pattern, clear the 13
// low bits): what cvt.rna.tf32.f32 computes for finite x, in two integer instructions where the cvt expands to four
// (add, |x| < inf test, select, mask) on sm_100
__device__ __forceinline__ float tf32_rna(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }

typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk2(f32x2 v, float &lo, float &hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float sum2(f32x2 v) { float lo, hi; upk2(v, lo, hi); return lo + hi; }
// leaky_relu of both halves: max(x, slope * x) equals the select form for every finite x (0 < slope < 1)
__device__ __forceinline__ f32x2 leaky2(f32x2 x) {
    float xl, xh, tl, th;
    upk2(x, xl, xh);
    upk2(mul2(x, pk2(kSlope, kSlope)), tl, th);
    return pk2(fmaxf(xl, tl), fmaxf(xh, th));
}
__device__ __forceinline__ f32x2 ld2(const float *p) { return *reinterpret_cast<const f32x2 *>(p); }      // 8-byte aligned
// sigmoid from the SFU approximations (ex2.approx, rcp.approx: ~2 ulp each, far inside the 1e-5 loss tolerance)
__device__ __forceinline__ float sigmoid_fast(float x) {
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * -1.4426950408889634f));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.f + e));
    return r;
}


// ------------------------------------------------------------------------------------------------
// Strided SIMT GEMM:  C[i, j] (op)= act( sum_k A(i, k) * B(k, j) + bias[j] )
//   A(i, k) = A[i * a_si + k * a_sk],  B(k, j) = B[k * b_sk + j * b_sj],  C[i * c_si + j]
// 64 x 64 x 16 tiles, 256 threads, 4 x 4 outputs per thread.  gridDim.z splits K (atomic accumulate).
// Covers X·W^T (forward), dY·W (input grad) and dY^T·X (weight grad, K = rows, split over z).
// (beats_conv, model.py:46-52, is evaluated over the occupied cells: mst_model_sparse.cuh.)
// ------------------------------------------------------------------------------------------------
struct GemmArgs {
    const float *A, *B, *bias;
    float *C;
    int M, N, K;
    long long a_si, a_sk, b_sk, b_sj, c_si;
    int act;          // 0 none, 1 leaky
    int accumulate;   // 0 store, 1 C += (plain), 2 atomicAdd (split-K or shared destination)
};

constexpr int kGemmTile = 64, kGemmK = 16, kGemmThreads = 256;

template <int AMODE>
__device__ __forceinline__ float gemm_load_a(const GemmArgs &g, int i, int k) {
    return g.A[(long long)i * g.a_si + (long long)k * g.a_sk];
}

template <int AMODE>
__global__ void __launch_bounds__(kGemmThreads)
k_gemm(GemmArgs g) {
    __shared__ float As[kGemmK][kGemmTile + 4];
    __shared__ float Bs[kGemmK][kGemmTile + 4];
    const int tid = threadIdx.x;
    const int i0 = blockIdx.y * kGemmTile, j0 = blockIdx.x * kGemmTile;
    const int kz = gridDim.z;
    const int k_per = ((g.K + kz - 1) / kz + kGemmK - 1) / kGemmK * kGemmK;
    const int k_lo = blockIdx.z * k_per, k_hi = min(g.K, k_lo + k_per);
    const int tx = tid & 15, ty = tid >> 4;
    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 4; b++) acc[a][b] = 0.f;
    // load mapping: pick the contiguous direction of each operand for coalescing
    const bool a_k_contig = g.a_sk == 1;
    const bool b_j_contig = (g.b_sj == 1);
    for (int k0 = k_lo; k0 < k_hi; k0 += kGemmK) {
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const int idx = tid + e * kGemmThreads;   // 0..1023 = 64 x 16
            int ii, kk;
            if (a_k_contig) { kk = idx & 15; ii = idx >> 4; } else { ii = idx & 63; kk = idx >> 6; }
            const int gi = i0 + ii, gk = k0 + kk;
            As[kk][ii] = (gi < g.M && gk < k_hi) ? gemm_load_a<AMODE>(g, gi, gk) : 0.f;
            int jj, kb;
            if (b_j_contig) { jj = idx & 63; kb = idx >> 6; } else { kb = idx & 15; jj = idx >> 4; }
            const int gj = j0 + jj, gkb = k0 + kb;
            Bs[kb][jj] = (gj < g.N && gkb < k_hi) ? g.B[(long long)gkb * g.b_sk + (long long)gj * g.b_sj] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < kGemmK; kk++) {
            float a[4], b[4];
#pragma unroll
            for (int x = 0; x < 4; x++) { a[x] = As[kk][ty * 4 + x]; b[x] = Bs[kk][tx * 4 + x]; }
#pragma unroll
            for (int x = 0; x < 4; x++)
#pragma unroll
                for (int y = 0; y < 4; y++) acc[x][y] = fmaf(a[x], b[y], acc[x][y]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int x = 0; x < 4; x++) {
        const int gi = i0 + ty * 4 + x;
        if (gi >= g.M) continue;
#pragma unroll
        for (int y = 0; y < 4; y++) {
            const int gj = j0 + tx * 4 + y;
            if (gj >= g.N) continue;
            float v = acc[x][y];
            if (g.bias && blockIdx.z == 0) v += g.bias[gj];
            if (g.act == 1) v = leaky(v);
            float *dst = g.C + (long long)gi * g.c_si + gj;
            if (g.accumulate == 0) *dst = v;
            else if (g.accumulate == 1) *dst += v;
            else atomicAdd(dst, v);
        }
    }
}

// Same contract for the problems whose 64 x 64 tiling would leave all but a handful of SMs idle (per-sample heads,
// per-channel instrument rows, their weight gradients): one warp per output element, lanes stride over k, shuffle
// reduction.  gridDim.y splits K like k_gemm's gridDim.z.
constexpr int kGemmWarpThreads = 256;
__global__ void __launch_bounds__(kGemmWarpThreads)
k_gemm_warp(GemmArgs g) {
    const int lane = threadIdx.x & 31;
    const long long out = (long long)blockIdx.x * (kGemmWarpThreads / 32) + (threadIdx.x >> 5);
    if (out >= (long long)g.M * g.N) return;
    const int i = (int)(out / g.N), j = (int)(out - (long long)i * g.N);
    const int k_per = (g.K + gridDim.y - 1) / gridDim.y;
    const int k_lo = blockIdx.y * k_per, k_hi = min(g.K, k_lo + k_per);
    const float *a = g.A + (long long)i * g.a_si, *b = g.B + (long long)j * g.b_sj;
    float acc0 = 0.f, acc1 = 0.f;
    int k = k_lo + lane;
    for (; k + 32 < k_hi; k += 64) {
        acc0 = fmaf(a[(long long)k * g.a_sk], b[(long long)k * g.b_sk], acc0);
        acc1 = fmaf(a[(long long)(k + 32) * g.a_sk], b[(long long)(k + 32) * g.b_sk], acc1);
    }
    if (k < k_hi) acc0 = fmaf(a[(long long)k * g.a_sk], b[(long long)k * g.b_sk], acc0);
    float v = acc0 + acc1;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if (lane != 0) return;
    if (g.bias && blockIdx.y == 0) v += g.bias[j];
    if (g.act == 1) v = leaky(v);
    float *dst = g.C + (long long)i * g.c_si + j;
    if (g.accumulate == 0) *dst = v;
    else if (g.accumulate == 1) *dst += v;
    else atomicAdd(dst, v);
}

// ------------------------------------------------------------------------------------------------
// Broadcast add + activation over the row space (n, c, b, t, f):  Y[row, :] = act(bias + sum_s Z_s[map_s(row), :])
// A source keeps a subset of the five coordinates (bit 4 = n, 3 = c, 2 = b, 1 = t, 0 = f); its rows are the
// row-major linearisation of the kept coordinates.  This is cat_with_broadcast + Linear, decomposed.
// ------------------------------------------------------------------------------------------------
struct Dims5 { int n, c, b, t, f; };
constexpr int kMaxSources = 6;
struct BcastArgs {
    const float *src[kMaxSources];
    int mask[kMaxSources];
    int n_src;
    const float *bias;   // may be null
    float *Y;
    Dims5 d;
    int cols;
    int act;             // 0 none, 1 leaky
};

__device__ __forceinline__ long long src_row(const Dims5 &d, int mask, int n, int c, int b, int t, int f) {
    long long r = 0;
    if (mask & 16) r = n;
    if (mask & 8) r = r * d.c + c;
    if (mask & 4) r = r * d.b + b;
    if (mask & 2) r = r * d.t + t;
    if (mask & 1) r = r * d.f + f;
    return r;
}

// One lane group (`lpr` lanes, a power of two) per (n, c, b, t) prefix: the source rows are decoded once per prefix, the
// sources that do not depend on f are summed once per column chunk and reused for the d.f rows of the prefix, so a row
// costs one load per f-dependent source and one store.  VEC = 4 moves 16 bytes per lane (cols % 4 == 0, 16-byte
// aligned sources and destination; the bias may sit at any float offset and is read with scalar loads).
template <int VEC>
__global__ void __launch_bounds__(256)
k_bcast_add(BcastArgs a, int lpr) {
    const long long prefixes = (long long)a.d.n * a.d.c * a.d.b * a.d.t;
    const int F = a.d.f, cv = a.cols / VEC;
    const int sub = threadIdx.x & (lpr - 1);
    const long long g0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) / lpr, n_groups = ((long long)gridDim.x * blockDim.x) / lpr;
    for (long long p = g0; p < prefixes; p += n_groups) {
        unsigned r = (unsigned)p;
        const int t = (int)(r % (unsigned)a.d.t); r /= (unsigned)a.d.t;
        const int b = (int)(r % (unsigned)a.d.b); r /= (unsigned)a.d.b;
        const int c = (int)(r % (unsigned)a.d.c); r /= (unsigned)a.d.c;
        const int n = (int)r;
        const float *sp[kMaxSources];
#pragma unroll
        for (int s = 0; s < kMaxSources; s++) {
            // row of the prefix in source s; an f-dependent source holds d.f consecutive rows from there
            const long long row = s < a.n_src ? src_row(a.d, a.mask[s] & ~1, n, c, b, t, 0) * ((a.mask[s] & 1) ? F : 1) : 0;
            sp[s] = s < a.n_src ? a.src[s] + row * a.cols : nullptr;
        }
        float *y = a.Y + p * F * a.cols;
        for (int j = sub; j < cv; j += lpr) {
            float base[VEC];
#pragma unroll
            for (int e = 0; e < VEC; e++) base[e] = a.bias ? a.bias[j * VEC + e] : 0.f;
#pragma unroll
            for (int s = 0; s < kMaxSources; s++)
                if (s < a.n_src && !(a.mask[s] & 1)) {
                    if (VEC == 4) { const float4 v = *reinterpret_cast<const float4 *>(sp[s] + j * 4); base[0] += v.x; base[1] += v.y; base[2] += v.z; base[3] += v.w; }
                    else base[0] += sp[s][j];
                }
#pragma unroll 2
            for (int f = 0; f < F; f++) {
                float v[VEC];
#pragma unroll
                for (int e = 0; e < VEC; e++) v[e] = base[e];
#pragma unroll
                for (int s = 0; s < kMaxSources; s++)
                    if (s < a.n_src && (a.mask[s] & 1)) {
                        const float *q = sp[s] + (long long)f * a.cols + j * VEC;
                        if (VEC == 4) { const float4 w = *reinterpret_cast<const float4 *>(q); v[0] += w.x; v[1] += w.y; v[2] += w.z; v[3] += w.w; }
                        else v[0] += q[0];
                    }
                if (a.act) {
#pragma unroll
                    for (int e = 0; e < VEC; e++) v[e] = leaky(v[e]);
                }
                float *dst = y + (long long)f * a.cols + j * VEC;
                if (VEC == 4) *reinterpret_cast<float4 *>(dst) = make_float4(v[0], v[1], v[2], v[3]);
                else dst[0] = v[0];
            }
        }
    }
}

// dPre = dY * act'(Y) elementwise (in place allowed)
__global__ void __launch_bounds__(256)
k_act_grad(const float *__restrict__ dY, const float *__restrict__ Y, float *__restrict__ dPre, long long n) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        dPre[i] = dY[i] * leaky_grad_from_out(Y[i]);
}

// out[a, i] (+)= sum_l in[a, l, i]   (sum over the middle axis; `inner` contiguous)
__global__ void __launch_bounds__(256)
k_reduce_mid(const float *__restrict__ in, float *__restrict__ out, long long A, int L, long long inner, int accumulate) {
    const long long total = A * inner;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const long long a = idx / inner, i = idx - a * inner;
        const float *p = in + a * L * inner + i;
        float s = 0.f;
        for (int l = 0; l < L; l++) s += p[(long long)l * inner];
        if (accumulate) out[idx] += s; else out[idx] = s;
    }
}

// out[a, i] += sum_l in[a, l, i] with the middle axis split over blockIdx.y (atomic accumulate into `out`)
__global__ void __launch_bounds__(256)
k_reduce_mid_split(const float *__restrict__ in, float *__restrict__ out, long long A, int L, long long inner, int l_per) {
    const long long total = A * inner;
    const int l0 = blockIdx.y * l_per, l1 = min(L, l0 + l_per);
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const long long a = idx / inner, i = idx - a * inner;
        const float *p = in + a * L * inner + i;
        float s = 0.f;
        for (int l = l0; l < l1; l++) s += p[(long long)l * inner];
        atomicAdd(out + idx, s);
    }
}

// beats_conv weight gradient: G[o][ch][tap] (57, 50, 14) += gT[perm(ch, tap)][o] with gT (700, 57) in im2col column order
__global__ void __launch_bounds__(256)
k_conv_wgrad_unpack(const float *__restrict__ gT, float *__restrict__ G);

// Adam (torch.optim.Adam defaults: no amsgrad, no weight decay) on the flat buffers + optional zero_grad.
//   m = b1 m + (1 - b1) g ; v = b2 v + (1 - b2) g^2 ; p -= lr / (1 - b1^t) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
__global__ void __launch_bounds__(256)
k_adam(float *__restrict__ p, float *__restrict__ g, float *__restrict__ m, float *__restrict__ v, long long n,
       float b1, float b2, float eps, float step_size, float inv_sqrt_bc2, float grad_scale, int zero_grad) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float gi = g[i] * grad_scale;
        const float mi = b1 * m[i] + (1.f - b1) * gi;
        const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
        m[i] = mi; v[i] = vi;
        p[i] -= step_size * (mi / (sqrtf(vi) * inv_sqrt_bc2 + eps));
        if (zero_grad) g[i] = 0.f;
    }
}

// column sums of a [rows, cols] matrix into out[cols] (atomic accumulate) — bias gradients
__global__ void __launch_bounds__(256)
k_colsum(const float *__restrict__ in, float *__restrict__ out, long long rows, int cols, float *__restrict__ out2 = nullptr) {
    // blockDim = 256 = 8 row-lanes x 32 col-lanes (cols handled in chunks of 32)
    __shared__ float part[8][33];
    const int cl = threadIdx.x & 31, rl = threadIdx.x >> 5;
    for (int c0 = blockIdx.y * 32; c0 < cols; c0 += gridDim.y * 32) {
        const int c = c0 + cl;
        float s = 0.f;
        if (c < cols)
            for (long long r = (long long)blockIdx.x * 8 + rl; r < rows; r += (long long)gridDim.x * 8) s += in[r * cols + c];
        part[rl][cl] = s;
        __syncthreads();
        if (rl == 0 && c < cols) {
            float t = 0.f;
#pragma unroll
            for (int k = 0; k < 8; k++) t += part[k][cl];
            atomicAdd(out + c, t);
            if (out2) atomicAdd(out2 + c, t);
        }
        __syncthreads();
    }
}

// dPre = dY * act'(Y) (in place allowed) and out[cols] += column sums of dPre, in ONE pass: the activation backward and
// the bias gradient of a Linear layer (same thread layout as k_colsum)
__global__ void __launch_bounds__(256)
k_act_grad_colsum(const float *__restrict__ dY, const float *__restrict__ Y, float *__restrict__ dPre, float *__restrict__ out,
                  long long rows, int cols) {
    __shared__ float part[8][33];
    const int cl = threadIdx.x & 31, rl = threadIdx.x >> 5;
    for (int c0 = blockIdx.y * 32; c0 < cols; c0 += gridDim.y * 32) {
        const int c = c0 + cl;
        float s = 0.f;
        if (c < cols)
            for (long long r = (long long)blockIdx.x * 8 + rl; r < rows; r += (long long)gridDim.x * 8) {
                const long long i = r * cols + c;
                const float v = dY[i] * leaky_grad_from_out(Y[i]);
                dPre[i] = v;
                s += v;
            }
        part[rl][cl] = s;
        __syncthreads();
        if (rl == 0 && c < cols) {
            float t = 0.f;
#pragma unroll
            for (int k = 0; k < 8; k++) t += part[k][cl];
            atomicAdd(out + c, t);
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256)
k_fill(float *p, long long n, float v) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) p[i] = v;
}

__global__ void __launch_bounds__(256)
k_axpy(float *__restrict__ y, const float *__restrict__ x, long long n) {   // y += x
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) y[i] += x[i];
}

// dst[r, perm(j)] = src[r, j] style column permutations of weights (and the inverse for their gradients)
//   mode 0: beats_conv  (57, 50, 14) [o][f*5+feat][tap]  <->  [o][f*70 + tap*5 + feat]
//   mode 1: PCE.linear  (64, 514)    [j][o*8+oct | 456+i]  <->  [j][oct*57+o | 456+i]
//   mode 2: UCE.linear  (64, 940)    [j][f*94 + feat*47 + note]  <->  [j][f*94 + note*2 + feat]
__device__ __forceinline__ int perm_col(int mode, int j) {
    if (mode == 0) { const int ch = j / 14, tap = j - ch * 14; const int f = ch / 5, feat = ch - f * 5; return f * 70 + tap * 5 + feat; }
    if (mode == 1) { if (j >= 456) return j; const int o = j >> 3, oct = j & 7; return oct * 57 + o; }
    const int f = j / 94, r = j - f * 94; const int feat = r / 47, note = r - feat * 47; return f * 94 + note * 2 + feat;
}
__global__ void __launch_bounds__(256)
k_conv_wgrad_unpack(const float *__restrict__ gT, float *__restrict__ G) {
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < 57 * 700; idx += gridDim.x * blockDim.x) {
        const int o = idx / 700, j = idx - o * 700;
        G[idx] += gT[perm_col(0, j) * 57 + o];
    }
}

__global__ void __launch_bounds__(256)
k_permute_cols(const float *__restrict__ src, float *__restrict__ dst, int rows, int cols, int mode, int inverse, int accumulate) {
    const int total = rows * cols;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
        const int r = idx / cols, j = idx - r * cols;
        const int pj = perm_col(mode, j);
        if (!inverse) dst[r * cols + pj] = src[idx];
        else if (accumulate) dst[idx] += src[r * cols + pj];
        else dst[idx] = src[r * cols + pj];
    }
}

}  // namespace mstm
