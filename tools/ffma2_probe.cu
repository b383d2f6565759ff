// Issue-rate probe for packed fp32 (FFMA2) against scalar FFMA on sm_100a (development probe).
// Each warp runs ILP independent accumulator chains; the multiplier comes from a register, a uniform register
// (kernel parameter) or the constant bank.  Prints warp-instructions per clock per SM.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__constant__ float c_w[64];

template <int MODE>
__global__ void __launch_bounds__(512) k(float *out, int iters, float w_param) {
    float acc[16];
    u64 acc2[8];
    float x = threadIdx.x * 1e-3f, w = out[threadIdx.x & 7];
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = i * 0.5f;
#pragma unroll
    for (int i = 0; i < 8; i++) acc2[i] = ((u64)__float_as_uint(i * 0.5f) << 32) | __float_as_uint(i * 0.25f);
    const u64 xx = ((u64)__float_as_uint(x) << 32) | __float_as_uint(x);
    u64 ww = ((u64)__float_as_uint(w) << 32) | __float_as_uint(w + 1.f);
    for (int it = 0; it < iters; it++) {
        if (MODE == 0) {            // scalar FFMA, 3 registers
#pragma unroll
            for (int i = 0; i < 16; i++) acc[i] = fmaf(acc[i], w, x);
        } else if (MODE == 1) {     // scalar FFMA, constant-bank multiplier
#pragma unroll
            for (int i = 0; i < 16; i++) acc[i] = fmaf(x, c_w[i], acc[i]);
        } else if (MODE == 2) {     // FFMA2, registers
#pragma unroll
            for (int i = 0; i < 8; i++) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc2[i]) : "l"(xx), "l"(ww));
        } else if (MODE == 3) {     // FFMA2 + an ALU-pipe instruction in between (FMNMX)
#pragma unroll
            for (int i = 0; i < 8; i++) {
                asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc2[i]) : "l"(xx), "l"(ww));
                acc[i] = fmaxf(acc[i], x + i);
                acc[i + 8] = fminf(acc[i + 8], w - i);
            }
        } else if (MODE == 4) {     // scalar FFMA const + 2 FMNMX per 2 FFMA (same ratio as MODE 3)
#pragma unroll
            for (int i = 0; i < 8; i++) {
                acc[i] = fmaf(x, c_w[i], acc[i]);
                acc[i + 8] = fmaf(x, c_w[i + 8], acc[i + 8]);
                x = fmaxf(x, acc[i]);
                w = fminf(w, acc[i + 8]);
            }
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) s += acc[i];
#pragma unroll
    for (int i = 0; i < 8; i++) s += __uint_as_float((uint32_t)acc2[i]) + __uint_as_float((uint32_t)(acc2[i] >> 32));
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + x + w;
}

template <int MODE> void run(const char *name, int fma_per_iter, int inst_per_iter) {
    float *d; cudaMalloc(&d, 148 * 512 * 4 * 4); cudaMemset(d, 0, 148 * 512 * 4 * 4);
    const int iters = 20000;
    k<MODE><<<148, 512>>>(d, 100, 1.f);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0); k<MODE><<<148, 512>>>(d, iters, 1.f); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double cycles = ms * 1e-3 * clk * 1e3;
    const double warps = 16;
    printf("%-44s %8.3f ms  fma lanes/clk/SM %7.1f  warp-inst/clk/SM %5.2f  (%s)\n", name, ms,
           warps * 32 * fma_per_iter * iters / cycles, warps * inst_per_iter * iters / cycles, cudaGetErrorString(cudaGetLastError()));
    cudaFree(d);
}
int main() {
    run<0>("FFMA reg,reg,reg", 16, 16);
    run<1>("FFMA reg,const,reg", 16, 16);
    run<2>("FFMA2 reg", 16, 8);
    run<3>("FFMA2 + 2 FMNMX", 16, 24);
    run<4>("2 FFMA const + 2 FMNMX", 16, 32);
    return 0;
}
