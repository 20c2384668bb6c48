// Micro-benchmark: issue cost of packed fp32x2 FMA-pipe instructions on sm_100a as a function of how many DISTINCT
// 64-bit source operands they read.  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a fma2_rate.cu -o fma2_rate
// One warp per SMSP (128 threads per CTA, 1 CTA per SM) or 8 warps per SMSP; prints cycles per warp-instruction per SMSP.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
#define FMA2(d, a, b, c) asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c))
#define MUL2(d, a, b) asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b))
#define MUFU(d, a) asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(d) : "f"(a))
#define FMA1(d, a, b, c) asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c))
constexpr int NACC = 8, ITERS = 4096;

template <int MODE>
__global__ void k(u64* out, long long* cyc, u64 seed) {
    u64 acc[NACC], x[NACC], y[NACC];
    float fa[NACC], fx[NACC], fy[NACC];
    for (int i = 0; i < NACC; ++i) {
        acc[i] = seed + threadIdx.x * 7 + i; x[i] = seed * 3 + i * 11 + threadIdx.x; y[i] = seed * 5 + i * 13;
        fa[i] = (float)(seed + i); fx[i] = 1.0f + 1e-7f * i; fy[i] = 1e-9f * threadIdx.x;
    }
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i) {
            if (MODE == 0) FMA2(acc[i], x[i], y[i], acc[i]);            // 3 distinct source pairs
            if (MODE == 1) FMA2(acc[i], acc[i], y[i], acc[i]);          // 2 distinct
            if (MODE == 2) FMA2(acc[i], acc[i], acc[i], acc[i]);        // 1 distinct
            if (MODE == 3) MUL2(acc[i], acc[i], y[i]);                  // FMUL2, 2 distinct
            if (MODE == 4) FMA1(fa[i], fx[i], fy[i], fa[i]);            // scalar FFMA, 3 distinct
            if (MODE == 5) FMA2(acc[i], x[i], y[(i + 1) % NACC], acc[i]);  // 3 distinct, other pairing
            if (MODE == 6) FMA2(acc[i], x[0], y[i], acc[i]);            // one shared multiplicand (like cos/sin)
            // do the packed and the scalar instructions share one pipe?  MODE 7: one FFMA2 + one scalar FFMA per slot
            // (2 instructions), MODE 8: one FFMA2 + two scalar FFMA (3 instructions); cycles are reported per SLOT
            if (MODE == 7 || MODE == 8) { FMA2(acc[i], x[i], y[i], acc[i]); FMA1(fa[i], fx[i], fy[i], fa[i]); }
            if (MODE == 8) FMA1(fx[i], fa[i], fy[i], fx[i]);
            if (MODE == 9) { FMA2(acc[i], x[i], y[i], acc[i]); MUFU(fa[i], fa[i]); }   // packed FMA + MUFU.RSQ per slot
        }
    }
    long long t1 = clock64();
    u64 s = 0; float fs = 0;
    for (int i = 0; i < NACC; ++i) { s += acc[i]; fs += fa[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + (u64)fs;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, int threads) {
    u64* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 8); cudaMalloc(&cyc, 148 * 8);
    k<MODE><<<148, threads>>>(out, cyc, 12345);
    k<MODE><<<148, threads>>>(out, cyc, 12345);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
    int warps_per_smsp = threads / 32 / 4;
    printf("%-44s threads/CTA %4d: %.3f cycles per warp-instruction per SMSP\n", name, threads,
           avg / ((double)ITERS * NACC * warps_per_smsp));
    cudaFree(out); cudaFree(cyc);
}
int main() {
    for (int threads : {128, 1024}) {
        run<0>("FFMA2 3 distinct src pairs", threads);
        run<5>("FFMA2 3 distinct src pairs (shifted)", threads);
        run<6>("FFMA2 shared multiplicand pair", threads);
        run<1>("FFMA2 2 distinct", threads);
        run<2>("FFMA2 1 distinct", threads);
        run<3>("FMUL2 2 distinct", threads);
        run<4>("FFMA scalar 3 distinct", threads);
        run<7>("slot = FFMA2 (3 distinct) + 1 scalar FFMA", threads);
        run<8>("slot = FFMA2 (3 distinct) + 2 scalar FFMA", threads);
        run<9>("slot = FFMA2 (3 distinct) + 1 MUFU.RSQ", threads);
    }
    return 0;
}
