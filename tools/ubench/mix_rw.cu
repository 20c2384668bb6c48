// Micro-benchmark: achievable HBM throughput on B200 as a function of the READ : WRITE mix, with plain 128-bit
// streaming loads / stores (no shared memory, no arithmetic).  The roofline denominator of this repository is the
// driver's 1 : 1 copy figure (MEASURED_PEAKS.json); the forward maps of the path write 36 bytes per record and read
// only 12-24, and this shows what that mix can reach at all.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/ubench/mix_rw.cu -o mix_rw ; run: ./mix_rw
#include <cstdio>
#include <cuda_runtime.h>

// every thread block streams `R` float4 reads and `W` float4 writes per item
template <int R, int W>
__global__ void __launch_bounds__(256) mix(const float4* __restrict__ in, float4* __restrict__ out, long long items) {
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < items; i += step) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const float4 v = __ldcs(in + (long long)r * items + i);
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
#pragma unroll
        for (int w = 0; w < W; ++w) __stcs(out + (long long)w * items + i, acc);
        if (W == 0 && acc.x == 123.456f) out[0] = acc;  // keep the loads alive
    }
}

template <int R, int W>
void run(const char* label, const float4* in, float4* out, long long items, int sms) {
    const int grid = sms * 8;
    for (int i = 0; i < 3; ++i) mix<R, W><<<grid, 256>>>(in, out, items);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    const int reps = 20;
    for (int i = 0; i < reps; ++i) mix<R, W><<<grid, 256>>>(in, out, items);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, a, b);
    ms /= reps;
    const double bytes = 16.0 * (R + W) * (double)items;
    printf("%-34s read %2d : write %2d   %8.4f ms   %8.1f GB/s\n", label, R, W, ms, bytes / ms / 1e6);
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const long long items = 1ll << 24;  // x 16 B x (R + W): 0.27 - 3.2 GB per launch, far beyond the 126 MB L2
    float4 *in, *out;
    cudaMalloc(&in, 16 * items * 9);
    cudaMalloc(&out, 16 * items * 9);
    cudaMemset(in, 0, 16 * items * 9);
    cudaMemset(out, 0, 16 * items * 9);
    run<1, 1>("copy", in, out, items, sms);
    run<9, 9>("copy (9 streams each way)", in, out, items, sms);
    run<1, 0>("read only", in, out, items, sms);
    run<9, 0>("read only (9 streams)", in, out, items, sms);
    run<0, 1>("write only", in, out, items, sms);
    run<0, 9>("write only (9 streams)", in, out, items, sms);
    run<1, 3>("euler / expmap fwd  (12 : 36 B)", in, out, items, sms);
    run<4, 9>("quat fwd           (16 : 36 B)", in, out, items, sms);
    run<2, 3>("ortho6d fwd        (24 : 36 B)", in, out, items, sms);
    run<2, 1>("frobenius bwd      (72 : 36 B)", in, out, items, sms);
    run<9, 3>("svd step           (72 : 36 B, 9:4.5)", in, out, items, sms);
    run<3, 1>("geodesic bwd-like  (3 : 1)", in, out, items, sms);
    cudaFree(in); cudaFree(out);
    return 0;
}
