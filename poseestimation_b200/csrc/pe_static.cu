// pe_frobenius_fwd_static: the Frobenius loss sum with NO allocation anywhere - for callers that cannot call cudaMalloc
// (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  It has no counterpart in the reference.
// Nothing else in the library uses this file and no existing code
// path was changed for it.
//
// The existing pe_frobenius_fwd meets its partial sums in a launch slot of the pool that pe_init cudaMalloc's.  This
// variant keeps its scratch in a file-scope `static __device__` array instead: a ticket word and one float64 partial per
// tile of 256 records, kPeStaticMaxTiles tiles = 256 KB + 256 B per GPU that loads the library, held until the process
// exits.  That size is the limit: n <= PE_STATIC_MAX_RECORDS (2^23), larger batches are rejected with PE_ERR_BAD_ARG.
//
// In HOST code the array's name is a host-side stand-in, not a device address.  The address the kernel gets comes from
// cudaGetSymbolAddress, asked for once per device with that device current (scratch_address below).
//
// Every call shares the one array, which keeps the previous call's contents: the ticket word is cleared on the caller's
// stream in front of each launch (and re-armed by the kernel), the partials are overwritten before they are read.  Two
// calls that OVERLAP on the GPU (different streams, or two host threads) race on it - serialise them.
//
// Side effect of the per-tile partials: the last CTA adds them in a fixed order, so *sum is bitwise reproducible from
// call to call (the slot-based kernels add per-CTA partials in completion order).
#include <mutex>

#include "pe_common.h"
#include "pe_math.cuh"
#include "pe_stage.cuh"

namespace pe {

constexpr int kPeStaticTile = 256;                 // records per tile = threads per CTA
constexpr int kPeStaticMaxTiles = 32768;           // x 256 records = PE_STATIC_MAX_RECORDS
constexpr int kPeStaticHeader = 256;               // ticket word, padded so that the partials stay 256-byte aligned
constexpr size_t kPeStaticBytes = kPeStaticHeader + sizeof(double) * (size_t)kPeStaticMaxTiles;
static_assert((long long)kPeStaticMaxTiles * kPeStaticTile == PE_STATIC_MAX_RECORDS, "header and kernel disagree on the limit");

static __device__ __align__(256) unsigned char g_pe_static_scratch[kPeStaticBytes];

// deterministic CTA sum (fixed shuffle tree, then the warps in order); result valid in thread 0
__device__ __forceinline__ double cta_sum_ordered(double v, double* s_warp) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();  // s_warp may still be read from the previous use
    if (lane == 0) s_warp[warp] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x == 0)
        for (int w = 0; w < kPeStaticTile / 32; ++w) t += s_warp[w];
    return t;
}

__global__ void __launch_bounds__(kPeStaticTile) frobenius_static_kernel(const float* __restrict__ a,
                                                                        const float* __restrict__ b, long long n,
                                                                        unsigned char* scratch, double* __restrict__ sum,
                                                                        int vec) {
    __shared__ __align__(16) float s_d[kPeStaticTile * 9];
    __shared__ double s_warp[kPeStaticTile / 32];
    __shared__ bool s_last;
    unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
    double* partial = reinterpret_cast<double*>(scratch + kPeStaticHeader);
    const long long ntiles = (n + kPeStaticTile - 1) / kPeStaticTile;
    const int tid = threadIdx.x;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long base = tile * kPeStaticTile;
        const int cnt = (int)((n - base) < kPeStaticTile ? (n - base) : kPeStaticTile);
        const float* pa = a + base * 9;
        const float* pb = b + base * 9;
        __syncthreads();  // the previous tile's differences have been consumed
        // coalesced: consecutive threads read consecutive words; the 36-byte record stride only exists in shared memory.
        // Full tiles of 16-byte aligned arrays (a tile is 9216 bytes) move as 128-bit streaming loads.
        if (vec && cnt == kPeStaticTile) {
            const float4* pa4 = reinterpret_cast<const float4*>(pa);
            const float4* pb4 = reinterpret_cast<const float4*>(pb);
            float4* s_d4 = reinterpret_cast<float4*>(s_d);
#pragma unroll
            for (int w = tid; w < kPeStaticTile * 9 / 4; w += kPeStaticTile) {
                const float4 x = __ldcs(pa4 + w), y = __ldcs(pb4 + w);
                s_d4[w] = make_float4(x.x - y.x, x.y - y.y, x.z - y.z, x.w - y.w);
            }
        } else {
            for (int w = tid; w < cnt * 9; w += kPeStaticTile) s_d[w] = __ldcs(pa + w) - __ldcs(pb + w);
        }
        __syncthreads();
        float f = 0.f;
        if (tid < cnt) {  // stride 9 words: conflict-free.  Same arithmetic as frob_dist (pe_math.cuh)
            float s = 0.f;
#pragma unroll
            for (int i = 0; i < 9; ++i) { const float d = s_d[9 * tid + i]; s = fmaf(d, d, s); }
            f = sqrtf(s);
        }
        const double t = cta_sum_ordered((double)f, s_warp);
        if (tid == 0) partial[tile] = t;
    }
    // ticket: the last CTA to arrive adds the tile partials in tile order
    if (tid == 0) {
        __threadfence();
        s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    double v = 0.0;
    for (long long t = tid; t < ntiles; t += kPeStaticTile) v += __ldcg(partial + t);  // L2: where the other CTAs wrote
    const double total = cta_sum_ordered(v, s_warp);
    if (tid == 0) {
        *sum = total;
        *ticket = 0u;  // re-armed for the next call
    }
}

// device address of the scratch array on the CURRENT device, resolved once per device
int scratch_address(unsigned char** out, int* sms_out) {
    static std::mutex mu;
    static unsigned char* addr[64] = {nullptr};
    static int sms[64] = {0};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail_cuda(e, "cudaGetDevice");
    if (dev < 0 || dev >= 64) return fail(PE_ERR_BAD_ARG, "device index out of range");
    std::lock_guard<std::mutex> lk(mu);
    if (!addr[dev]) {
        int major = 0, n_sm = 0;
        e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
        if (e == cudaSuccess) e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return fail_cuda(e, "cudaDeviceGetAttribute");
        if (major < 10) return fail(PE_ERR_NO_DEVICE, "this library is built for sm_100a (B200) only");
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_pe_static_scratch);  // `dev` is current: the symbol's instance on THIS device
        if (e != cudaSuccess) return fail_cuda(e, "cudaGetSymbolAddress(static scratch)");
        if (!p) return fail(PE_ERR_INTERNAL, "cudaGetSymbolAddress returned a null address");
        sms[dev] = n_sm;
        addr[dev] = static_cast<unsigned char*>(p);
    }
    *out = addr[dev];
    *sms_out = sms[dev];
    return 0;
}

}  // namespace pe

using namespace pe;

extern "C" int pe_frobenius_fwd_static(const float* a, const float* b, double* sum, int64_t n, void* stream) {
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (n < 0) return fail(PE_ERR_BAD_ARG, "negative record count");
    if (n > (int64_t)PE_STATIC_MAX_RECORDS)
        return fail(PE_ERR_BAD_ARG, "pe_frobenius_fwd_static: n exceeds PE_STATIC_MAX_RECORDS (8388608 matrices: the static "
                                    "scratch holds one partial per 256 records); use pe_frobenius_fwd");
    if (!sum) return fail(PE_ERR_BAD_ARG, "null accumulator pointer");
    if (reinterpret_cast<uintptr_t>(sum) & 7u) return fail(PE_ERR_BAD_ARG, "float64 accumulator not aligned to 8 bytes");
    if ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 3u)
        return fail(PE_ERR_BAD_ARG, "array not aligned to 4 bytes");
    if (n == 0) {  // an empty batch still defines its sum
        cudaError_t e = cudaMemsetAsync(sum, 0, sizeof(double), st);
        if (e != cudaSuccess) return fail_cuda(e, "cudaMemsetAsync(sum)");
        return 0;
    }
    if (!a || !b) return fail(PE_ERR_BAD_ARG, "null input pointer");
    unsigned char* scratch = nullptr;
    int sms = 0;
    if (int rc = scratch_address(&scratch, &sms)) return rc;
    // the array keeps the last call's contents: the ticket must start at zero (also after a call that was aborted)
    cudaError_t e = cudaMemsetAsync(scratch, 0, sizeof(unsigned int), st);
    if (e != cudaSuccess) return fail_cuda(e, "cudaMemsetAsync(ticket)");
    const long long ntiles = (n + kPeStaticTile - 1) / kPeStaticTile;
    long long grid = (long long)sms * 8;  // 8 x 256 threads per SM (9.3 KB of shared memory each)
    if (grid > ntiles) grid = ntiles;
    const int vec = ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15u) == 0;
    frobenius_static_kernel<<<(unsigned)grid, kPeStaticTile, 0, st>>>(a, b, (long long)n, scratch, sum, vec);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, "frobenius_static_kernel");
    count_launch();
    return 0;
}
