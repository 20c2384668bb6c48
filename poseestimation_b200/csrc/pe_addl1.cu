// ADD-L1 mesh losses of the iterative refiner:
//   compute_ADD_L1_loss              /root/reference/Iterative/loss.py:10-28
//   compute_disentangled_ADD_L1_loss /root/reference/Iterative/loss.py:31-48   (rot + transl + depth terms)
//   transform_pts                    /root/reference/Iterative/loss.py:51-69
//
// The reference transforms the (B, N, 3) vertex array once per pose (6 transforms for the disentangled loss: 6 x 24 B
// per vertex plus the temporaries of abs/mean).  Algebraically
//   T_gt p - T_pred p = (R_gt - R_pred) p + (t_gt - t_pred),
// and in the disentangled loss only the rotation term depends on the vertices at all (the other two poses differ
// from T_gt in the translation only: their per-vertex distance is the constant |t_gt - t'|).  So ONE pass over the
// vertices (12 B per vertex) gives the loss and, from the same pass, the sums that make up its gradient w.r.t.
// T_pred:  d/dR_ij = -sum_k sgn(d_ki) p_kj / (3N),  d/dt_i = -sum_k sgn(d_ki) / (3N).
//
// Layout: one WARP per item (pose pair + its vertices), lane 0 streams the vertices through a warp-private TMA ring,
// lanes read vertices at a 3-word stride (conflict-free), 13 running sums per lane, butterfly reduction.
#include "pe_common.h"
#include "pe_stage.cuh"

namespace pe {

constexpr int kAWarps = 32;   // 12 B per vertex: twice the vertex rate of the Kabsch kernel for the same GB/s; the kernel
constexpr int kAStages = 2;   // is latency-bound on its 13 accumulator chains, so it runs the full 32 warps per SM
constexpr int kAChunk = 256;  // vertices per stage (3 KB)
constexpr int kAStageWords = kAChunk * 3;
constexpr int kASmem = 1024 + kAWarps * kAStages * kAStageWords * 4;

struct AddL1Params {
    const float* t_gt;    // (b, 16)
    const float* t_pred;  // (b, 16)
    const float* pts;     // (b, npts, 3)
    float* item;          // (b) per-item loss, nullable
    double* sum;          // sum of the per-item losses, nullable
    float* grad;          // (b, 12): d item / d T_pred[:3, :4], nullable
    long long b;
    int npts;
    int disentangled;
    int use_tma;
};

__device__ __forceinline__ float wsum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float sgn(float v) { return v > 0.f ? 1.f : (v < 0.f ? -1.f : 0.f); }

template <bool TMA>
__global__ void __launch_bounds__(kAWarps * 32, 1) add_l1_kernel(const AddL1Params prm) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw) + warp * kAStages;
    float* ring = reinterpret_cast<float*>(smem_raw + 1024) + warp * kAStages * kAStageWords;

    const long long nwarps = (long long)gridDim.x * kAWarps;
    const long long gw = (long long)blockIdx.x * kAWarps + warp;
    const int npts = prm.npts;
    const int cpi = (npts + kAChunk - 1) / kAChunk;  // chunks per item
    const long long my_items = gw < prm.b ? (prm.b - gw + nwarps - 1) / nwarps : 0;
    const long long steps = my_items * cpi;
    uint64_t policy = 0;

    auto issue = [&](int s, long long j) {  // lane 0: chunk (j % cpi) of this warp's item number (j / cpi)
        long long it = gw + (j / cpi) * nwarps;
        int ch = (int)(j % cpi);
        int cnt = min(kAChunk, npts - ch * kAChunk);
        uint32_t bytes = (uint32_t)cnt * 12u;
        mbar_arrive_expect_tx(&bars[s], bytes);
        bulk_g2s(ring + s * kAStageWords, prm.pts + (it * npts + (long long)ch * kAChunk) * 3, bytes, &bars[s], policy);
    };
    if constexpr (TMA) {
        if (lane == 0) {
            policy = l2_evict_first_policy();
            for (int s = 0; s < kAStages; ++s) mbar_init(&bars[s], 1);
            mbar_init_fence();
            for (int s = 0; s < kAStages; ++s)
                if (s < steps) issue(s, s);
        }
        __syncwarp();
    }

    double warp_total = 0.0;  // lane 0 only
    long long j = 0;          // running chunk counter of this warp (ring position)
    for (long long it = gw; it < prm.b; it += nwarps) {
        // R_gt - R_pred and t_gt - t_pred (the latter is dropped from the vertex term of the disentangled loss)
        float d[9], dt[3], tdelta[3];
        {
            const float* a = prm.t_gt + it * 16;
            const float* c = prm.t_pred + it * 16;
#pragma unroll
            for (int i = 0; i < 3; ++i) {
#pragma unroll
                for (int k = 0; k < 3; ++k) d[3 * i + k] = __ldg(a + 4 * i + k) - __ldg(c + 4 * i + k);
                tdelta[i] = __ldg(a + 4 * i + 3) - __ldg(c + 4 * i + 3);
                dt[i] = prm.disentangled ? 0.f : tdelta[i];
            }
        }
        float acc = 0.f, S[9], sv[3];
#pragma unroll
        for (int i = 0; i < 9; ++i) S[i] = 0.f;
        sv[0] = sv[1] = sv[2] = 0.f;

        for (int ch = 0; ch < cpi; ++ch, ++j) {
            const int s = (int)(j % kAStages);
            const int cnt = min(kAChunk, npts - ch * kAChunk);
            const float* sp;
            if constexpr (TMA) {
                mbar_wait(&bars[s], (uint32_t)((j / kAStages) & 1));
                sp = ring + s * kAStageWords;
            } else {
                sp = prm.pts + (it * npts + (long long)ch * kAChunk) * 3;
            }
#pragma unroll 4
            for (int k = lane; k < cnt; k += 32) {
                float x = sp[3 * k], y = sp[3 * k + 1], z = sp[3 * k + 2];
                float e0 = fmaf(d[2], z, fmaf(d[1], y, fmaf(d[0], x, dt[0])));
                float e1 = fmaf(d[5], z, fmaf(d[4], y, fmaf(d[3], x, dt[1])));
                float e2 = fmaf(d[8], z, fmaf(d[7], y, fmaf(d[6], x, dt[2])));
                acc += fabsf(e0) + fabsf(e1) + fabsf(e2);
                float g0 = sgn(e0), g1 = sgn(e1), g2 = sgn(e2);
                S[0] = fmaf(g0, x, S[0]); S[1] = fmaf(g0, y, S[1]); S[2] = fmaf(g0, z, S[2]);
                S[3] = fmaf(g1, x, S[3]); S[4] = fmaf(g1, y, S[4]); S[5] = fmaf(g1, z, S[5]);
                S[6] = fmaf(g2, x, S[6]); S[7] = fmaf(g2, y, S[7]); S[8] = fmaf(g2, z, S[8]);
                sv[0] += g0; sv[1] += g1; sv[2] += g2;
            }
            if constexpr (TMA) {
                __syncwarp();
                if (lane == 0 && j + kAStages < steps) issue(s, j + kAStages);
            }
        }
        __syncwarp();
        acc = wsum(acc);
#pragma unroll
        for (int i = 0; i < 9; ++i) S[i] = wsum(S[i]);
        sv[0] = wsum(sv[0]); sv[1] = wsum(sv[1]); sv[2] = wsum(sv[2]);
        if (lane == 0) {
            const float k3n = 1.0f / (3.0f * (float)npts);
            float loss = acc * k3n;
            float gt3[3] = {-sv[0] * k3n, -sv[1] * k3n, -sv[2] * k3n};
            if (prm.disentangled) {
                // transl pose: (x, y) of T_pred, depth pose: z of T_pred - constant per-vertex distances
                loss += (fabsf(tdelta[0]) + fabsf(tdelta[1]) + fabsf(tdelta[2])) * (1.0f / 3.0f);
                gt3[0] = -sgn(tdelta[0]) * (1.0f / 3.0f);
                gt3[1] = -sgn(tdelta[1]) * (1.0f / 3.0f);
                gt3[2] = -sgn(tdelta[2]) * (1.0f / 3.0f);
            }
            if (prm.item) prm.item[it] = loss;
            warp_total += (double)loss;
            if (prm.grad) {
                float* gd = prm.grad + it * 12;
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    gd[4 * i] = -S[3 * i] * k3n; gd[4 * i + 1] = -S[3 * i + 1] * k3n; gd[4 * i + 2] = -S[3 * i + 2] * k3n;
                    gd[4 * i + 3] = gt3[i];
                }
            }
        }
    }
    if (prm.sum && lane == 0 && my_items > 0) atomicAdd(prm.sum, warp_total);
}

}  // namespace pe

using namespace pe;

extern "C" int pe_add_l1(const float* t_gt, const float* t_pred, const float* pts, int64_t b, int64_t npts,
                         int disentangled, float* item, double* sum, float* grad, void* stream) {
    if (b < 0 || npts < 1 || npts > (1 << 24)) return fail(PE_ERR_BAD_ARG, "bad item count or vertex count");
    if (reinterpret_cast<uintptr_t>(sum) & 7u) return fail(PE_ERR_BAD_ARG, "float64 accumulator not aligned to 8 bytes");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (sum) {
        cudaError_t e0 = cudaMemsetAsync(sum, 0, sizeof(double), st);
        if (e0 != cudaSuccess) return fail_cuda(e0, "cudaMemsetAsync(sum)");
    }
    if (b == 0) return 0;
    if (!t_gt || !t_pred || !pts) return fail(PE_ERR_BAD_ARG, "null pointer");
    if (!item && !sum && !grad) return fail(PE_ERR_BAD_ARG, "all outputs are null");
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail_cuda(e, "cudaGetDevice");
    static int sms[64] = {0};
    if (dev < 0 || dev >= 64) return fail(PE_ERR_BAD_ARG, "device index out of range");
    if (sms[dev] == 0) {
        cudaDeviceProp prop;
        e = cudaGetDeviceProperties(&prop, dev);
        if (e != cudaSuccess) return fail_cuda(e, "cudaGetDeviceProperties");
        if (prop.major < 10) return fail(PE_ERR_NO_DEVICE, "this library is built for sm_100a (B200) only");
        e = cudaFuncSetAttribute(add_l1_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kASmem);
        if (e != cudaSuccess) return fail_cuda(e, "cudaFuncSetAttribute(smem)");
        sms[dev] = prop.multiProcessorCount;
    }
    AddL1Params prm{t_gt, t_pred, pts, item, sum, grad, (long long)b, (int)npts, disentangled, 0};
    prm.use_tma = (npts % 4 == 0) && (reinterpret_cast<uintptr_t>(pts) & 15u) == 0 && !force_direct();
    long long grid = (b + kAWarps - 1) / kAWarps;
    if (grid > sms[dev]) grid = sms[dev];
    if (prm.use_tma) add_l1_kernel<true><<<(unsigned)grid, kAWarps * 32, kASmem, st>>>(prm);
    else add_l1_kernel<false><<<(unsigned)grid, kAWarps * 32, 0, st>>>(prm);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, "add_l1_kernel");
    count_launch();
    return 0;
}
