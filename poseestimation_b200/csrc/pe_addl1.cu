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
// Layout: one WARP per item (pose pair + its vertices), lane 0 streams the vertices through a warp-private TMA ring.
// The loop is issue-bound before it is HBM-bound (13 accumulator chains per vertex), so each lane takes TWO vertices
// per step (v and v + 32 of a 64-vertex group, 3-word lane stride: conflict-free) and runs the whole update in packed
// fp32x2 arithmetic: 9 FFMA2 for the residuals, |e| = sgn(e) e as one FFMA2 per row, 9 FFMA2 + 3 FADD2 for the
// gradient sums.  The 13 sums are reduced across the warp with a transposing butterfly (16 + 8 + 4 + 2 + 1 shuffles
// instead of 13 x 5) that leaves sum j in lane 2j, and those lanes write the outputs.
#include <mutex>

#include "pe_common.h"
#include "pe_lanes.cuh"
#include "pe_stage.cuh"

namespace pe {

#ifndef PE_AWARPS
#define PE_AWARPS 16
#endif
#ifndef PE_ASTAGES
#define PE_ASTAGES 3
#endif
constexpr int kAWarps = PE_AWARPS;
constexpr int kAStages = PE_ASTAGES;
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
    int lg_split;         // each item is shared by 2^lg_split warps (small batches); outputs are then added atomically
};

// sign with sgn(0) = 0, as torch.sign in the backward of abs(): LOP3 + FSETP + FSEL
__device__ __forceinline__ float sgn(float v) {
    const float s = __int_as_float((__float_as_int(v) & 0x80000000) | 0x3f800000);
    return v != 0.f ? s : 0.f;
}

struct AddL1Acc {
    F2 s[9], sv[3], acc;
};

// two vertices (x, y, z hold one coordinate of each); MASKED: lanes of a ragged tail switch a vertex off through va / vb
template <bool MASKED>
__device__ __forceinline__ void add_l1_pair(const F2* d, const F2* dt, F2 x, F2 y, F2 z, bool va, bool vb, AddL1Acc& a) {
    F2 e0 = t_fma(d[2], z, t_fma(d[1], y, t_fma(d[0], x, dt[0])));
    F2 e1 = t_fma(d[5], z, t_fma(d[4], y, t_fma(d[3], x, dt[1])));
    F2 e2 = t_fma(d[8], z, t_fma(d[7], y, t_fma(d[6], x, dt[2])));
    F2 g0{sgn(e0.x), sgn(e0.y)}, g1{sgn(e1.x), sgn(e1.y)}, g2{sgn(e2.x), sgn(e2.y)};
    if constexpr (MASKED) {
        if (!va) { g0.x = 0.f; g1.x = 0.f; g2.x = 0.f; }
        if (!vb) { g0.y = 0.f; g1.y = 0.f; g2.y = 0.f; }
    }
    a.acc = t_fma(g0, e0, a.acc); a.acc = t_fma(g1, e1, a.acc); a.acc = t_fma(g2, e2, a.acc);  // |e| = sgn(e) e
    a.s[0] = t_fma(g0, x, a.s[0]); a.s[1] = t_fma(g0, y, a.s[1]); a.s[2] = t_fma(g0, z, a.s[2]);
    a.s[3] = t_fma(g1, x, a.s[3]); a.s[4] = t_fma(g1, y, a.s[4]); a.s[5] = t_fma(g1, z, a.s[5]);
    a.s[6] = t_fma(g2, x, a.s[6]); a.s[7] = t_fma(g2, y, a.s[7]); a.s[8] = t_fma(g2, z, a.s[8]);
    a.sv[0] = t_add(a.sv[0], g0); a.sv[1] = t_add(a.sv[1], g1); a.sv[2] = t_add(a.sv[2], g2);
}

// v[0..15] per lane -> the warp total of v[j] in lanes 2j and 2j + 1 (returned): 31 shuffles for 16 sums
__device__ __forceinline__ float transpose_reduce16(float* v, int lane) {
#pragma unroll
    for (int o = 16; o >= 2; o >>= 1) {
        const int half = o >> 1;  // values kept after this step
        const bool up = (lane & o) != 0;
#pragma unroll
        for (int i = 0; i < half; ++i) {
            const float send = up ? v[i] : v[i + half];
            const float keep = up ? v[i + half] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
    return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 1);
}

template <bool TMA>
__global__ void __launch_bounds__(kAWarps * 32, 1) add_l1_kernel(const AddL1Params prm) {
    pdl_wait();               // programmatic dependent launch (pe_common.h)
    pdl_launch_dependents();
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw) + warp * kAStages;
    float* ring = reinterpret_cast<float*>(smem_raw + 1024) + warp * kAStages * kAStageWords;

    const long long nwarps = (long long)gridDim.x * kAWarps;
    const long long gw = (long long)blockIdx.x * kAWarps + warp;
    const int npts = prm.npts;
    const int cpi = (npts + kAChunk - 1) / kAChunk;  // chunks per item
    // Work units: (item, part).  A small batch (fewer items than warps on the device) is split 2^lg ways, part p of an
    // item being its chunks [p cpi / 2^lg, (p + 1) cpi / 2^lg).  The number of warps is a multiple of 2^lg, so a warp
    // keeps the same part for all of its units: unit u = gw + k nwarps -> item u >> lg, part gw & (2^lg - 1).
    const int lg = prm.lg_split;
    const int part = (int)(gw & ((1 << lg) - 1));
    const int ch0 = (part * cpi) >> lg, ch1 = ((part + 1) * cpi) >> lg;
    const long long units = prm.b << lg;
    const long long my_units = gw < units ? (units - gw + nwarps - 1) / nwarps : 0;
    const long long steps = my_units * (ch1 - ch0);
    const float k3n = 1.0f / (3.0f * (float)npts);
    uint64_t policy = 0;

    // lane 0: the next chunk to fetch is chunk iss_ch of item iss_it (kept incrementally: a 64-bit division per
    // chunk costs more instructions than the chunk's arithmetic)
    const long long item_step = nwarps >> lg;
    long long iss_it = gw >> lg;
    int iss_ch = ch0;
    auto issue = [&](int s) {
        int cnt = min(kAChunk, npts - iss_ch * kAChunk);
        uint32_t bytes = (uint32_t)cnt * 12u;
        mbar_arrive_expect_tx(&bars[s], bytes);
        bulk_g2s(ring + s * kAStageWords, prm.pts + (iss_it * npts + (long long)iss_ch * kAChunk) * 3, bytes, &bars[s],
                 policy);
        if (++iss_ch == ch1) { iss_ch = ch0; iss_it += item_step; }
    };
    if constexpr (TMA) {
        if (lane == 0) {
            policy = l2_evict_first_policy();
            for (int s = 0; s < kAStages; ++s) mbar_init(&bars[s], 1);
            mbar_init_fence();
            for (int s = 0; s < kAStages; ++s)
                if (s < steps) issue(s);
        }
        __syncwarp();
    }

    // lanes 0..11 fetch one entry each of (T_gt - T_pred)[:3, :4] of an item: 0..8 the rotation block row-major, 9..11
    // the translation column.  The fetch for the NEXT item is issued at the top of the current one, so its latency
    // hides behind a whole item of vertex work instead of stalling the warp while its ring sits full.
    const int poff = lane < 9 ? 4 * (lane / 3) + lane % 3 : 4 * (lane - 9) + 3;
    // (the two loaded words are only subtracted when the item starts: a subtraction at fetch time would wait for them)
    auto fetch_pose = [&](long long item, float& a, float& c) {
        a = c = 0.f;
        if (lane < 12 && item < prm.b) {
            a = __ldg(prm.t_gt + item * 16 + poff);
            c = __ldg(prm.t_pred + item * 16 + poff);
        }
    };
    float cur_a, cur_c;
    fetch_pose(gw >> lg, cur_a, cur_c);

    double warp_total = 0.0;  // lane 24 only (the lane that ends up with the loss)
    long long j = 0;          // running chunk counter of this warp
    int s = 0;                // ring position of chunk j (j % kAStages)
    uint32_t phase = 0;       // its mbarrier parity ((j / kAStages) & 1)
    for (long long it = gw >> lg; it < prm.b && ch0 < ch1; it += item_step) {
        const float cur = cur_a - cur_c;
        fetch_pose(it + item_step, cur_a, cur_c);  // in flight for the whole item
        // R_gt - R_pred and t_gt - t_pred (the latter is dropped from the vertex term of the disentangled loss)
        F2 d[9], dt[3];
        float tdelta[3];
#pragma unroll
        for (int i = 0; i < 9; ++i) {
            const float v = __shfl_sync(0xffffffffu, cur, i);
            d[i] = F2{v, v};
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            tdelta[i] = __shfl_sync(0xffffffffu, cur, 9 + i);
            const float v = prm.disentangled ? 0.f : tdelta[i];
            dt[i] = F2{v, v};
        }
        AddL1Acc acc;
#pragma unroll
        for (int i = 0; i < 9; ++i) acc.s[i] = F2{0.f, 0.f};
        acc.sv[0] = acc.sv[1] = acc.sv[2] = acc.acc = F2{0.f, 0.f};

        for (int ch = ch0; ch < ch1; ++ch, ++j) {
            const int cnt = min(kAChunk, npts - ch * kAChunk);
            const float* sp;
            if constexpr (TMA) {
                mbar_wait(&bars[s], phase);
                sp = ring + s * kAStageWords;
            } else {
                sp = prm.pts + (it * npts + (long long)ch * kAChunk) * 3;
            }
            int k = 0;
#pragma unroll 2
            for (; k + 64 <= cnt; k += 64) {
                const float* pa = sp + 3 * (k + lane);
                const float* pb = pa + 96;
                add_l1_pair<false>(d, dt, F2{pa[0], pb[0]}, F2{pa[1], pb[1]}, F2{pa[2], pb[2]}, true, true, acc);
            }
            if (k < cnt) {  // ragged tail of the item: fewer than 64 vertices left
                const bool va = k + lane < cnt, vb = k + 32 + lane < cnt;
                const float* pa = sp + 3 * (k + lane);
                const float* pb = pa + 96;
                F2 x{va ? pa[0] : 0.f, vb ? pb[0] : 0.f}, y{va ? pa[1] : 0.f, vb ? pb[1] : 0.f},
                    z{va ? pa[2] : 0.f, vb ? pb[2] : 0.f};
                add_l1_pair<true>(d, dt, x, y, z, va, vb, acc);
            }
            if constexpr (TMA) {
                __syncwarp();
                if (lane == 0 && j + kAStages < steps) issue(s);
            }
            if (++s == kAStages) { s = 0; phase ^= 1u; }
        }
        __syncwarp();
        float v[16];
#pragma unroll
        for (int i = 0; i < 9; ++i) v[i] = acc.s[i].x + acc.s[i].y;
#pragma unroll
        for (int i = 0; i < 3; ++i) v[9 + i] = acc.sv[i].x + acc.sv[i].y;
        v[12] = acc.acc.x + acc.acc.y;
        v[13] = v[14] = v[15] = 0.f;
        const float tot = transpose_reduce16(v, lane);
        const int jv = lane >> 1;  // which sum this lane holds
        if ((lane & 1) == 0 && jv < 13) {
            // a split item: every part adds its share (the outputs were zeroed by the host side); the terms that do not
            // depend on the vertices come from part 0 only
            const bool first = part == 0;
            if (jv == 12) {
                float loss = tot * k3n;
                if (prm.disentangled && first)  // transl pose: (x, y) of T_pred, depth pose: z of T_pred - constants
                    loss += (fabsf(tdelta[0]) + fabsf(tdelta[1]) + fabsf(tdelta[2])) * (1.0f / 3.0f);
                if (prm.item) {
                    if (lg == 0) prm.item[it] = loss;
                    else atomicAdd(prm.item + it, loss);
                }
                warp_total += (double)loss;
            } else if (prm.grad) {
                float* gd = prm.grad + it * 12;
                float val;
                int off;
                if (jv < 9) {
                    off = 4 * (jv / 3) + jv % 3;
                    val = -tot * k3n;
                } else {
                    const int i = jv - 9;
                    const float td = i == 0 ? tdelta[0] : (i == 1 ? tdelta[1] : tdelta[2]);
                    off = 4 * i + 3;
                    val = prm.disentangled ? (first ? -sgn(td) * (1.0f / 3.0f) : 0.f) : -tot * k3n;
                }
                if (lg == 0) gd[off] = val;
                else atomicAdd(gd + off, val);
            }
        }
    }
    if (prm.sum && lane == 24 && my_units > 0) atomicAdd(prm.sum, warp_total);
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
    static std::mutex mu;  // the per-device attribute setup runs once, also with several host threads
    if (dev < 0 || dev >= 64) return fail(PE_ERR_BAD_ARG, "device index out of range");
    std::unique_lock<std::mutex> lk(mu);
    if (sms[dev] == 0) {
        cudaDeviceProp prop;
        e = cudaGetDeviceProperties(&prop, dev);
        if (e != cudaSuccess) return fail_cuda(e, "cudaGetDeviceProperties");
        if (prop.major < 10) return fail(PE_ERR_NO_DEVICE, "this library is built for sm_100a (B200) only");
        e = cudaFuncSetAttribute(add_l1_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kASmem);
        if (e != cudaSuccess) return fail_cuda(e, "cudaFuncSetAttribute(smem)");
        sms[dev] = prop.multiProcessorCount;
    }
    const int nsm = sms[dev];
    lk.unlock();
    AddL1Params prm{t_gt, t_pred, pts, item, sum, grad, (long long)b, (int)npts, disentangled, 0, 0};
    prm.use_tma = (npts % 4 == 0) && (reinterpret_cast<uintptr_t>(pts) & 15u) == 0 && !force_direct();
    // small batches (the reference trains with 64 meshes per step): share an item among up to 16 warps
    const long long all_warps = (long long)nsm * kAWarps;
    const long long cpi = (npts + kAChunk - 1) / kAChunk;
    // (each part keeps at least four chunks: below that the two extra memsets and the atomics cost more than the split saves)
    while (prm.lg_split < 4 && (b << (prm.lg_split + 1)) <= all_warps && (cpi >> (prm.lg_split + 1)) >= 4) ++prm.lg_split;
    if (prm.lg_split > 0) {
        cudaError_t e1 = cudaSuccess;
        if (item) e1 = cudaMemsetAsync(item, 0, sizeof(float) * (size_t)b, st);
        if (e1 == cudaSuccess && grad) e1 = cudaMemsetAsync(grad, 0, sizeof(float) * 12 * (size_t)b, st);
        if (e1 != cudaSuccess) return fail_cuda(e1, "cudaMemsetAsync(split outputs)");
    }
    long long grid = ((b << prm.lg_split) + kAWarps - 1) / kAWarps;
    if (grid > nsm) grid = nsm;
    if (prm.use_tma) launch_pdl(add_l1_kernel<true>, (unsigned)grid, (unsigned)(kAWarps * 32), (size_t)kASmem, st, prm);
    else launch_pdl(add_l1_kernel<false>, (unsigned)grid, (unsigned)(kAWarps * 32), (size_t)0, st, prm);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, "add_l1_kernel");
    count_launch();
    return 0;
}
