// Kabsch alignment of cloud pairs.  The reference has no such function: it regresses the
// rotation with PointNet++ (/root/reference/point_cloud/main.py:158-203); this op solves the same (pc1, pc_out)
// pair of lines 171-181 in closed form and ends in the reference's own projection,
// symmetric_orthogonalization (rotation_representation.py:192-206), via pe::svd_orth.
//
// Layout: one WARP per cloud pair.  Lane 0 streams the pair's two (npts, 3) arrays through a warp-private ring of
// shared-memory stages with TMA bulk copies (kChunk points = 3 KB per array per stage, kStages deep); the 32 lanes
// read points at a 3-word stride (conflict-free) and each keeps 15 running sums (sum p', sum q', sum q' p'^T, all
// relative to the pair's first point so that far-from-origin clouds do not cancel).  A butterfly shuffle reduces
// them; the 3x3 covariance of pair k is parked in lane k % 32, and every 32 pairs the whole warp runs 32
// register-resident SVDs at once instead of one lane running them one by one.
//
// Weighted variant (what a correspondence-weighting network feeds): per-point weights w_k >= 0,
//   pbar = sum w p / W,  qbar = sum w q / W,  C = sum w (q - qbar)(p - pbar)^T,  W = sum w,
// a third 1 KB stream per stage (28 B per point pair instead of 24); W == 0 gives C = 0 (svd_orth's rotation of the
// zero matrix) rather than NaN.  The context record then carries W as a 16th word.
#include <mutex>

#include "pe_common.h"
#include "pe_math.cuh"
#include "pe_stage.cuh"

namespace pe {

constexpr int kKWarps = 8;     // warps per CTA, one CTA per SM
constexpr int kKStages = 3;    // ring depth per warp
constexpr int kChunk = 256;    // points per stage per array
template <bool WT>
struct KGeo {
    static constexpr int kStageWords = 2 * kChunk * 3 + (WT ? kChunk : 0);
    static constexpr int kSmem = 1024 + kKWarps * kKStages * kStageWords * 4;
    static constexpr int kCtx = WT ? 16 : 15;
};

struct KabschParams {
    const float* p;
    const float* q;
    const float* w;  // (b, npts) weights, weighted instantiation only
    float* r;
    float* t;  // nullable
    float* ctx;  // nullable: per pair H (9), pbar (3), qbar (3) - what the backward needs besides R
    int raw;     // 1: no alignment; r = sum_k q_k p_k^T (uncentred), t = sum_k q_k  (reduction of an outer product)
    long long b;
    int npts;
    int use_tma;
};

__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <bool WT>
__global__ void __launch_bounds__(kKWarps * 32, 1) kabsch_kernel(const KabschParams prm) {
    pdl_wait();               // programmatic dependent launch (pe_common.h)
    pdl_launch_dependents();
    constexpr int kKStageWords = KGeo<WT>::kStageWords;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw) + warp * kKStages;
    float* ring = reinterpret_cast<float*>(smem_raw + 1024) + warp * kKStages * kKStageWords;

    const long long nwarps = (long long)gridDim.x * kKWarps;
    const long long gw = (long long)blockIdx.x * kKWarps + warp;
    const int npts = prm.npts;
    const int cpp = (npts + kChunk - 1) / kChunk;  // chunks per pair
    const long long my_pairs = gw < prm.b ? (prm.b - gw + nwarps - 1) / nwarps : 0;
    const long long items = my_pairs * cpp;
    const bool tma = prm.use_tma != 0;
    uint64_t policy = 0;

    if (lane == 0) {
        policy = l2_evict_first_policy();
        for (int s = 0; s < kKStages; ++s) mbar_init(&bars[s], 1);
        mbar_init_fence();
    }
    __syncwarp();

    // lane 0: the next chunk to fetch is chunk iss_ch of pair iss_pair (kept incrementally: no 64-bit divisions)
    long long iss_pair = gw;
    int iss_ch = 0;
    auto issue = [&](int s) {
        const long long pair = iss_pair;
        const int ch = iss_ch;
        if (++iss_ch == cpp) { iss_ch = 0; iss_pair += nwarps; }
        int cnt = min(kChunk, npts - ch * kChunk);
        uint32_t bytes = (uint32_t)cnt * 12u;
        long long off = (pair * npts + (long long)ch * kChunk) * 3;
        float* dst = ring + s * kKStageWords;
        mbar_arrive_expect_tx(&bars[s], 2 * bytes + (WT ? (uint32_t)cnt * 4u : 0u));
        bulk_g2s(dst, prm.p + off, bytes, &bars[s], policy);
        bulk_g2s(dst + kChunk * 3, prm.q + off, bytes, &bars[s], policy);
        if constexpr (WT)
            bulk_g2s(dst + 2 * kChunk * 3, prm.w + pair * npts + (long long)ch * kChunk, (uint32_t)cnt * 4u, &bars[s],
                     policy);
    };
    if (tma && lane == 0) {
        for (int s = 0; s < kKStages; ++s)
            if (s < items) issue(s);
    }

    float sp0 = 0, sp1 = 0, sp2 = 0, sq0 = 0, sq1 = 0, sq2 = 0;
    float c00 = 0, c01 = 0, c02 = 0, c10 = 0, c11 = 0, c12 = 0, c20 = 0, c21 = 0, c22 = 0;
    float p00 = 0, p01 = 0, p02 = 0, q00 = 0, q01 = 0, q02 = 0;  // the pair's first point (the shift)
    float sw = 0.f, kw = 0.f;  // running / parked sum of weights
    float keep[9], kp[3], kq[3];
    long long kpair = -1;
#pragma unroll
    for (int i = 0; i < 9; ++i) keep[i] = 0.f;
    kp[0] = kp[1] = kp[2] = kq[0] = kq[1] = kq[2] = 0.f;
    int parked = 0;

    auto flush = [&]() {
        if (!WT && lane < parked && prm.raw) {
            // sum q p^T = sum (q - qbar)(p - pbar)^T + N qbar pbar^T ;  sum q = N qbar
            const float nn = (float)npts;
            float* dst = prm.r + kpair * 9;
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j < 3; ++j) dst[3 * i + j] = fmaf(nn * kq[i], kp[j], keep[3 * i + j]);
            if (prm.t) {
                float* td = prm.t + kpair * 3;
                td[0] = nn * kq[0]; td[1] = nn * kq[1]; td[2] = nn * kq[2];
            }
        } else if (lane < parked) {
            float r[9], ds;
            svd_orth_fast(keep, r, ds);
            if (prm.ctx) {
                float* cd = prm.ctx + kpair * KGeo<WT>::kCtx;
#pragma unroll
                for (int i = 0; i < 9; ++i) cd[i] = keep[i];
                cd[9] = kp[0]; cd[10] = kp[1]; cd[11] = kp[2];
                cd[12] = kq[0]; cd[13] = kq[1]; cd[14] = kq[2];
                if constexpr (WT) cd[15] = kw;
            }
            float* dst = prm.r + kpair * 9;
#pragma unroll
            for (int i = 0; i < 9; ++i) dst[i] = r[i];
            if (prm.t) {
                float* td = prm.t + kpair * 3;
                td[0] = kq[0] - (r[0] * kp[0] + r[1] * kp[1] + r[2] * kp[2]);
                td[1] = kq[1] - (r[3] * kp[0] + r[4] * kp[1] + r[5] * kp[2]);
                td[2] = kq[2] - (r[6] * kp[0] + r[7] * kp[1] + r[8] * kp[2]);
            }
        }
        parked = 0;
    };

    int s = 0, ch = 0;        // ring position and chunk-in-pair of item j, kept incrementally
    uint32_t parity = 0;
    long long pair = gw;
    for (long long j = 0; j < items; ++j) {
        const int cnt = min(kChunk, npts - ch * kChunk);
        const float* sP;
        const float* sQ;
        const float* sW = nullptr;
        if (tma) {
            mbar_wait(&bars[s], parity);
            sP = ring + s * kKStageWords;
            sQ = sP + kChunk * 3;
            if constexpr (WT) sW = sP + 2 * kChunk * 3;
        } else {  // unaligned clouds (npts % 4 != 0 or odd base): read HBM directly, same arithmetic
            long long off = (pair * npts + (long long)ch * kChunk) * 3;
            sP = prm.p + off;
            sQ = prm.q + off;
            if constexpr (WT) sW = prm.w + pair * npts + (long long)ch * kChunk;
        }
        if (ch == 0) {
            int k0 = 0;
            if constexpr (WT) {  // shift by a point that counts: the first of the leading 32 with a non-zero weight
                const unsigned live = __ballot_sync(0xffffffffu, lane < cnt && sW[lane] > 0.f);
                k0 = live ? __ffs(live) - 1 : 0;
            }
            p00 = sP[3 * k0]; p01 = sP[3 * k0 + 1]; p02 = sP[3 * k0 + 2];
            q00 = sQ[3 * k0]; q01 = sQ[3 * k0 + 1]; q02 = sQ[3 * k0 + 2];
        }
#pragma unroll 4
        for (int i = lane; i < cnt; i += 32) {
            float a0 = sP[3 * i] - p00, a1 = sP[3 * i + 1] - p01, a2 = sP[3 * i + 2] - p02;
            float b0 = sQ[3 * i] - q00, b1 = sQ[3 * i + 1] - q01, b2 = sQ[3 * i + 2] - q02;
            if constexpr (WT) {
                const float wk = sW[i];
                sw += wk;
                a0 *= wk; a1 *= wk; a2 *= wk;  // a = w (p - p0) from here on
                sq0 = fmaf(wk, b0, sq0); sq1 = fmaf(wk, b1, sq1); sq2 = fmaf(wk, b2, sq2);
            } else {
                sq0 += b0; sq1 += b1; sq2 += b2;
            }
            sp0 += a0; sp1 += a1; sp2 += a2;
            c00 = fmaf(b0, a0, c00); c01 = fmaf(b0, a1, c01); c02 = fmaf(b0, a2, c02);
            c10 = fmaf(b1, a0, c10); c11 = fmaf(b1, a1, c11); c12 = fmaf(b1, a2, c12);
            c20 = fmaf(b2, a0, c20); c21 = fmaf(b2, a1, c21); c22 = fmaf(b2, a2, c22);
        }
        __syncwarp();  // every lane is done with stage s
        if (tma && lane == 0 && j + kKStages < items) issue(s);
        const bool last_chunk = ch == cpp - 1;

        if (last_chunk) {
            sp0 = warp_sum_f(sp0); sp1 = warp_sum_f(sp1); sp2 = warp_sum_f(sp2);
            sq0 = warp_sum_f(sq0); sq1 = warp_sum_f(sq1); sq2 = warp_sum_f(sq2);
            c00 = warp_sum_f(c00); c01 = warp_sum_f(c01); c02 = warp_sum_f(c02);
            c10 = warp_sum_f(c10); c11 = warp_sum_f(c11); c12 = warp_sum_f(c12);
            c20 = warp_sum_f(c20); c21 = warp_sum_f(c21); c22 = warp_sum_f(c22);
            float inv = 1.0f / (float)npts;
            if constexpr (WT) {
                sw = warp_sum_f(sw);
                inv = sw > 0.f ? 1.0f / sw : 0.f;
            }
            if (lane == parked) {
                kw = sw;
                // C = sum q' p'^T - (sum q')(sum p')^T / N   (= sum (q - qbar)(p - pbar)^T)
                keep[0] = fmaf(-sq0 * inv, sp0, c00); keep[1] = fmaf(-sq0 * inv, sp1, c01); keep[2] = fmaf(-sq0 * inv, sp2, c02);
                keep[3] = fmaf(-sq1 * inv, sp0, c10); keep[4] = fmaf(-sq1 * inv, sp1, c11); keep[5] = fmaf(-sq1 * inv, sp2, c12);
                keep[6] = fmaf(-sq2 * inv, sp0, c20); keep[7] = fmaf(-sq2 * inv, sp1, c21); keep[8] = fmaf(-sq2 * inv, sp2, c22);
                kp[0] = fmaf(sp0, inv, p00); kp[1] = fmaf(sp1, inv, p01); kp[2] = fmaf(sp2, inv, p02);
                kq[0] = fmaf(sq0, inv, q00); kq[1] = fmaf(sq1, inv, q01); kq[2] = fmaf(sq2, inv, q02);
                kpair = pair;
            }
            ++parked;
            sp0 = sp1 = sp2 = sq0 = sq1 = sq2 = sw = 0.f;
            c00 = c01 = c02 = c10 = c11 = c12 = c20 = c21 = c22 = 0.f;
            if (parked == 32) flush();
        }
        if (++s == kKStages) { s = 0; parity ^= 1u; }
        if (++ch == cpp) { ch = 0; pair += nwarps; }
    }
    if (parked > 0) flush();
}

}  // namespace pe

using namespace pe;

namespace {
int kabsch_launch(KabschParams prm, void* stream) {
    if (prm.b < 0 || prm.npts < 1 || prm.npts > (1 << 24)) return fail(PE_ERR_BAD_ARG, "bad cloud count or cloud size");
    if (prm.b == 0) return 0;
    if (!prm.p || !prm.q || !prm.r) return fail(PE_ERR_BAD_ARG, "null pointer");
    const bool weighted = prm.w != nullptr;
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
        e = cudaFuncSetAttribute(kabsch_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, KGeo<false>::kSmem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(kabsch_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, KGeo<true>::kSmem);
        if (e != cudaSuccess) return fail_cuda(e, "cudaFuncSetAttribute(smem)");
        sms[dev] = prop.multiProcessorCount;
    }
    const int nsm = sms[dev];
    lk.unlock();
    prm.use_tma = (prm.npts % 4 == 0) &&
                  ((reinterpret_cast<uintptr_t>(prm.p) | reinterpret_cast<uintptr_t>(prm.q) |
                    reinterpret_cast<uintptr_t>(prm.w)) & 15u) == 0 && !force_direct();
    long long grid = (prm.b + kKWarps - 1) / kKWarps;
    if (grid > nsm) grid = nsm;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (weighted) launch_pdl(kabsch_kernel<true>, (unsigned)grid, (unsigned)(kKWarps * 32), (size_t)KGeo<true>::kSmem, st, prm);
    else launch_pdl(kabsch_kernel<false>, (unsigned)grid, (unsigned)(kKWarps * 32), (size_t)KGeo<false>::kSmem, st, prm);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, "kabsch_kernel");
    count_launch();
    return 0;
}
}  // namespace

extern "C" int pe_kabsch_fwd(const float* p, const float* q, int64_t b, int64_t npts, float* r, float* t,
                             void* stream) {
    return kabsch_launch(KabschParams{p, q, nullptr, r, t, nullptr, 0, (long long)b, (int)npts, 0}, stream);
}
extern "C" int pe_kabsch_fwd_ctx(const float* p, const float* q, int64_t b, int64_t npts, float* r, float* t,
                                 float* ctx, void* stream) {
    if (!ctx) return fail(PE_ERR_BAD_ARG, "null context pointer");
    return kabsch_launch(KabschParams{p, q, nullptr, r, t, ctx, 0, (long long)b, (int)npts, 0}, stream);
}
extern "C" int pe_kabsch_weighted_fwd(const float* p, const float* q, const float* w, int64_t b, int64_t npts, float* r,
                                      float* t, float* ctx, void* stream) {
    if (!w && b > 0) return fail(PE_ERR_BAD_ARG, "null weight pointer");
    return kabsch_launch(KabschParams{p, q, w, r, t, ctx, 0, (long long)b, (int)npts, 0}, stream);
}
extern "C" int pe_cloud_outer_sum(const float* g, const float* p, int64_t b, int64_t npts, float* m, float* gsum,
                                  void* stream) {
    return kabsch_launch(KabschParams{p, g, nullptr, m, gsum, nullptr, 1, (long long)b, (int)npts, 0}, stream);
}
