// Per-cloud affine maps of xyz points:  out[b][k] = A_b p[b][k] + t_b   (A_b 3x3 row-major, optionally transposed).
//
// This one streaming kernel serves
//   * the cloud-pair construction of /root/reference/point_cloud/main.py:173-181 (pc2 = gt_rmat @ pc1), where the
//     reference expands gt_rmat to (B*N, 3, 3) - a 36 KB/pair temporary - to rotate 12 KB of points;
//   * transform_pts of /root/reference/Iterative/loss.py:51-69 (T[:3,:3] p + T[:3,3]);
//   * both halves of the Kabsch backward (dQ = G_H (p - pbar) + ..., dP = G_H^T (q - qbar) + ...), and the
//     point-gradient of the two maps above (A^T g).
// 24 algorithmic bytes per point (12 read, 12 written).  Four consecutive points are 48 bytes = three 128-bit
// words, so a thread moves its points with three LDG.128 and three STG.128, fully coalesced, no shared memory; the
// 12 coefficients of the cloud are broadcast loads (all lanes of a warp normally sit in the same cloud).
#include "pe_common.h"

namespace pe {

struct AffineParams {
    const float* a;   // (b, 9)  or (b, a_stride) rows with the 3x3 block at the front
    const float* t;   // (b, 3) with stride t_stride, nullable
    const float* p;   // (b, npts, 3)
    float* out;       // (b, npts, 3)
    long long b;
    long long npts;
    long long a_stride, t_stride;  // floats between consecutive clouds
    int a_row, t_elem;             // floats between rows of A (3, or 4 for the top-left block of a 4x4 transform) and
                                   // between components of t (1, or 4 for the last column of a 4x4 transform)
    int transpose;
};

__device__ __forceinline__ void load_coeffs(const AffineParams& prm, long long cloud, float* m, float* t) {
    const float* a = prm.a + cloud * prm.a_stride;
    if (prm.transpose) {
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) m[3 * i + j] = __ldg(a + prm.a_row * j + i);
    } else {
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) m[3 * i + j] = __ldg(a + prm.a_row * i + j);
    }
    if (prm.t) {
        const float* tt = prm.t + cloud * prm.t_stride;
        t[0] = __ldg(tt); t[1] = __ldg(tt + prm.t_elem); t[2] = __ldg(tt + 2 * prm.t_elem);
    } else {
        t[0] = t[1] = t[2] = 0.f;
    }
}

__device__ __forceinline__ void apply(const float* m, const float* t, float x, float y, float z, float& ox, float& oy,
                                      float& oz) {
    ox = fmaf(m[2], z, fmaf(m[1], y, fmaf(m[0], x, t[0])));
    oy = fmaf(m[5], z, fmaf(m[4], y, fmaf(m[3], x, t[1])));
    oz = fmaf(m[8], z, fmaf(m[7], y, fmaf(m[6], x, t[2])));
}

// Grid-stride walk over (cloud, position in cloud) without a 64-bit division per step (each is ~100 instructions):
// one division when the thread starts, then cloud += step / per, pos += step % per with a carry.
struct CloudWalk {
    long long cloud, pos, dq, dr, per;
    __device__ __forceinline__ CloudWalk(long long first, long long step, long long per_) : per(per_) {
        cloud = first / per; pos = first - cloud * per;
        dq = step / per; dr = step - dq * per;
    }
    __device__ __forceinline__ void next() {
        cloud += dq; pos += dr;
        if (pos >= per) { pos -= per; ++cloud; }
    }
};

// npts % 4 == 0 and 16-byte aligned arrays: one thread = 4 points of one cloud
__global__ void __launch_bounds__(256) affine_points_vec4_kernel(const AffineParams prm) {
    pdl_wait();               // programmatic dependent launch (pe_common.h)
    pdl_launch_dependents();
    const long long groups = prm.b * prm.npts / 4;
    const long long gpc = prm.npts / 4;  // groups per cloud
    const long long step = (long long)gridDim.x * blockDim.x;
    const float4* src = reinterpret_cast<const float4*>(prm.p);
    float4* dst = reinterpret_cast<float4*>(prm.out);
    const long long first = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    CloudWalk walk(first, step, gpc);
    for (long long g = first; g < groups; g += step, walk.next()) {
        float m[9], t[3];
        load_coeffs(prm, walk.cloud, m, t);
        float4 v0 = __ldcs(src + 3 * g), v1 = __ldcs(src + 3 * g + 1), v2 = __ldcs(src + 3 * g + 2);
        float4 o0, o1, o2;
        apply(m, t, v0.x, v0.y, v0.z, o0.x, o0.y, o0.z);
        apply(m, t, v0.w, v1.x, v1.y, o0.w, o1.x, o1.y);
        apply(m, t, v1.z, v1.w, v2.x, o1.z, o1.w, o2.x);
        apply(m, t, v2.y, v2.z, v2.w, o2.y, o2.z, o2.w);
        __stcs(dst + 3 * g, o0); __stcs(dst + 3 * g + 1, o1); __stcs(dst + 3 * g + 2, o2);
    }
}

// any npts / alignment: one thread = one point
__global__ void __launch_bounds__(256) affine_points_scalar_kernel(const AffineParams prm) {
    pdl_wait();               // programmatic dependent launch (pe_common.h)
    pdl_launch_dependents();
    const long long total = prm.b * prm.npts;
    const long long step = (long long)gridDim.x * blockDim.x;
    const long long first = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    CloudWalk walk(first, step, prm.npts);
    for (long long k = first; k < total; k += step, walk.next()) {
        float m[9], t[3];
        load_coeffs(prm, walk.cloud, m, t);
        float ox, oy, oz;
        apply(m, t, prm.p[3 * k], prm.p[3 * k + 1], prm.p[3 * k + 2], ox, oy, oz);
        prm.out[3 * k] = ox; prm.out[3 * k + 1] = oy; prm.out[3 * k + 2] = oz;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Point part of the WEIGHTED Kabsch backward.  With (A_q | b_q), (A_p | b_p) from
// pe_kabsch_weighted_bwd_coeffs (A_p = A_q^T = G_C^T, the 1/N of the plain case replaced by 1/W):
//   u_k = A_q p_k + b_q,   v_k = A_p q_k + b_p,
//   dq_k = w_k u_k,   dp_k = w_k v_k,
//   dw_k = (q_k - qbar) . u_k - e . (p_k - pbar),   e = R^T g_t / W = -(b_p + A_p qbar)
// (dw: the weight enters C directly, and pbar, qbar - hence t = qbar - R pbar - through (p_k - pbar)/W,
// (q_k - qbar)/W; the centroid terms of dC vanish because sum_k w_k (p_k - pbar) = 0.)
// One pass: reads p, q, w (28 B per point), writes whichever of dp, dq, dw are asked for (up to 28 B).
struct WBwdParams {
    const float* p;
    const float* q;
    const float* w;
    const float* cq;   // (b, 12)
    const float* cp;   // (b, 12)
    const float* ctx;  // (b, 16): pbar at 9..11, qbar at 12..14
    float* dp;         // nullable
    float* dq;         // nullable
    float* dw;         // nullable
    long long b, npts;
};

struct WBwdCoeffs {
    float a[9], bq[3], bp[3], pb[3], qb[3], e[3];
};

__device__ __forceinline__ void wbwd_load(const WBwdParams& prm, long long cloud, WBwdCoeffs& c) {
    const float* cq = prm.cq + cloud * 12;
    const float* cp = prm.cp + cloud * 12;
    const float* cx = prm.ctx + cloud * 16;
#pragma unroll
    for (int i = 0; i < 9; ++i) c.a[i] = __ldg(cq + i);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        c.bq[i] = __ldg(cq + 9 + i);
        c.bp[i] = __ldg(cp + 9 + i);
        c.pb[i] = __ldg(cx + 9 + i);
        c.qb[i] = __ldg(cx + 12 + i);
    }
#pragma unroll
    for (int i = 0; i < 3; ++i)
        c.e[i] = -(c.bp[i] + c.a[i] * c.qb[0] + c.a[3 + i] * c.qb[1] + c.a[6 + i] * c.qb[2]);
}

__device__ __forceinline__ void wbwd_point(const WBwdCoeffs& c, float w, float px, float py, float pz, float qx, float qy,
                                           float qz, float* dp, float* dq, float& dw) {
    const float u0 = fmaf(c.a[2], pz, fmaf(c.a[1], py, fmaf(c.a[0], px, c.bq[0])));
    const float u1 = fmaf(c.a[5], pz, fmaf(c.a[4], py, fmaf(c.a[3], px, c.bq[1])));
    const float u2 = fmaf(c.a[8], pz, fmaf(c.a[7], py, fmaf(c.a[6], px, c.bq[2])));
    const float v0 = fmaf(c.a[6], qz, fmaf(c.a[3], qy, fmaf(c.a[0], qx, c.bp[0])));
    const float v1 = fmaf(c.a[7], qz, fmaf(c.a[4], qy, fmaf(c.a[1], qx, c.bp[1])));
    const float v2 = fmaf(c.a[8], qz, fmaf(c.a[5], qy, fmaf(c.a[2], qx, c.bp[2])));
    dq[0] = w * u0; dq[1] = w * u1; dq[2] = w * u2;
    dp[0] = w * v0; dp[1] = w * v1; dp[2] = w * v2;
    dw = (qx - c.qb[0]) * u0 + (qy - c.qb[1]) * u1 + (qz - c.qb[2]) * u2 -
         (c.e[0] * (px - c.pb[0]) + c.e[1] * (py - c.pb[1]) + c.e[2] * (pz - c.pb[2]));
}

// npts % 4 == 0 and 16-byte aligned arrays: one thread = 4 points of one cloud (7 LDG.128, up to 7 STG.128)
__global__ void __launch_bounds__(256) kabsch_weighted_bwd_vec4_kernel(const WBwdParams prm) {
    pdl_wait();               // programmatic dependent launch (pe_common.h)
    pdl_launch_dependents();
    const long long groups = prm.b * prm.npts / 4;
    const long long gpc = prm.npts / 4;
    const long long step = (long long)gridDim.x * blockDim.x;
    const float4* sp = reinterpret_cast<const float4*>(prm.p);
    const float4* sq = reinterpret_cast<const float4*>(prm.q);
    const float4* swt = reinterpret_cast<const float4*>(prm.w);
    const long long first = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    CloudWalk walk(first, step, gpc);
    for (long long g = first; g < groups; g += step, walk.next()) {
        WBwdCoeffs c;
        wbwd_load(prm, walk.cloud, c);
        const float4 p0 = __ldcs(sp + 3 * g), p1 = __ldcs(sp + 3 * g + 1), p2 = __ldcs(sp + 3 * g + 2);
        const float4 q0 = __ldcs(sq + 3 * g), q1 = __ldcs(sq + 3 * g + 1), q2 = __ldcs(sq + 3 * g + 2);
        const float4 wv = __ldcs(swt + g);
        float dp[12], dq[12], dw[4];
        wbwd_point(c, wv.x, p0.x, p0.y, p0.z, q0.x, q0.y, q0.z, dp, dq, dw[0]);
        wbwd_point(c, wv.y, p0.w, p1.x, p1.y, q0.w, q1.x, q1.y, dp + 3, dq + 3, dw[1]);
        wbwd_point(c, wv.z, p1.z, p1.w, p2.x, q1.z, q1.w, q2.x, dp + 6, dq + 6, dw[2]);
        wbwd_point(c, wv.w, p2.y, p2.z, p2.w, q2.y, q2.z, q2.w, dp + 9, dq + 9, dw[3]);
        if (prm.dp) {
            float4* d = reinterpret_cast<float4*>(prm.dp) + 3 * g;
            __stcs(d, make_float4(dp[0], dp[1], dp[2], dp[3]));
            __stcs(d + 1, make_float4(dp[4], dp[5], dp[6], dp[7]));
            __stcs(d + 2, make_float4(dp[8], dp[9], dp[10], dp[11]));
        }
        if (prm.dq) {
            float4* d = reinterpret_cast<float4*>(prm.dq) + 3 * g;
            __stcs(d, make_float4(dq[0], dq[1], dq[2], dq[3]));
            __stcs(d + 1, make_float4(dq[4], dq[5], dq[6], dq[7]));
            __stcs(d + 2, make_float4(dq[8], dq[9], dq[10], dq[11]));
        }
        if (prm.dw) __stcs(reinterpret_cast<float4*>(prm.dw) + g, make_float4(dw[0], dw[1], dw[2], dw[3]));
    }
}

// any npts / alignment: one thread = one point
__global__ void __launch_bounds__(256) kabsch_weighted_bwd_scalar_kernel(const WBwdParams prm) {
    pdl_wait();               // programmatic dependent launch (pe_common.h)
    pdl_launch_dependents();
    const long long total = prm.b * prm.npts;
    const long long step = (long long)gridDim.x * blockDim.x;
    const long long first = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    CloudWalk walk(first, step, prm.npts);
    for (long long k = first; k < total; k += step, walk.next()) {
        WBwdCoeffs c;
        wbwd_load(prm, walk.cloud, c);
        float dp[3], dq[3], dw;
        wbwd_point(c, prm.w[k], prm.p[3 * k], prm.p[3 * k + 1], prm.p[3 * k + 2], prm.q[3 * k], prm.q[3 * k + 1],
                   prm.q[3 * k + 2], dp, dq, dw);
        if (prm.dp) { prm.dp[3 * k] = dp[0]; prm.dp[3 * k + 1] = dp[1]; prm.dp[3 * k + 2] = dp[2]; }
        if (prm.dq) { prm.dq[3 * k] = dq[0]; prm.dq[3 * k + 1] = dq[1]; prm.dq[3 * k + 2] = dq[2]; }
        if (prm.dw) prm.dw[k] = dw;
    }
}

}  // namespace pe

using namespace pe;

extern "C" int pe_kabsch_weighted_bwd_points(const float* p, const float* q, const float* w, const float* coef_q,
                                             const float* coef_p, const float* ctx, int64_t b, int64_t npts, float* dp,
                                             float* dq, float* dw, void* stream) {
    if (b < 0 || npts < 0) return fail(PE_ERR_BAD_ARG, "negative cloud count or cloud size");
    if (b == 0 || npts == 0) return 0;
    if (!p || !q || !w || !coef_q || !coef_p || !ctx) return fail(PE_ERR_BAD_ARG, "null pointer");
    if (!dp && !dq && !dw) return fail(PE_ERR_BAD_ARG, "all outputs are null");
    int dev = 0, sms = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return fail_cuda(e, "cudaDeviceGetAttribute");
    WBwdParams prm{p, q, w, coef_q, coef_p, ctx, dp, dq, dw, (long long)b, (long long)npts};
    const uintptr_t all = reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(w) |
                          reinterpret_cast<uintptr_t>(dp) | reinterpret_cast<uintptr_t>(dq) |
                          reinterpret_cast<uintptr_t>(dw);
    const bool vec = (npts % 4 == 0) && (all & 15u) == 0 && !force_direct();
    const long long work = vec ? b * npts / 4 : b * npts;
    long long grid = (work + 255) / 256;
    if (grid > (long long)sms * 8) grid = (long long)sms * 8;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (vec) launch_pdl(kabsch_weighted_bwd_vec4_kernel, (unsigned)grid, 256u, (size_t)0, st, prm);
    else launch_pdl(kabsch_weighted_bwd_scalar_kernel, (unsigned)grid, 256u, (size_t)0, st, prm);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, "kabsch_weighted_bwd_points");
    count_launch();
    return 0;
}

extern "C" int pe_affine_points(const float* a, int64_t a_stride, int a_row, int transpose, const float* t,
                                int64_t t_stride, int t_elem, const float* p, int64_t b, int64_t npts, float* out,
                                void* stream) {
    if (b < 0 || npts < 0) return fail(PE_ERR_BAD_ARG, "negative cloud count or cloud size");
    if (b == 0 || npts == 0) return 0;
    if (!a || !p || !out) return fail(PE_ERR_BAD_ARG, "null pointer");
    if (a_row < 3 || a_stride < 2 * a_row + 3 || (t && (t_elem < 1 || t_stride < 2 * t_elem + 1)))
        return fail(PE_ERR_BAD_ARG, "coefficient stride smaller than the record");
    int dev = 0, sms = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return fail_cuda(e, "cudaDeviceGetAttribute");
    AffineParams prm{a, t, p, out, (long long)b, (long long)npts, (long long)a_stride, (long long)t_stride, a_row,
                     t_elem, transpose};
    const bool vec = (npts % 4 == 0) && ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0 &&
                     !force_direct();
    const long long work = vec ? b * npts / 4 : b * npts;
    long long grid = (work + 255) / 256;
    if (grid > (long long)sms * 8) grid = (long long)sms * 8;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (vec) launch_pdl(affine_points_vec4_kernel, (unsigned)grid, 256u, (size_t)0, st, prm);
    else launch_pdl(affine_points_scalar_kernel, (unsigned)grid, 256u, (size_t)0, st, prm);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, "affine_points");
    count_launch();
    return 0;
}
