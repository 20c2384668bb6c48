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

// npts % 4 == 0 and 16-byte aligned arrays: one thread = 4 points of one cloud
__global__ void __launch_bounds__(256) affine_points_vec4_kernel(const AffineParams prm) {
    const long long groups = prm.b * prm.npts / 4;
    const long long gpc = prm.npts / 4;  // groups per cloud
    const long long step = (long long)gridDim.x * blockDim.x;
    const float4* src = reinterpret_cast<const float4*>(prm.p);
    float4* dst = reinterpret_cast<float4*>(prm.out);
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += step) {
        float m[9], t[3];
        load_coeffs(prm, g / gpc, m, t);
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
    const long long total = prm.b * prm.npts;
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += step) {
        float m[9], t[3];
        load_coeffs(prm, k / prm.npts, m, t);
        float ox, oy, oz;
        apply(m, t, prm.p[3 * k], prm.p[3 * k + 1], prm.p[3 * k + 2], ox, oy, oz);
        prm.out[3 * k] = ox; prm.out[3 * k + 1] = oy; prm.out[3 * k + 2] = oz;
    }
}

}  // namespace pe

using namespace pe;

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
    if (vec) affine_points_vec4_kernel<<<(unsigned)grid, 256, 0, st>>>(prm);
    else affine_points_scalar_kernel<<<(unsigned)grid, 256, 0, st>>>(prm);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, "affine_points");
    count_launch();
    return 0;
}
