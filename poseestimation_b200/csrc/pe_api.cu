// extern "C" entry points of libpe_rotation.so (declared in include/pe_rotation.h): library state, the light per-record
// ops (maps, losses, angle_error, pose update, Kabsch backward coefficients) and the scale kernels.  The SVD entry points
// are in pe_api_svd.cu, the single-pass step in pe_api_step.cu, pe_map_metric_fwd in pe_api_metric.cu.
#include "pe_api_run.cuh"

using namespace pe;

namespace {

// y = x * (*g): the backward of a single-pass step (the gradient for a unit upstream gradient times the upstream
// scalar, read on the device so that no host synchronisation is needed).  Pure stream: 8 B per word.
__global__ void __launch_bounds__(256) scale_by_scalar_kernel(const float4* __restrict__ x, const float* __restrict__ g,
                                                              float4* __restrict__ y, long long n4,
                                                              const float* __restrict__ xt, float* __restrict__ yt,
                                                              int tail) {
    pdl_wait();
    pdl_launch_dependents();
    const float k = __ldg(g);
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += step) {
        float4 v = __ldcs(x + i);
        v.x *= k; v.y *= k; v.z *= k; v.w *= k;
        __stcs(y + i, v);
    }
    if (blockIdx.x == 0 && (int)threadIdx.x < tail) yt[threadIdx.x] = xt[threadIdx.x] * k;
}

// x *= (*g) in place, and NOTHING when *g == 1: the backward of the single-pass step inside `loss.backward()`, where
// the upstream gradient of the loss is exactly 1 (every call site of the reference: Comparison/main.py:62,
// UPNA/main.py:58, 3D-Pose/main.py:86).  The test is made on the device, so there is no host synchronisation; when it
// holds the launch costs a few microseconds instead of a 72 B/matrix pass.
__global__ void __launch_bounds__(1024) scale_inplace_kernel(float4* __restrict__ x, const float* __restrict__ g,
                                                            long long n4, float* __restrict__ xt, int tail) {
    pdl_wait();
    pdl_launch_dependents();
    const float k = __ldg(g);
    if (k == 1.0f) return;
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += step) {
        float4 v = x[i];
        v.x *= k; v.y *= k; v.z *= k; v.w *= k;
        x[i] = v;
    }
    if (blockIdx.x == 0 && (int)threadIdx.x < tail) xt[threadIdx.x] *= k;
}

}  // namespace

extern "C" {

int pe_version(void) { return PE_VERSION; }
const char* pe_last_error(void) { return last_error_buf(); }
long long pe_launch_count(void) { return launch_counter().load(); }
int pe_init(int device) {
    int dev = device;
    if (dev < 0) {
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return fail_cuda(e, "cudaGetDevice");
    }
    return slots_init(dev);
}

#define PE_MAP_API(fn, Tag)                                                                                      \
    int pe_##fn##_fwd(const float* x, int64_t stride, float* r, int64_t n, void* stream) {                       \
        return run<Tag##Fwd>(x, stride, nullptr, nullptr, r, nullptr, n, nullptr, nullptr, {}, stream);          \
    }                                                                                                            \
    int pe_##fn##_bwd(const float* x, int64_t stride, const float* g, float* dx, int64_t n, void* stream) {      \
        return run<Tag##Bwd>(x, stride, g, nullptr, dx, nullptr, n, nullptr, nullptr, {}, stream);               \
    }
PE_MAP_API(ortho6d, Ortho6d)
PE_MAP_API(quat, Quat)
PE_MAP_API(euler, Euler)
PE_MAP_API(ortho5d, Ortho5d)
PE_MAP_API(expmap, Expmap)
int pe_expmap_eps_fwd(const float* v, int64_t stride, float eps, float* r, int64_t n, void* stream) {
    if (!(eps > 0.f)) return fail(PE_ERR_BAD_ARG, "eps must be positive");
    return run<ExpmapEpsFwd>(v, stride, nullptr, nullptr, r, nullptr, n, nullptr, nullptr, {eps}, stream);
}
int pe_expmap_eps_bwd(const float* v, int64_t stride, const float* g, float eps, float* dv, int64_t n, void* stream) {
    if (!(eps > 0.f)) return fail(PE_ERR_BAD_ARG, "eps must be positive");
    return run<ExpmapEpsBwd>(v, stride, g, nullptr, dv, nullptr, n, nullptr, nullptr, {eps}, stream);
}

int pe_frobenius_fwd(const float* a, const float* b, double* sum, int64_t n, void* stream) {
    return run<FrobFwd>(a, 9, b, nullptr, nullptr, nullptr, n, sum, nullptr, {}, stream);
}
int pe_frobenius_mean_fwd(const float* a, const float* b, double* sum, float* mean, int64_t n, void* stream) {
    if (!mean) return fail(PE_ERR_BAD_ARG, "null mean pointer");
    return run<FrobFwd>(a, 9, b, nullptr, nullptr, nullptr, n, sum, nullptr, {}, stream, mean,
                        n > 0 ? 1.0 / (double)n : 0.0);
}
int pe_frobenius_bwd(const float* a, const float* b, const float* gout, float inv_n, float* da, float* db,
                     int64_t n, void* stream) {
    if (!gout) return fail(PE_ERR_BAD_ARG, "null upstream-gradient pointer");
    if (da && db) return run<FrobBwd<true>>(a, 9, b, nullptr, da, db, n, nullptr, nullptr, {gout, inv_n}, stream);
    if (da) return run<FrobBwd<false>>(a, 9, b, nullptr, da, nullptr, n, nullptr, nullptr, {gout, inv_n}, stream);
    if (db) return run<FrobBwd<false>>(a, 9, b, nullptr, db, nullptr, n, nullptr, nullptr, {gout, -inv_n}, stream);
    return fail(PE_ERR_BAD_ARG, "both gradient outputs are null");
}

int pe_geodesic_fwd(const float* m1, const float* m2, float lo, float hi, float* theta, double* sum, int64_t n,
                    void* stream) {
    Clamp c{lo, hi};
    if (theta && sum) return run<GeoFwd<true, true>>(m1, 9, m2, nullptr, theta, nullptr, n, sum, nullptr, c, stream);
    if (theta) return run<GeoFwd<true, false>>(m1, 9, m2, nullptr, theta, nullptr, n, nullptr, nullptr, c, stream);
    if (sum) return run<GeoFwd<false, true>>(m1, 9, m2, nullptr, nullptr, nullptr, n, sum, nullptr, c, stream);
    return fail(PE_ERR_BAD_ARG, "both outputs are null");
}
int pe_geodesic_reduce_fwd(const float* m1, const float* m2, float lo, float hi, double* sum, float* out,
                           double scale, int64_t n, void* stream) {
    if (!out) return fail(PE_ERR_BAD_ARG, "null output pointer");
    Clamp c{lo, hi};
    return run<GeoFwd<false, true>>(m1, 9, m2, nullptr, nullptr, nullptr, n, sum, nullptr, c, stream, out, scale);
}
int pe_geodesic_bwd(const float* m1, const float* m2, float lo, float hi, const float* g, int g_is_scalar,
                    float gmul, float* d1, float* d2, int64_t n, void* stream) {
    if (!g) return fail(PE_ERR_BAD_ARG, "null upstream-gradient pointer");
    GeoGrad ex{lo, hi, gmul, g};
    const float* gv = g_is_scalar ? nullptr : g;
#define PE_GEO(MODE, O0, O1)                                                                                   \
    (g_is_scalar ? run<GeoBwd<MODE, true>>(m1, 9, m2, nullptr, O0, O1, n, nullptr, nullptr, ex, stream)        \
                 : run<GeoBwd<MODE, false>>(m1, 9, m2, gv, O0, O1, n, nullptr, nullptr, ex, stream))
    if (d1 && d2) return PE_GEO(2, d1, d2);
    if (d1) return PE_GEO(0, d1, nullptr);
    if (d2) return PE_GEO(1, d2, nullptr);
#undef PE_GEO
    return fail(PE_ERR_BAD_ARG, "both gradient outputs are null");
}

int pe_angle_error_fwd(const float* r1, const float* r2, double* deg, double* sum, int* flag, int64_t n,
                       void* stream) {
    if (reinterpret_cast<uintptr_t>(deg) & 7u) return fail(PE_ERR_BAD_ARG, "float64 output not aligned to 8 bytes");
    float* d = reinterpret_cast<float*>(deg);
    if (deg && sum) return run<AngleFwd<true, true>>(r1, 9, r2, nullptr, d, nullptr, n, sum, flag, {}, stream);
    if (deg) return run<AngleFwd<true, false>>(r1, 9, r2, nullptr, d, nullptr, n, nullptr, flag, {}, stream);
    if (sum) return run<AngleFwd<false, true>>(r1, 9, r2, nullptr, nullptr, nullptr, n, sum, flag, {}, stream);
    return fail(PE_ERR_BAD_ARG, "both outputs are null");
}

int pe_axis_angle_fwd(const float* theta, const float* axis, float* r, int64_t n, void* stream) {
    return run<AxisAngleFwd>(theta, 1, axis, nullptr, r, nullptr, n, nullptr, nullptr, {}, stream);
}
int pe_kabsch_bwd_coeffs(const float* ctx, const float* r, const float* g, int64_t npts, float* coef_q, float* coef_p,
                         int64_t b, void* stream) {
    if (npts < 1) return fail(PE_ERR_BAD_ARG, "bad cloud size");
    return run<KabschBwdCoeffs>(ctx, 15, r, g, coef_q, coef_p, b, nullptr, nullptr, {1.0f / (float)npts}, stream);
}
int pe_kabsch_weighted_bwd_coeffs(const float* ctx, const float* r, const float* g, float* coef_q, float* coef_p,
                                  int64_t b, void* stream) {
    return run<KabschBwdCoeffsW>(ctx, 16, r, g, coef_q, coef_p, b, nullptr, nullptr, {}, stream);
}
int pe_pose_update_fwd(const float* out, int64_t out_stride, const float* t_init, float fx, float fy, float* t_pred,
                       int64_t n, void* stream) {
    return run<PoseUpdateFwd>(out, out_stride, t_init, nullptr, t_pred, nullptr, n, nullptr, nullptr,
                              {1.0f / fx, 1.0f / fy}, stream);
}
int pe_pose_update_bwd(const float* out, int64_t out_stride, const float* t_init, const float* g_t, float fx, float fy,
                       float* d_out, int64_t n, void* stream) {
    return run<PoseUpdateBwd>(out, out_stride, t_init, g_t, d_out, nullptr, n, nullptr, nullptr,
                              {1.0f / fx, 1.0f / fy}, stream);
}
int pe_scale_by_scalar(const float* x, const float* g, float* y, int64_t nwords, void* stream) {
    if (nwords < 0) return fail(PE_ERR_BAD_ARG, "negative word count");
    if (nwords == 0) return 0;
    if (!x || !g || !y) return fail(PE_ERR_BAD_ARG, "null pointer");
    if (!aligned16(x) || !aligned16(y)) return fail(PE_ERR_BAD_ARG, "pe_scale_by_scalar needs 16-byte aligned arrays");
    int dev = 0, sms = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return fail_cuda(e, "cudaDeviceGetAttribute");
    const long long n4 = nwords / 4;
    long long grid = (n4 + 255) / 256;
    if (grid > (long long)sms * 8) grid = (long long)sms * 8;
    if (grid < 1) grid = 1;
    scale_by_scalar_kernel<<<(unsigned)grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const float4*>(x), g, reinterpret_cast<float4*>(y), n4, x + 4 * n4, y + 4 * n4,
        (int)(nwords - 4 * n4));
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, "scale_by_scalar");
    count_launch();
    return 0;
}

int pe_scale_by_scalar_inplace(float* x, const float* g, int64_t nwords, void* stream) {
    if (nwords < 0) return fail(PE_ERR_BAD_ARG, "negative word count");
    if (nwords == 0) return 0;
    if (!x || !g) return fail(PE_ERR_BAD_ARG, "null pointer");
    if (!aligned16(x)) return fail(PE_ERR_BAD_ARG, "pe_scale_by_scalar_inplace needs a 16-byte aligned array");
    int dev = 0, sms = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return fail_cuda(e, "cudaDeviceGetAttribute");
    const long long n4 = nwords / 4;
    // few, fat CTAs: when *g == 1 the launch is pure overhead, and that is the CTA count (2 x 1024 threads still fill an SM)
    long long grid = (n4 + 1023) / 1024;
    if (grid > (long long)sms * 2) grid = (long long)sms * 2;
    if (grid < 1) grid = 1;
    e = launch_pdl(scale_inplace_kernel, (unsigned)grid, 1024u, 0, static_cast<cudaStream_t>(stream),
                   reinterpret_cast<float4*>(x), g, (long long)n4, x + 4 * n4, (int)(nwords - 4 * n4));
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, "scale_inplace");
    count_launch();
    return 0;
}

}  // extern "C"
