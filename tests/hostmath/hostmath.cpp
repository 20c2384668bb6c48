// TEST INFRASTRUCTURE: the per-record formulas of poseestimation_b200/csrc/pe_math.cuh compiled for the HOST with
// g++, so that the CPU test-suite can check the exact arithmetic the kernels run against the oracle without a GPU.
// Nothing in the package loads this library.
#include "../../poseestimation_b200/csrc/pe_lanes.cuh"

#define LOOP for (long i = 0; i < n; ++i)
extern "C" {
void hm_svd_orth(const float* m, float* r, float* d, long n) { LOOP pe::svd_orth_fast(m + 9 * i, r + 9 * i, d[i]); }
// the 4-sweep, V-accumulating variant the production one was derived from (kept for A/B tests)
void hm_svd_orth_4sweep(const float* m, float* r, float* d, long n) { LOOP pe::svd_orth<pe::kSweeps>(m + 9 * i, r + 9 * i, d[i]); }
void hm_svd_orth_bwd(const float* m, const float* r, const float* g, float* dm, long n) {
    LOOP pe::svd_orth_pullback(m + 9 * i, r + 9 * i, g + 9 * i, dm + 9 * i);
}
void hm_ortho6d(const float* x, float* r, long n) { LOOP pe::ortho6d(x + 6 * i, r + 9 * i); }
void hm_ortho6d_bwd(const float* x, const float* g, float* dx, long n) { LOOP pe::ortho6d_pullback(x + 6 * i, g + 9 * i, dx + 6 * i); }
void hm_quat(const float* x, float* r, long n) { LOOP pe::quat_map(x + 4 * i, r + 9 * i); }
void hm_quat_bwd(const float* x, const float* g, float* dx, long n) { LOOP pe::quat_pullback(x + 4 * i, g + 9 * i, dx + 4 * i); }
void hm_euler(const float* x, float* r, long n) { LOOP pe::euler_map(x + 3 * i, r + 9 * i); }
void hm_euler_bwd(const float* x, const float* g, float* dx, long n) { LOOP pe::euler_pullback(x + 3 * i, g + 9 * i, dx + 3 * i); }
void hm_ortho5d(const float* x, float* r, long n) { LOOP pe::ortho5d(x + 5 * i, r + 9 * i); }
void hm_ortho5d_bwd(const float* x, const float* g, float* dx, long n) { LOOP pe::ortho5d_pullback(x + 5 * i, g + 9 * i, dx + 5 * i); }
void hm_expmap(const float* x, float* r, long n) { LOOP pe::expmap(x + 3 * i, r + 9 * i); }
void hm_expmap_bwd(const float* x, const float* g, float* dx, long n) { LOOP pe::expmap_pullback(x + 3 * i, g + 9 * i, dx + 3 * i); }
void hm_expmap_eps(const float* x, const float* eps, float* r, long n) { LOOP pe::expmap(x + 3 * i, r + 9 * i, eps[0]); }
void hm_expmap_eps_bwd(const float* x, const float* g, const float* eps, float* dx, long n) {
    LOOP pe::expmap_pullback(x + 3 * i, g + 9 * i, dx + 3 * i, eps[0]);
}
// angle_error's float64 tail: the Newton acos (degrees) and the full per-pair angle with its range flag
void hm_acos_deg(const double* c, double* deg, long n) { LOOP deg[i] = pe::acos_deg_fast(c[i]); }
void hm_angle_deg(const float* a, const float* b, double* deg, int* flag, long n) {
    LOOP { int f = 0; deg[i] = pe::angle_deg_f64(a + 9 * i, b + 9 * i, f); flag[i] = f; }
}
// the fused sums' angle: float64 trace + float32 tail (scalar kernels), and the error-free float32 trace + float32 tail
// in the packed two-lane form of the SVD kernels (record i rides with record i + n/2)
void hm_angle_sum_term(const float* a, const float* b, double* deg, int* flag, long n) {
    LOOP { int f = 0; deg[i] = pe::angle_deg_sum_term(pe::dot9_f64(a + 9 * i, b + 9 * i), f); flag[i] = f; }
}
void hm_angle_split_scalar(const float* a, const float* b, double* deg, float* ahi_lo, long n) {
    LOOP {
        float ahi, lo, a32;
        pe::trace_minus_one_split_scalar(a + 9 * i, b + 9 * i, ahi, lo);
        bool neg;
        float t = pe::half_versine_from_split(ahi, lo, neg, a32);
        deg[i] = pe::angle_deg_tail32_t(t, neg);
        ahi_lo[2 * i] = ahi; ahi_lo[2 * i + 1] = lo;
    }
}
void hm_angle_split_x2(const float* a, const float* b, double* deg, float* ahi_lo, long n) {
    long h = n / 2;
    for (long i = 0; i < h; ++i) {
        pe::F2 aa[9], bb[9], ahi, lo, a32;
        for (int k = 0; k < 9; ++k) { aa[k] = pe::F2{a[9 * i + k], a[9 * (i + h) + k]}; bb[k] = pe::F2{b[9 * i + k], b[9 * (i + h) + k]}; }
        pe::trace_minus_one_split(aa, bb, ahi, lo);
        pe::B2 neg;
        pe::F2 t = pe::half_versine_from_split(ahi, lo, neg, a32);
        pe::F2 d = pe::angle_deg_tail32_t(t, neg);
        deg[i] = d.x; deg[i + h] = d.y;
        ahi_lo[2 * i] = ahi.x; ahi_lo[2 * i + 1] = lo.x; ahi_lo[2 * (i + h)] = ahi.y; ahi_lo[2 * (i + h) + 1] = lo.y;
    }
}
// sampler map, Kabsch backward coefficients, CosyPose pose update (fx = fy = 4000/9, utility.py:73-88)
void hm_axis_angle(const float* th, const float* ax, float* r, long n) { LOOP pe::axis_angle_matrix(th[i], ax + 3 * i, r + 9 * i); }
void hm_kabsch_bwd_coeffs(const float* ctx, const float* r, const float* g, const float* inv_n, float* cq, float* cp, long n) {
    LOOP pe::kabsch_bwd_coeffs(ctx + 15 * i, r + 9 * i, g + 12 * i, inv_n[i], cq + 12 * i, cp + 12 * i);
}
void hm_pose_update(const float* o, const float* ti, float* to, long n) {
    LOOP pe::pose_update(o + 12 * i, ti + 16 * i, 9.0f / 4000.0f, 9.0f / 4000.0f, to + 16 * i);
}
void hm_pose_update_bwd(const float* o, const float* ti, const float* g, float* d, long n) {
    LOOP pe::pose_update_pullback(o + 12 * i, ti + 16 * i, g + 16 * i, 9.0f / 4000.0f, 9.0f / 4000.0f, d + 12 * i);
}
// the two-matrices-per-thread (packed fp32x2) instantiation: record i rides with record i + n/2 (n even)
void hm_svd_orth_x2(const float* m, float* r, float* d, long n) {
    long h = n / 2;
    for (long i = 0; i < h; ++i) {
        pe::F2 mm[9], rr[9], ds;
        for (int k = 0; k < 9; ++k) mm[k] = pe::F2{m[9 * i + k], m[9 * (i + h) + k]};
        pe::svd_orth_t(mm, rr, ds);
        for (int k = 0; k < 9; ++k) { r[9 * i + k] = rr[k].x; r[9 * (i + h) + k] = rr[k].y; }
        d[i] = ds.x; d[i + h] = ds.y;
    }
}
// pullback from the decomposition's own (V, singular values), as the recompute / fused kernels run it
void hm_svd_orth_bwd_ctx_x2(const float* m, const float* g, float* dm, long n) {
    long h = n / 2;
    for (long i = 0; i < h; ++i) {
        pe::F2 mm[9], rr[9], gg[9], dd[9], ds;
        for (int k = 0; k < 9; ++k) {
            mm[k] = pe::F2{m[9 * i + k], m[9 * (i + h) + k]};
            gg[k] = pe::F2{g[9 * i + k], g[9 * (i + h) + k]};
        }
        pe::SvdCtx<pe::F2> ctx;
        auto fetch = [&mm](pe::F2* dst) { for (int k = 0; k < 9; ++k) dst[k] = mm[k]; };
        pe::svd_orth_f<pe::F2>(fetch, rr, ds, &ctx);
        pe::F2 z0, z1, z2;
        pe::skew_axial_rtg(rr, gg, z0, z1, z2);
        pe::svd_orth_pullback_ctx<pe::F2>(ctx, fetch, rr, z0, z1, z2, dd);
        for (int k = 0; k < 9; ++k) { dm[9 * i + k] = dd[k].x; dm[9 * (i + h) + k] = dd[k].y; }
    }
}
void hm_svd_orth_bwd_x2(const float* m, const float* r, const float* g, float* dm, long n) {
    long h = n / 2;
    for (long i = 0; i < h; ++i) {
        pe::F2 mm[9], rr[9], gg[9], dd[9];
        for (int k = 0; k < 9; ++k) {
            mm[k] = pe::F2{m[9 * i + k], m[9 * (i + h) + k]};
            rr[k] = pe::F2{r[9 * i + k], r[9 * (i + h) + k]};
            gg[k] = pe::F2{g[9 * i + k], g[9 * (i + h) + k]};
        }
        pe::svd_orth_pullback_t(mm, rr, gg, dd);
        for (int k = 0; k < 9; ++k) { dm[9 * i + k] = dd[k].x; dm[9 * (i + h) + k] = dd[k].y; }
    }
}
}
