// Lane abstraction for the SVD math: the same formulas instantiate for T = float (one matrix per thread) and
// T = F2 (TWO matrices per thread, one in each half of a 64-bit register pair).  On sm_100 the F2 arithmetic maps
// to the packed fp32x2 instructions FMUL2 / FFMA2 / FADD2: one issue slot does the work for both matrices.  The SVD
// kernels are issue-bound, not HBM-bound, so this is the lever that matters.
#pragma once
#include "pe_math.cuh"

namespace pe {

struct B2 {
    bool x, y;
};

// ---- float ---------------------------------------------------------------------------------------------------
PE_DEV float t_set(float, float v) { return v; }  // first argument only selects the overload
PE_DEV float t_mul(float a, float b) { return a * b; }
PE_DEV float t_fma(float a, float b, float c) { return fmaf(a, b, c); }
PE_DEV float t_add(float a, float b) { return a + b; }
PE_DEV float t_sub(float a, float b) { return a - b; }
PE_DEV float t_neg(float a) { return -a; }
PE_DEV float t_abs(float a) { return fabsf(a); }
PE_DEV float t_max(float a, float b) { return fmaxf(a, b); }
PE_DEV float t_rsq(float a) { return rsqrt_fast(a); }
PE_DEV float t_sqrt(float a) { return sqrtf(a); }
PE_DEV float t_sqrt_fast(float a) { return sqrt_fast(a); }
PE_DEV bool t_ltz(float a) { return a < 0.f; }
PE_DEV float t_sel(bool m, float a, float b) { return m ? a : b; }
PE_DEV float t_prescale(const float* m) { return pow2_prescale(m); }

// ---- F2 ------------------------------------------------------------------------------------------------------
PE_DEV F2 t_set(F2, float v) { return F2{v, v}; }
PE_DEV F2 t_mul(F2 a, F2 b) { return f2_mul(a, b); }
PE_DEV F2 t_fma(F2 a, F2 b, F2 c) { return f2_fma(a, b, c); }
PE_DEV F2 t_add(F2 a, F2 b) {
#ifdef __CUDA_ARCH__
    float2 r = __fadd2_rn(make_float2(a.x, a.y), make_float2(b.x, b.y));
    return F2{r.x, r.y};
#else
    return F2{a.x + b.x, a.y + b.y};
#endif
}
PE_DEV F2 t_neg(F2 a) { return F2{-a.x, -a.y}; }
PE_DEV F2 t_sub(F2 a, F2 b) { return t_add(a, t_neg(b)); }
PE_DEV F2 t_abs(F2 a) { return F2{fabsf(a.x), fabsf(a.y)}; }
PE_DEV F2 t_max(F2 a, F2 b) { return F2{fmaxf(a.x, b.x), fmaxf(a.y, b.y)}; }
PE_DEV F2 t_rsq(F2 a) { return F2{rsqrt_fast(a.x), rsqrt_fast(a.y)}; }
PE_DEV F2 t_sqrt(F2 a) { return F2{sqrtf(a.x), sqrtf(a.y)}; }
PE_DEV F2 t_sqrt_fast(F2 a) { return F2{sqrt_fast(a.x), sqrt_fast(a.y)}; }
PE_DEV B2 t_ltz(F2 a) { return B2{a.x < 0.f, a.y < 0.f}; }
PE_DEV F2 t_sel(B2 m, F2 a, F2 b) { return F2{m.x ? a.x : b.x, m.y ? a.y : b.y}; }
PE_DEV F2 t_prescale(const F2* m) {
    float a[9], b[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) { a[i] = m[i].x; b[i] = m[i].y; }
    return F2{pow2_prescale(a), pow2_prescale(b)};
}
PE_DEV float lane_get(float v, int) { return v; }
PE_DEV float lane_get(F2 v, int k) { return k ? v.y : v.x; }
PE_DEV void lane_put(float& v, int, float s) { v = s; }
PE_DEV void lane_put(F2& v, int k, float s) { if (k) v.y = s; else v.x = s; }
template <class T> struct LaneCount { static constexpr int value = 1; };
template <> struct LaneCount<F2> { static constexpr int value = 2; };

// ---- the Jacobi rotation, one (or two) matrices, columns as three separate rows ----------------------------------
template <bool FRESH, class T>
PE_DEV void hestenes_t(T& p0, T& p1, T& p2, T& q0, T& q1, T& q2, T& np, T& nq) {
    const T z = p0;  // overload selector
    // squared norms: tracked from the previous rotations, or recomputed (see hestenes_a in pe_math.cuh)
    T al = np, be = nq;
    if constexpr (FRESH) {
        al = t_fma(p2, p2, t_fma(p1, p1, t_mul(p0, p0)));
        be = t_fma(q2, q2, t_fma(q1, q1, t_mul(q0, q0)));
    }
    T ga = t_fma(p2, q2, t_fma(p1, q1, t_mul(p0, q0)));
    T dh = t_mul(t_set(z, 0.5f), t_sub(al, be));
    T ad = t_max(t_abs(dh), t_set(z, 1e-18f));
    // tan(theta) = ga / (rho + |dh|) with rho = sqrt(dh^2 + ga^2)  (= sin 2theta / (1 + cos 2theta)): no cancellation,
    // two MUFU ops, and (c, s) = (u, ga) / sqrt(u^2 + ga^2) share one reciprocal square root
    T g2 = t_mul(ga, ga);
    T rho = t_sqrt_fast(t_fma(ad, ad, g2));
    T u = t_add(rho, ad);
    T k = t_rsq(t_fma(u, u, g2));
    T c = t_mul(u, k), s = t_mul(ga, k);
    T mean = t_sub(al, dh);
    np = t_add(mean, rho); nq = t_sub(mean, rho);
    auto neg = t_ltz(dh);
    T cp = t_sel(neg, s, c), sp = t_sel(neg, c, s), ns = t_neg(sp);
    T t;
    t = p0; p0 = t_fma(cp, t, t_mul(sp, q0)); q0 = t_fma(cp, q0, t_mul(ns, t));
    t = p1; p1 = t_fma(cp, t, t_mul(sp, q1)); q1 = t_fma(cp, q1, t_mul(ns, t));
    t = p2; p2 = t_fma(cp, t, t_mul(sp, q2)); q2 = t_fma(cp, q2, t_mul(ns, t));
}

// Same algorithm as svd_orth_fast (pe_math.cuh), lane-generic.  Degenerate lanes (rank <= 1, zero, NaN) are
// detected and redone with the scalar routine, which owns the special cases.
//
// `fetch(T* dst)` delivers the nine entries of M.  It is called twice (three times for a degenerate lane): once for
// the Jacobi iteration and once more, afterwards, for the recovery of V - so M does not sit in registers while the
// rotations run (the kernels re-read it from shared memory; see PairSrc in pe_stream.cuh).
// What the decomposition leaves for a pullback in the same thread (12 registers per lane pair): the symmetric matrix
// A = (tr(P) I - P)^-1 (regularised, see svd_orth_pullback_ctx) as a00, a01, a02, a11, a12, a22, and a bit mask of lanes
// that took the scalar route (their A is not valid).
template <class T>
struct SvdCtx {
    T a[6];
    int deg;
};

// The scalar routine for degenerate lanes.  PE_FALLBACK_NOINLINE=1 keeps it out of line (one copy per kernel instead of
// one per lane, its registers not part of the hot path's allocation); A/B switch, see DESIGN.md 4.2.
#ifndef PE_FALLBACK_NOINLINE
#define PE_FALLBACK_NOINLINE 0
#endif
#if PE_FALLBACK_NOINLINE && defined(__CUDACC__)
static __device__ __noinline__ void svd_orth_fast_cold(const float* m, float* r, float* ds) {
    float d;
    svd_orth_fast(m, r, d);
    *ds = d;
}
#endif
PE_DEV void svd_orth_degenerate_lane(const float* m, float* r, float& ds) {
#if PE_FALLBACK_NOINLINE && defined(__CUDA_ARCH__)
    svd_orth_fast_cold(m, r, &ds);
#else
    svd_orth_fast(m, r, ds);
#endif
}

template <class T, class Fetch>
PE_DEV void svd_orth_f(const Fetch& fetch, T* r, T& dsign, SvdCtx<T>* ctx = nullptr) {
    T a[9];
    fetch(a);
    const T z = a[0];
    const T sc = t_prescale(a);
    // column j = (c[j], c[3+j], c[6+j])
    T c00 = t_mul(a[0], sc), c01 = t_mul(a[1], sc), c02 = t_mul(a[2], sc), c10 = t_mul(a[3], sc),
      c11 = t_mul(a[4], sc), c12 = t_mul(a[5], sc), c20 = t_mul(a[6], sc), c21 = t_mul(a[7], sc),
      c22 = t_mul(a[8], sc);
    T q0n = t_fma(c20, c20, t_fma(c10, c10, t_mul(c00, c00)));
    T q1n = t_fma(c21, c21, t_fma(c11, c11, t_mul(c01, c01)));
    T q2n = t_fma(c22, c22, t_fma(c12, c12, t_mul(c02, c02)));
    hestenes_t<false>(c00, c10, c20, c01, c11, c21, q0n, q1n);
    hestenes_t<false>(c00, c10, c20, c02, c12, c22, q0n, q2n);
    hestenes_t<false>(c01, c11, c21, c02, c12, c22, q1n, q2n);
    hestenes_t<false>(c00, c10, c20, c02, c12, c22, q0n, q2n);
    hestenes_t<false>(c00, c10, c20, c01, c11, c21, q0n, q1n);
    hestenes_t<false>(c01, c11, c21, c02, c12, c22, q1n, q2n);
    hestenes_t<false>(c00, c10, c20, c02, c12, c22, q0n, q2n);
    hestenes_t<(kFreshTail >= 3)>(c00, c10, c20, c01, c11, c21, q0n, q1n);
    hestenes_t<(kFreshTail >= 2)>(c00, c10, c20, c02, c12, c22, q0n, q2n);
    hestenes_t<(kFreshTail >= 1)>(c01, c11, c21, c02, c12, c22, q1n, q2n);
    T al0 = t_fma(c20, c20, t_fma(c10, c10, t_mul(c00, c00)));
    T al1 = t_fma(c21, c21, t_fma(c11, c11, t_mul(c01, c01)));
    T k0 = t_rsq(al0), k1 = t_rsq(al1);
    T u00 = t_mul(c00, k0), u10 = t_mul(c10, k0), u20 = t_mul(c20, k0);
    T u01 = t_mul(c01, k1), u11 = t_mul(c11, k1), u21 = t_mul(c21, k1);
    T u02 = t_fma(u10, u21, t_neg(t_mul(u20, u11)));
    T u12 = t_fma(u20, u01, t_neg(t_mul(u00, u21)));
    T u22 = t_fma(u00, u11, t_neg(t_mul(u10, u01)));
    T dd = t_fma(c02, u02, t_fma(c12, u12, t_mul(c22, u22)));
    dsign = t_sel(t_ltz(dd), t_set(z, -1.f), t_set(z, 1.f));
    T l2 = z, e0 = z, e1 = z;  // ctx only: regularised inverse eigenvalues of tr(P) I - P (see svd_orth_pullback_ctx)
    if (ctx) {
        // singular values of the prescaled matrix, the sign of det(M) on the third: P = R^T M = V diag(s0, s1, s2) V^T
        const T s0 = t_mul(al0, k0), s1 = t_mul(al1, k1), s2 = dd;
        T a = t_add(s1, s2), b = t_add(s0, s2), cc = t_add(s0, s1);
        T ab = t_mul(a, b), bc = t_mul(b, cc), ac = t_mul(a, cc);
        T det = t_mul(ab, cc);
        T tr3 = t_mul(t_add(t_add(a, b), cc), t_set(z, 1.0f / 3.0f));
        T fl = t_mul(t_set(z, kDetFloor), t_abs(t_mul(t_mul(tr3, tr3), tr3)));
        det = t_max(det, fl);
        T inv = z;
#pragma unroll
        for (int k = 0; k < LaneCount<T>::value; ++k) {
            float d = lane_get(det, k);
            lane_put(inv, k, d > 1e-30f ? lane_get(sc, k) * rcp_fast(d) : 0.f);
        }
        l2 = t_mul(ab, inv);
        e0 = t_sub(t_mul(bc, inv), l2); e1 = t_sub(t_mul(ac, inv), l2);
    }
    // right vectors from the left ones: w_j = (sc M)^T u_j, Gram-Schmidt, cross product
    fetch(a);
#pragma unroll
    for (int i = 0; i < 9; ++i) a[i] = t_mul(a[i], sc);
    T w00 = t_fma(a[6], u20, t_fma(a[3], u10, t_mul(a[0], u00)));
    T w10 = t_fma(a[7], u20, t_fma(a[4], u10, t_mul(a[1], u00)));
    T w20 = t_fma(a[8], u20, t_fma(a[5], u10, t_mul(a[2], u00)));
    T w01 = t_fma(a[6], u21, t_fma(a[3], u11, t_mul(a[0], u01)));
    T w11 = t_fma(a[7], u21, t_fma(a[4], u11, t_mul(a[1], u01)));
    T w21 = t_fma(a[8], u21, t_fma(a[5], u11, t_mul(a[2], u01)));
    // Gram-Schmidt without waiting for the first normalisation: n0 w1 - (w1 . w0) w0 has the direction of
    // w1 - (w1 . v0) v0, so the two reciprocal square roots are independent (this tail is latency-bound)
    T n0 = t_fma(w20, w20, t_fma(w10, w10, t_mul(w00, w00)));
    T pr = t_neg(t_fma(w21, w20, t_fma(w11, w10, t_mul(w01, w00))));
    T j0 = t_rsq(n0);
    w01 = t_fma(pr, w00, t_mul(n0, w01)); w11 = t_fma(pr, w10, t_mul(n0, w11)); w21 = t_fma(pr, w20, t_mul(n0, w21));
    T v00 = t_mul(w00, j0), v10 = t_mul(w10, j0), v20 = t_mul(w20, j0);
    T n1 = t_fma(w21, w21, t_fma(w11, w11, t_mul(w01, w01)));
    T j1 = t_rsq(n1);
    T v01 = t_mul(w01, j1), v11 = t_mul(w11, j1), v21 = t_mul(w21, j1);
    T v02 = t_fma(v10, v21, t_neg(t_mul(v20, v11)));
    T v12 = t_fma(v20, v01, t_neg(t_mul(v00, v21)));
    T v22 = t_fma(v00, v11, t_neg(t_mul(v10, v01)));
    r[0] = t_fma(u02, v02, t_fma(u01, v01, t_mul(u00, v00)));
    r[1] = t_fma(u02, v12, t_fma(u01, v11, t_mul(u00, v10)));
    r[2] = t_fma(u02, v22, t_fma(u01, v21, t_mul(u00, v20)));
    r[3] = t_fma(u12, v02, t_fma(u11, v01, t_mul(u10, v00)));
    r[4] = t_fma(u12, v12, t_fma(u11, v11, t_mul(u10, v10)));
    r[5] = t_fma(u12, v22, t_fma(u11, v21, t_mul(u10, v20)));
    r[6] = t_fma(u22, v02, t_fma(u21, v01, t_mul(u20, v00)));
    r[7] = t_fma(u22, v12, t_fma(u21, v11, t_mul(u20, v10)));
    r[8] = t_fma(u22, v22, t_fma(u21, v21, t_mul(u20, v20)));
    if (ctx) {
        // A = V diag(l0, l1, l2) V^T with (l0, l1, l2) = (bc, ac, ab) / max(abc, floor); since sum_j v_j v_j^T = I,
        // A = l2 I + (l0 - l2) v0 v0^T + (l1 - l2) v1 v1^T
        const T p00 = t_mul(e0, v00), p10 = t_mul(e0, v10), p20 = t_mul(e0, v20);
        const T p01 = t_mul(e1, v01), p11 = t_mul(e1, v11), p21 = t_mul(e1, v21);
        ctx->a[0] = t_fma(p01, v01, t_fma(p00, v00, l2));
        ctx->a[1] = t_fma(p01, v11, t_mul(p00, v10));
        ctx->a[2] = t_fma(p01, v21, t_mul(p00, v20));
        ctx->a[3] = t_fma(p11, v11, t_fma(p10, v10, l2));
        ctx->a[4] = t_fma(p11, v21, t_mul(p10, v20));
        ctx->a[5] = t_fma(p21, v21, t_fma(p20, v20, l2));
        ctx->deg = 0;
    }
    // degenerate lanes: hand them to the scalar routine (rare: rank <= 1, all-zero, non-finite)
#pragma unroll
    for (int k = 0; k < LaneCount<T>::value; ++k) {
        float A0 = lane_get(al0, k), A1 = lane_get(al1, k), N0 = lane_get(n0, k), N1 = lane_get(n1, k);
        if (__builtin_expect(!(A1 > 1e-24f * A0) || !(A0 > 0.f) || !(N0 > 0.f) || !(N1 > 1e-30f), 0)) {
            float ms[9], rs[9], ds;
            fetch(a);
#pragma unroll
            for (int i = 0; i < 9; ++i) ms[i] = lane_get(a[i], k);
            svd_orth_degenerate_lane(ms, rs, ds);
#pragma unroll
            for (int i = 0; i < 9; ++i) lane_put(r[i], k, rs[i]);
            lane_put(dsign, k, ds);
            if (ctx) ctx->deg |= 1 << k;
        }
    }
}

template <class T>
PE_DEV void svd_orth_t(const T* m, T* r, T& dsign) {
    svd_orth_f<T>([m](T* dst) {
#pragma unroll
        for (int i = 0; i < 9; ++i) dst[i] = m[i];
    }, r, dsign);
}

// Pullback of R = svd_orth(M), lane-generic (formulas and guards: svd_orth_pullback in pe_math.cuh).
// H = tr(P) I - P for P = sym(R^T M), scaled by the power-of-two prescale `sc`, and the solve y = H^-1 z, dM = R [y]_x.
// Shared by the generic pullback and by the fused Frobenius step (whose z comes from R^T R_true directly).
template <class T>
PE_DEV void svd_orth_pullback_solve(const T* m, const T* r, T z0, T z1, T z2, T* dm) {
    const T z = m[0];
    const T sc = t_prescale(m);
    T n[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            n[3 * i + j] = t_fma(r[6 + i], m[6 + j], t_fma(r[3 + i], m[3 + j], t_mul(r[i], m[j])));
    const T hs = t_mul(t_set(z, -0.5f), sc);
    T h01 = t_mul(hs, t_add(n[1], n[3])), h02 = t_mul(hs, t_add(n[2], n[6])), h12 = t_mul(hs, t_add(n[5], n[7]));
    T h00 = t_mul(sc, t_add(n[4], n[8])), h11 = t_mul(sc, t_add(n[0], n[8])), h22 = t_mul(sc, t_add(n[0], n[4]));
    T c00 = t_fma(h11, h22, t_neg(t_mul(h12, h12))), c01 = t_fma(h02, h12, t_neg(t_mul(h01, h22))),
      c02 = t_fma(h01, h12, t_neg(t_mul(h02, h11)));
    T c11 = t_fma(h00, h22, t_neg(t_mul(h02, h02))), c12 = t_fma(h01, h02, t_neg(t_mul(h00, h12))),
      c22 = t_fma(h00, h11, t_neg(t_mul(h01, h01)));
    T det = t_fma(h00, c00, t_fma(h01, c01, t_mul(h02, c02)));
    T tr3 = t_mul(t_add(t_add(h00, h11), h22), t_set(z, 1.0f / 3.0f));
    T fl = t_mul(t_set(z, kDetFloor), t_abs(t_mul(t_mul(tr3, tr3), tr3)));
    det = t_max(det, fl);
    T inv = z;
#pragma unroll
    for (int k = 0; k < LaneCount<T>::value; ++k) {
        float d = lane_get(det, k);
        lane_put(inv, k, d > 1e-30f ? lane_get(sc, k) * rcp_fast(d) : 0.f);
    }
    T y0 = t_mul(inv, t_fma(c00, z0, t_fma(c01, z1, t_mul(c02, z2))));
    T y1 = t_mul(inv, t_fma(c01, z0, t_fma(c11, z1, t_mul(c12, z2))));
    T y2 = t_mul(inv, t_fma(c02, z0, t_fma(c12, z1, t_mul(c22, z2))));
    T ny0 = t_neg(y0), ny1 = t_neg(y1), ny2 = t_neg(y2);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        T r0 = r[3 * i], r1 = r[3 * i + 1], r2 = r[3 * i + 2];
        dm[3 * i + 0] = t_fma(r1, y2, t_mul(r2, ny1));
        dm[3 * i + 1] = t_fma(r2, y0, t_mul(r0, ny2));
        dm[3 * i + 2] = t_fma(r0, y1, t_mul(r1, ny0));
    }
}

// The same solve from the decomposition itself (SvdCtx), for kernels that have just computed it: with
// P = R^T M = V diag(s0, s1, s2) V^T the matrix tr(P) I - P is V diag(a, b, c) V^T, a = s1 + s2, b = s0 + s2, c = s0 + s1,
// so y = V diag(bc, ac, ab) V^T z / max(abc, floor) - the adjugate form, i.e. exactly the regularisation of the
// cofactor solve above - built once, next to the decomposition, as the symmetric A (6 words); the pullback is then
// y = A z and dM = R [y]_x: 61 packed instructions in all instead of 89, and M is not read again.
// Lanes flagged in ctx.deg (scalar route, no valid V) are redone by the generic solve from M.
template <class T, class Fetch>
PE_DEV void svd_orth_pullback_ctx(const SvdCtx<T>& c, const Fetch& fetch, const T* r, T z0, T z1, T z2, T* dm) {
    T y0 = t_fma(c.a[2], z2, t_fma(c.a[1], z1, t_mul(c.a[0], z0)));
    T y1 = t_fma(c.a[4], z2, t_fma(c.a[3], z1, t_mul(c.a[1], z0)));
    T y2 = t_fma(c.a[5], z2, t_fma(c.a[4], z1, t_mul(c.a[2], z0)));
    T ny0 = t_neg(y0), ny1 = t_neg(y1), ny2 = t_neg(y2);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        T r0 = r[3 * i], r1 = r[3 * i + 1], r2 = r[3 * i + 2];
        dm[3 * i + 0] = t_fma(r1, y2, t_mul(r2, ny1));
        dm[3 * i + 1] = t_fma(r2, y0, t_mul(r0, ny2));
        dm[3 * i + 2] = t_fma(r0, y1, t_mul(r1, ny0));
    }
    if (__builtin_expect(c.deg != 0, 0)) {
        T m[9];
        fetch(m);
#pragma unroll
        for (int k = 0; k < LaneCount<T>::value; ++k) {
            if (c.deg & (1 << k)) {
                float ms[9], rs[9], ds[9];
#pragma unroll
                for (int i = 0; i < 9; ++i) { ms[i] = lane_get(m[i], k); rs[i] = lane_get(r[i], k); }
                svd_orth_pullback_solve<float>(ms, rs, lane_get(z0, k), lane_get(z1, k), lane_get(z2, k), ds);
#pragma unroll
                for (int i = 0; i < 9; ++i) lane_put(dm[i], k, ds[i]);
            }
        }
    }
}

// axial vector of T - T^T for T = R^T G (only the six off-diagonal entries of T are formed)
template <class T>
PE_DEV void skew_axial_rtg(const T* r, const T* g, T& z0, T& z1, T& z2) {
    T t01 = t_fma(r[6], g[7], t_fma(r[3], g[4], t_mul(r[0], g[1])));
    T t02 = t_fma(r[6], g[8], t_fma(r[3], g[5], t_mul(r[0], g[2])));
    T t10 = t_fma(r[7], g[6], t_fma(r[4], g[3], t_mul(r[1], g[0])));
    T t12 = t_fma(r[7], g[8], t_fma(r[4], g[5], t_mul(r[1], g[2])));
    T t20 = t_fma(r[8], g[6], t_fma(r[5], g[3], t_mul(r[2], g[0])));
    T t21 = t_fma(r[8], g[7], t_fma(r[5], g[4], t_mul(r[2], g[1])));
    z0 = t_sub(t21, t12); z1 = t_sub(t02, t20); z2 = t_sub(t10, t01);
}

// Pullback of R = svd_orth(M), lane-generic (formulas and guards: svd_orth_pullback in pe_math.cuh).
template <class T>
PE_DEV void svd_orth_pullback_t(const T* m, const T* r, const T* g, T* dm) {
    T z0, z1, z2;
    skew_axial_rtg(r, g, z0, z1, z2);
    svd_orth_pullback_solve(m, r, z0, z1, z2, dm);
}

// ---- angle_error for the fused sums, lane-generic ---------------------------------------------------------------------
// tr(R^T R_t) - 1 (= 2 cos) as an UNEVALUATED float32 pair ahi + lo, error-free to ~1e-11: every product is added to a
// big accumulator (96 = 1.5 * 2^6, one ulp = 2^-17) by ONE fused multiply-add, whose rounding error is recovered
// exactly by two more (acc - acc' is exact, r m - (acc' - acc) is the error, <= 2^-18) and summed on the side.  Four
// packed instructions per product and no conversion, against 2 conversions + 1 DFMA per product and lane in float64.
// acc stays inside [64, 128) for |trace| <= 32, i.e. for anything that is remotely a pair of rotations.
template <class T>
PE_DEV void trace_minus_one_split(const T* r, const T* m, T& ahi, T& lo) {
    const T z = r[0];
    const T minus1 = t_set(z, -1.0f);
    T acc = t_set(z, 96.0f);
    lo = t_set(z, 0.0f);
#pragma unroll
    for (int i = 0; i < 9; ++i) {
        const T an = t_fma(r[i], m[i], acc);
        const T dl = t_fma(an, minus1, acc);
        lo = t_add(lo, t_fma(r[i], m[i], dl));
        acc = an;
    }
    ahi = t_add(acc, t_set(z, -97.0f));
}
// one matrix pair per thread (the one-record-per-thread kernels): the nine products ride in the two halves of packed
// registers (elements 0-4 and 5-8), two accumulators, merged exactly at the end
PE_DEV void trace_minus_one_split_scalar(const float* r, const float* m, float& ahi, float& lo) {
    const F2 minus1{-1.0f, -1.0f};
    F2 acc{96.0f, 96.0f}, l{0.0f, 0.0f};
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        const F2 a{r[i], i + 5 < 9 ? r[i + 5] : 0.0f}, b{m[i], i + 5 < 9 ? m[i + 5] : 0.0f};
        const F2 an = f2_fma(a, b, acc);
        const F2 dl = f2_fma(an, minus1, acc);
        l = t_add(l, f2_fma(a, b, dl));
        acc = an;
    }
    ahi = (acc.x - 96.0f) + (acc.y - 97.0f);  // both on the 2^-17 grid, |sum| < 16: exact
    lo = l.x + l.y;
}
// t = (1 - |cos|)/2 from that pair (exact up to the final rounding) and the sign of the cosine
template <class T, class M>
PE_DEV T half_versine_from_split(T ahi, T lo, M& neg, T& a32) {
    const T z = ahi;
    a32 = t_add(ahi, lo);
    neg = t_ltz(a32);
    const T sg = t_sel(neg, t_set(z, 0.25f), t_set(z, -0.25f));  // -sign(cos) / 4
    return t_fma(sg, lo, t_fma(t_set(z, -0.25f), t_abs(ahi), t_set(z, 0.5f)));
}
// degrees from t = sin^2(angle/2 or (180 - angle)/2): angle_deg_tail32 of pe_math.cuh for one or two lanes
template <class T, class M>
PE_DEV T angle_deg_tail32_t(T t, M neg) {
    const T z = t;
    t = t_sel(t_ltz(t), t_set(z, 0.0f), t);
    const T h = t_sqrt_fast(t);
    const T u = t_mul(h, h);
    T q = t_set(z, kAsinC6);
    q = t_fma(q, u, t_set(z, kAsinC5));
    q = t_fma(q, u, t_set(z, kAsinC4));
    q = t_fma(q, u, t_set(z, kAsinC3));
    q = t_fma(q, u, t_set(z, kAsinC2));
    q = t_fma(q, u, t_set(z, kAsinC1));
    q = t_fma(q, u, t_set(z, kAsinC0));
    const T psi = t_fma(t_mul(h, u), q, h);
    return t_fma(psi, t_sel(neg, t_set(z, -kTwoRadToDeg), t_set(z, kTwoRadToDeg)), t_sel(neg, t_set(z, 180.0f), t_set(z, 0.0f)));
}

// d = a - b and |d|_F^2, lane-generic
template <class T>
PE_DEV T diff9_sq(const T* a, const T* b, T* d) {
    d[0] = t_sub(a[0], b[0]);
    T s = t_mul(d[0], d[0]);
#pragma unroll
    for (int i = 1; i < 9; ++i) { d[i] = t_sub(a[i], b[i]); s = t_fma(d[i], d[i], s); }
    return s;
}

}  // namespace pe
