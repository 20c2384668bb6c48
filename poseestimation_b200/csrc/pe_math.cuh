// Register-resident per-record math for the rotation-mapping hot path (sm_100a).
//
// Every function here works on ONE record held in registers by ONE thread: a 3x3 matrix is nine floats in
// row-major order (m[3*i+j]), exactly the layout `x.view(-1, 3, 3)` gives the reference
// (/root/reference/rotation_representation.py:199).  No function touches memory; the staging kernels in
// pe_stage.cuh move records between HBM, shared memory and registers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>
#include <math.h>

namespace pe {

// __host__ as well, so tests/hostmath can run the very same formulas on the CPU against the oracle (test
// infrastructure only: nothing in the package calls the host instantiation).
#define PE_DEV __host__ __device__ __forceinline__

PE_DEV float rsqrt_fast(float x) {  // MUFU.RSQ, ~2^-22 relative error, flush-to-zero
#ifdef __CUDA_ARCH__
    float y;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
#else
    return 1.0f / sqrtf(x);
#endif
}
PE_DEV float sqrt_fast(float x) {  // MUFU.SQRT, ~2^-22 relative error, flush-to-zero
#ifdef __CUDA_ARCH__
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
#else
    return sqrtf(x);
#endif
}
PE_DEV float rcp_fast(float x) {  // MUFU.RCP, ~1 ulp, flush-to-zero
#ifdef __CUDA_ARCH__
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
#else
    return 1.0f / x;
#endif
}
PE_DEV int f2i(float x) {
#ifdef __CUDA_ARCH__
    return __float_as_int(x);
#else
    int i; memcpy(&i, &x, 4); return i;
#endif
}
PE_DEV float i2f(int i) {
#ifdef __CUDA_ARCH__
    return __int_as_float(i);
#else
    float x; memcpy(&x, &i, 4); return x;
#endif
}
PE_DEV void sincos_f(float a, float* s, float* c) {
#ifdef __CUDA_ARCH__
    sincosf(a, s, c);
#else
    *s = sinf(a); *c = cosf(a);
#endif
}
// rsqrt with one Newton step: full fp32 accuracy, used where a single rounding matters (elementwise maps).
PE_DEV float rsqrt_nr(float x) {
    float y = rsqrt_fast(x);
    float t = x * y;
    return y * fmaf(-0.5f * t, y, 1.5f);
}

// 2^-e for e = exponent of the largest |entry| : an exact power-of-two prescale, so the nine mantissas are
// untouched and Gram entries stay inside [0, 12].  Exponent is clamped so zero/denormal and huge inputs do
// not produce a zero or non-finite scale.
PE_DEV float pow2_prescale(const float* m) {
    float mx = fmaxf(fmaxf(fmaxf(fabsf(m[0]), fabsf(m[1])), fmaxf(fabsf(m[2]), fabsf(m[3]))),
                     fmaxf(fmaxf(fabsf(m[4]), fabsf(m[5])), fmaxf(fabsf(m[6]), fmaxf(fabsf(m[7]), fabsf(m[8])))));
    int e = (f2i(mx) >> 23) & 0xff;
    e = e < 1 ? 1 : (e > 253 ? 253 : e);
    return i2f((254 - e) << 23);
}

// One Hestenes rotation of columns (p, q) of A (3 rows) and V (3 rows).  The angle is the exact Jacobi angle
// (two MUFU.RSQ, no division); when column q is the longer one a quarter turn is composed in, so that after the
// rotation |a_p| >= |a_q| ("sorting rotation").  After the last (1,2) rotation of a sweep column 2 is the
// shortest, which is what the cross-product completion in svd_orth() relies on.
PE_DEV void hestenes(float& ap0, float& ap1, float& ap2, float& aq0, float& aq1, float& aq2,
                     float& vp0, float& vp1, float& vp2, float& vq0, float& vq1, float& vq2) {
    float al = fmaf(ap2, ap2, fmaf(ap1, ap1, ap0 * ap0));
    float be = fmaf(aq2, aq2, fmaf(aq1, aq1, aq0 * aq0));
    float ga = fmaf(ap2, aq2, fmaf(ap1, aq1, ap0 * aq0));
    float dh = 0.5f * (al - be);
    float ad = fmaxf(fabsf(dh), 1e-18f);
    float r = rsqrt_fast(fmaf(ad, ad, ga * ga));
    float c2 = ad * r;                   // cos 2θ  in [0, 1]
    float s2 = ga * r;                   // sin 2θ  (sign fixed below)
    float w = fmaf(2.0f, c2, 2.0f);      // 2 + 2 cos 2θ  in [2, 4]
    float h = rsqrt_fast(w);
    float c = 0.5f * w * h;              // cos θ
    float s = s2 * h;                    // sin θ  for dh > 0
    // dh < 0: the inner angle changes sign, then a quarter turn swaps the roles of p and q.
    bool neg = dh < 0.0f;
    float cp = neg ? s : c;              // (-(-s)) = s
    float sp = neg ? c : s;
    // p' = cp*p + sp*q ; q' = cp*q - sp*p
    float t;
    t = ap0; ap0 = fmaf(cp, t, sp * aq0); aq0 = fmaf(cp, aq0, -sp * t);
    t = ap1; ap1 = fmaf(cp, t, sp * aq1); aq1 = fmaf(cp, aq1, -sp * t);
    t = ap2; ap2 = fmaf(cp, t, sp * aq2); aq2 = fmaf(cp, aq2, -sp * t);
    t = vp0; vp0 = fmaf(cp, t, sp * vq0); vq0 = fmaf(cp, vq0, -sp * t);
    t = vp1; vp1 = fmaf(cp, t, sp * vq1); vq1 = fmaf(cp, vq1, -sp * t);
    t = vp2; vp2 = fmaf(cp, t, sp * vq2); vq2 = fmaf(cp, vq2, -sp * t);
}

// Nearest rotation to M in the Frobenius norm:  R = U diag(1, 1, det(U V^T)) V^T  with  M = U S V^T
// (reference: symmetric_orthogonalization, /root/reference/rotation_representation.py:192-206).
//
// Method (not the reference's LAPACK path): one-sided Jacobi on the columns of A = M V, V accumulated from
// proper rotations only (det V = +1).  At convergence A = U S.  With u0, u1 the two longest normalised columns,
//     R = u0 v0^T + u1 v1^T + (u0 x u1) v2^T
// is the det-fixed answer for either sign of det(M): the true u2 is +-(u0 x u1) and the sign is exactly the
// reference's det(U V^T).  The smallest singular direction is never normalised, so rank-2 inputs are exact,
// and nothing is squared (no M^T M), so the error is at the conditioning limit eps*s1/(s2 + d*s3).
//
// `dsign` receives det(U V^T) = sign(det M) as +-1 (what the reference multiplies the last row of V^T by).
template <int SWEEPS>
PE_DEV void svd_orth(const float* m, float* r, float& dsign) {
    const float sc = pow2_prescale(m);
    float a00 = m[0] * sc, a01 = m[1] * sc, a02 = m[2] * sc;
    float a10 = m[3] * sc, a11 = m[4] * sc, a12 = m[5] * sc;
    float a20 = m[6] * sc, a21 = m[7] * sc, a22 = m[8] * sc;
    float v00 = 1.f, v01 = 0.f, v02 = 0.f, v10 = 0.f, v11 = 1.f, v12 = 0.f, v20 = 0.f, v21 = 0.f, v22 = 1.f;
#pragma unroll
    for (int sw = 0; sw < SWEEPS; ++sw) {
        hestenes(a00, a10, a20, a01, a11, a21, v00, v10, v20, v01, v11, v21);
        hestenes(a00, a10, a20, a02, a12, a22, v00, v10, v20, v02, v12, v22);
        hestenes(a01, a11, a21, a02, a12, a22, v01, v11, v21, v02, v12, v22);
    }
    float al0 = fmaf(a20, a20, fmaf(a10, a10, a00 * a00));
    float al1 = fmaf(a21, a21, fmaf(a11, a11, a01 * a01));
    float k0 = rsqrt_fast(al0), k1 = rsqrt_fast(al1);
    float u00 = a00 * k0, u10 = a10 * k0, u20 = a20 * k0;
    float u01 = a01 * k1, u11 = a11 * k1, u21 = a21 * k1;
    if (__builtin_expect(!(al1 > 1e-24f * al0) || !(al0 > 0.f), 0)) {
        // rank <= 1 (or the zero matrix): the reference returns an arbitrary completion; pick a finite one.
        if (!(al0 > 0.f)) {
            if (al0 == 0.f) { u00 = 1.f; u10 = 0.f; u20 = 0.f; }  // zero matrix -> identity (V = I too)
        }
        if (al0 == al0) {  // not NaN
            float ax = fabsf(u00), ay = fabsf(u10), az = fabsf(u20);
            float e0 = (ax <= ay && ax <= az) ? 1.f : 0.f;
            float e1 = (e0 == 0.f && ay <= az) ? 1.f : 0.f;
            float e2 = (e0 == 0.f && e1 == 0.f) ? 1.f : 0.f;
            // u1 = normalise(u0 x e) ... any unit vector orthogonal to u0
            float c0 = u10 * e2 - u20 * e1, c1 = u20 * e0 - u00 * e2, c2 = u00 * e1 - u10 * e0;
            float kk = rsqrt_nr(c0 * c0 + c1 * c1 + c2 * c2);
            u01 = c0 * kk; u11 = c1 * kk; u21 = c2 * kk;
        }
    }
    // third left vector by cross product (this is where det(U V^T) is applied)
    float u02 = fmaf(u10, u21, -u20 * u11);
    float u12 = fmaf(u20, u01, -u00 * u21);
    float u22 = fmaf(u00, u11, -u10 * u01);
    dsign = (fmaf(a02, u02, fmaf(a12, u12, a22 * u22)) < 0.f) ? -1.f : 1.f;
    // V's columns carry the accumulated (1 + O(2^-22)) scale of the approximate rsqrt's: normalise once.
    float n0 = rsqrt_fast(fmaf(v20, v20, fmaf(v10, v10, v00 * v00)));
    float n1 = rsqrt_fast(fmaf(v21, v21, fmaf(v11, v11, v01 * v01)));
    float n2 = rsqrt_fast(fmaf(v22, v22, fmaf(v12, v12, v02 * v02)));
    v00 *= n0; v10 *= n0; v20 *= n0;
    v01 *= n1; v11 *= n1; v21 *= n1;
    v02 *= n2; v12 *= n2; v22 *= n2;
    // R_ik = sum_j u_ij v_kj
    r[0] = fmaf(u02, v02, fmaf(u01, v01, u00 * v00));
    r[1] = fmaf(u02, v12, fmaf(u01, v11, u00 * v10));
    r[2] = fmaf(u02, v22, fmaf(u01, v21, u00 * v20));
    r[3] = fmaf(u12, v02, fmaf(u11, v01, u10 * v00));
    r[4] = fmaf(u12, v12, fmaf(u11, v11, u10 * v10));
    r[5] = fmaf(u12, v22, fmaf(u11, v21, u10 * v20));
    r[6] = fmaf(u22, v02, fmaf(u21, v01, u20 * v00));
    r[7] = fmaf(u22, v12, fmaf(u21, v11, u20 * v10));
    r[8] = fmaf(u22, v22, fmaf(u21, v21, u20 * v20));
}

constexpr int kSweeps = 4;  // 4 sweeps reach the fp32 conditioning limit on Gaussian inputs

// ---------------------------------------------------------------------------------------------------------
// Production variant of svd_orth (same result, ~35 % fewer issue slots):
//  * V is NOT accumulated.  At convergence A = M V = U S, so the right vectors follow from the left ones:
//    v_j ~ M^T u_j.  v0 = normalise(M^T u0); v1 = normalise(M^T u1 - (.,v0) v0) (the Gram-Schmidt step removes
//    exactly the eps*s0/s1-amplified part of u1's error); v2 = v0 x v1.  R = u0 v0^T + u1 v1^T + (u0 x u1)(v0 x v1)^T.
//  * 10 rotations in the order (0,1)(0,2)(1,2)(0,2)(0,1)(1,2)(0,2)(0,1)(0,2)(1,2) instead of 4 cyclic sweeps (12):
//    found by exhaustive search over pair sequences; same err*cond as 12 on 2^19 Gaussian matrices.
//  * rows 0 and 1 of every column live in one 64-bit register pair and are rotated with packed FMUL2/FFMA2
//    (sm_100 fp32x2): half the issue slots for the rotation application.
struct F2 {
    float x, y;
};
PE_DEV F2 f2_fma(F2 a, F2 b, F2 c) {
#ifdef __CUDA_ARCH__
    float2 r = __ffma2_rn(make_float2(a.x, a.y), make_float2(b.x, b.y), make_float2(c.x, c.y));
    return F2{r.x, r.y};
#else
    return F2{fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)};
#endif
}
PE_DEV F2 f2_mul(F2 a, F2 b) {
#ifdef __CUDA_ARCH__
    float2 r = __fmul2_rn(make_float2(a.x, a.y), make_float2(b.x, b.y));
    return F2{r.x, r.y};
#else
    return F2{a.x * b.x, a.y * b.y};
#endif
}
// one sorting Hestenes rotation of columns p, q held as (rows 0-1 packed, row 2 scalar).
// np, nq carry the squared column norms from rotation to rotation: the rotated pair's new norms are the eigenvalues
// mean +- rho of its 2x2 Gram matrix, which the angle computation has already formed, so a TRACKED rotation needs one
// dot product (3 FMA-pipe instructions) instead of three (9).  FRESH recomputes both norms from the columns; the last
// kFreshTail rotations are fresh, so rounding drift of the tracked values cannot reach the result.
#ifndef PE_FRESH_TAIL
#define PE_FRESH_TAIL 1
#endif
constexpr int kFreshTail = PE_FRESH_TAIL;
template <bool FRESH>
PE_DEV void hestenes_a(F2& p01, float& p2, F2& q01, float& q2, float& np, float& nq) {
    float al = FRESH ? fmaf(p2, p2, fmaf(p01.y, p01.y, p01.x * p01.x)) : np;
    float be = FRESH ? fmaf(q2, q2, fmaf(q01.y, q01.y, q01.x * q01.x)) : nq;
    float ga = fmaf(p2, q2, fmaf(p01.y, q01.y, p01.x * q01.x));
    float dh = 0.5f * (al - be);
    float ad = fmaxf(fabsf(dh), 1e-18f);
    // tan(theta) = ga / (rho + |dh|), rho = sqrt(dh^2 + ga^2); (c, s) = (u, ga) / sqrt(u^2 + ga^2)  [pe_lanes.cuh]
    float g2 = ga * ga;
    float rho = sqrt_fast(fmaf(ad, ad, g2));
    float u = rho + ad;
    float k = rsqrt_fast(fmaf(u, u, g2));
    float c = u * k, s = ga * k;
    float mean = al - dh;
    np = mean + rho; nq = mean - rho;
    bool neg = dh < 0.0f;
    float cp = neg ? s : c, sp = neg ? c : s;
    F2 C{cp, cp}, S{sp, sp}, N{-sp, -sp};
    F2 pn = f2_fma(C, p01, f2_mul(S, q01));
    F2 qn = f2_fma(C, q01, f2_mul(N, p01));
    float t = p2;
    p2 = fmaf(cp, t, sp * q2);
    q2 = fmaf(cp, q2, -sp * t);
    p01 = pn; q01 = qn;
}
// a unit vector orthogonal to the unit vector u (any one; used only for rank-deficient inputs)
PE_DEV void any_orthogonal(float u0, float u1, float u2, float& o0, float& o1, float& o2) {
    float ax = fabsf(u0), ay = fabsf(u1), az = fabsf(u2);
    float e0 = (ax <= ay && ax <= az) ? 1.f : 0.f;
    float e1 = (e0 == 0.f && ay <= az) ? 1.f : 0.f;
    float e2 = (e0 == 0.f && e1 == 0.f) ? 1.f : 0.f;
    float c0 = u1 * e2 - u2 * e1, c1 = u2 * e0 - u0 * e2, c2 = u0 * e1 - u1 * e0;
    float kk = rsqrt_nr(c0 * c0 + c1 * c1 + c2 * c2);
    o0 = c0 * kk; o1 = c1 * kk; o2 = c2 * kk;
}
PE_DEV void svd_orth_fast(const float* m, float* r, float& dsign) {
    const float sc = pow2_prescale(m);
    // column j: rows 0,1 packed, row 2 scalar
    F2 c0{m[0] * sc, m[3] * sc}, c1{m[1] * sc, m[4] * sc}, c2{m[2] * sc, m[5] * sc};
    float z0 = m[6] * sc, z1 = m[7] * sc, z2 = m[8] * sc;
    float q0n = fmaf(z0, z0, fmaf(c0.y, c0.y, c0.x * c0.x));
    float q1n = fmaf(z1, z1, fmaf(c1.y, c1.y, c1.x * c1.x));
    float q2n = fmaf(z2, z2, fmaf(c2.y, c2.y, c2.x * c2.x));
    hestenes_a<false>(c0, z0, c1, z1, q0n, q1n);
    hestenes_a<false>(c0, z0, c2, z2, q0n, q2n);
    hestenes_a<false>(c1, z1, c2, z2, q1n, q2n);
    hestenes_a<false>(c0, z0, c2, z2, q0n, q2n);
    hestenes_a<false>(c0, z0, c1, z1, q0n, q1n);
    hestenes_a<false>(c1, z1, c2, z2, q1n, q2n);
    hestenes_a<false>(c0, z0, c2, z2, q0n, q2n);
    hestenes_a<(kFreshTail >= 3)>(c0, z0, c1, z1, q0n, q1n);
    hestenes_a<(kFreshTail >= 2)>(c0, z0, c2, z2, q0n, q2n);
    hestenes_a<(kFreshTail >= 1)>(c1, z1, c2, z2, q1n, q2n);
    float al0 = fmaf(z0, z0, fmaf(c0.y, c0.y, c0.x * c0.x));
    float al1 = fmaf(z1, z1, fmaf(c1.y, c1.y, c1.x * c1.x));
    float k0 = rsqrt_fast(al0), k1 = rsqrt_fast(al1);
    float u00 = c0.x * k0, u10 = c0.y * k0, u20 = z0 * k0;
    float u01 = c1.x * k1, u11 = c1.y * k1, u21 = z1 * k1;
    if (__builtin_expect(!(al1 > 1e-24f * al0) || !(al0 > 0.f), 0)) {  // rank <= 1, or the zero matrix
        if (al0 == 0.f) { u00 = 1.f; u10 = 0.f; u20 = 0.f; }
        if (al0 == al0) any_orthogonal(u00, u10, u20, u01, u11, u21);
    }
    float u02 = fmaf(u10, u21, -u20 * u11);
    float u12 = fmaf(u20, u01, -u00 * u21);
    float u22 = fmaf(u00, u11, -u10 * u01);
    dsign = (fmaf(c2.x, u02, fmaf(c2.y, u12, z2 * u22)) < 0.f) ? -1.f : 1.f;
    // right vectors from the left ones: w_j = (sc M)^T u_j
    float m0 = m[0] * sc, m1 = m[1] * sc, m2 = m[2] * sc, m3 = m[3] * sc, m4 = m[4] * sc, m5 = m[5] * sc,
          m6 = m[6] * sc, m7 = m[7] * sc, m8 = m[8] * sc;
    float w00 = fmaf(m6, u20, fmaf(m3, u10, m0 * u00));
    float w10 = fmaf(m7, u20, fmaf(m4, u10, m1 * u00));
    float w20 = fmaf(m8, u20, fmaf(m5, u10, m2 * u00));
    float w01 = fmaf(m6, u21, fmaf(m3, u11, m0 * u01));
    float w11 = fmaf(m7, u21, fmaf(m4, u11, m1 * u01));
    float w21 = fmaf(m8, u21, fmaf(m5, u11, m2 * u01));
    float n0 = fmaf(w20, w20, fmaf(w10, w10, w00 * w00));
    float j0 = rsqrt_fast(n0);
    float v00 = w00 * j0, v10 = w10 * j0, v20 = w20 * j0;
    if (__builtin_expect(!(n0 > 0.f), 0)) {
        if (n0 == 0.f) { v00 = 1.f; v10 = 0.f; v20 = 0.f; }
    }
    float pr = fmaf(w21, v20, fmaf(w11, v10, w01 * v00));
    w01 = fmaf(-pr, v00, w01); w11 = fmaf(-pr, v10, w11); w21 = fmaf(-pr, v20, w21);
    float n1 = fmaf(w21, w21, fmaf(w11, w11, w01 * w01));
    float j1 = rsqrt_fast(n1);
    float v01 = w01 * j1, v11 = w11 * j1, v21 = w21 * j1;
    if (__builtin_expect(!(n1 > 1e-30f), 0)) {
        if (n1 == n1) any_orthogonal(v00, v10, v20, v01, v11, v21);
    }
    float v02 = fmaf(v10, v21, -v20 * v11);
    float v12 = fmaf(v20, v01, -v00 * v21);
    float v22 = fmaf(v00, v11, -v10 * v01);
    r[0] = fmaf(u02, v02, fmaf(u01, v01, u00 * v00));
    r[1] = fmaf(u02, v12, fmaf(u01, v11, u00 * v10));
    r[2] = fmaf(u02, v22, fmaf(u01, v21, u00 * v20));
    r[3] = fmaf(u12, v02, fmaf(u11, v01, u10 * v00));
    r[4] = fmaf(u12, v12, fmaf(u11, v11, u10 * v10));
    r[5] = fmaf(u12, v22, fmaf(u11, v21, u10 * v20));
    r[6] = fmaf(u22, v02, fmaf(u21, v01, u20 * v00));
    r[7] = fmaf(u22, v12, fmaf(u21, v11, u20 * v10));
    r[8] = fmaf(u22, v22, fmaf(u21, v21, u20 * v20));
}

// Pullback of R = svd_orth(M):  given M, R and G = dL/dR, return dL/dM.
//
// Polar form of the closed formula in SURVEY.md 7.4 (B = D U^T G V, W_ij = (B_ij - B_ji)/(s'_i + s'_j),
// dM = U D W V^T): with P = R^T M = V diag(s') V^T (symmetric) and Z = R^T G - G^T R (skew, axial vector z),
// the skew Y = V W V^T solves  Y P + P Y = Z, which for 3x3 is the symmetric linear system
//     (tr(P) I - P) y = z,       dM = R [y]_x .
// The eigenvalues of tr(P) I - P are exactly the s'_i + s'_j, so the only degenerate direction is
// s2 + d*s3 -> 0.  Guard: det is floored at kDetFloor * (tr/3)^3, which bounds the gradient.
// This replaces autograd through torch.svd / det / cat / matmul (rotation_representation.py:200-205).
constexpr float kDetFloor = 1e-7f;
PE_DEV void svd_orth_pullback(const float* m, const float* r, const float* g, float* dm) {
    const float sc = pow2_prescale(m);
    // N = R^T M (unscaled); the power-of-two scale goes onto the six entries of H instead of the nine of N
    float n[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            n[3 * i + j] = fmaf(r[6 + i], m[6 + j], fmaf(r[3 + i], m[3 + j], r[i] * m[j]));
    const float hs = -0.5f * sc;
    float h01 = hs * (n[1] + n[3]), h02 = hs * (n[2] + n[6]), h12 = hs * (n[5] + n[7]);
    float h00 = sc * (n[4] + n[8]), h11 = sc * (n[0] + n[8]), h22 = sc * (n[0] + n[4]);
    // T = R^T G ; z = axial(T - T^T)
    float t01 = fmaf(r[6], g[7], fmaf(r[3], g[4], r[0] * g[1]));
    float t02 = fmaf(r[6], g[8], fmaf(r[3], g[5], r[0] * g[2]));
    float t10 = fmaf(r[7], g[6], fmaf(r[4], g[3], r[1] * g[0]));
    float t12 = fmaf(r[7], g[8], fmaf(r[4], g[5], r[1] * g[2]));
    float t20 = fmaf(r[8], g[6], fmaf(r[5], g[3], r[2] * g[0]));
    float t21 = fmaf(r[8], g[7], fmaf(r[5], g[4], r[2] * g[1]));
    float z0 = t21 - t12, z1 = t02 - t20, z2 = t10 - t01;
    // y = adj(H) z / det(H)
    float c00 = fmaf(h11, h22, -h12 * h12), c01 = fmaf(h02, h12, -h01 * h22), c02 = fmaf(h01, h12, -h02 * h11);
    float c11 = fmaf(h00, h22, -h02 * h02), c12 = fmaf(h01, h02, -h00 * h12), c22 = fmaf(h00, h11, -h01 * h01);
    float det = fmaf(h00, c00, fmaf(h01, c01, h02 * c02));
    float tr3 = (h00 + h11 + h22) * (1.0f / 3.0f);
    float fl = kDetFloor * fabsf(tr3 * tr3 * tr3);
    det = fmaxf(det, fl);
    float inv = det > 1e-30f ? sc * rcp_fast(det) : 0.f;  // det = 0 only for the zero matrix: define its gradient as 0
    float y0 = inv * fmaf(c00, z0, fmaf(c01, z1, c02 * z2));
    float y1 = inv * fmaf(c01, z0, fmaf(c11, z1, c12 * z2));
    float y2 = inv * fmaf(c02, z0, fmaf(c12, z1, c22 * z2));
    // dM = R [y]_x
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float r0 = r[3 * i], r1 = r[3 * i + 1], r2 = r[3 * i + 2];
        dm[3 * i + 0] = fmaf(r1, y2, -r2 * y1);
        dm[3 * i + 1] = fmaf(r2, y0, -r0 * y2);
        dm[3 * i + 2] = fmaf(r0, y1, -r1 * y0);
    }
}

// ---------------------------------------------------------------------------------------------------------
// Gram-Schmidt 6D map (reference: compute_rotation_matrix_from_ortho6d, rotation_representation.py:21-36).
// x = a/|a| ; z = (x X b)/|x X b| ; y = z X x ; the three vectors are the COLUMNS of R.  No epsilon: a zero or
// collinear input gives NaN, as in the reference.
PE_DEV void ortho6d(const float* p, float* r) {
    float ka = rsqrt_nr(fmaf(p[2], p[2], fmaf(p[1], p[1], p[0] * p[0])));
    float x0 = p[0] * ka, x1 = p[1] * ka, x2 = p[2] * ka;
    float z0 = fmaf(x1, p[5], -x2 * p[4]), z1 = fmaf(x2, p[3], -x0 * p[5]), z2 = fmaf(x0, p[4], -x1 * p[3]);
    float kz = rsqrt_nr(fmaf(z2, z2, fmaf(z1, z1, z0 * z0)));
    z0 *= kz; z1 *= kz; z2 *= kz;
    float y0 = fmaf(z1, x2, -z2 * x1), y1 = fmaf(z2, x0, -z0 * x2), y2 = fmaf(z0, x1, -z1 * x0);
    r[0] = x0; r[1] = y0; r[2] = z0;
    r[3] = x1; r[4] = y1; r[5] = z1;
    r[6] = x2; r[7] = y2; r[8] = z2;
}

// Pullback of ortho6d: recomputes the forward from p (24 B) instead of saving intermediates.
PE_DEV void ortho6d_pullback(const float* p, const float* g, float* dp) {
    float na2 = fmaf(p[2], p[2], fmaf(p[1], p[1], p[0] * p[0]));
    float ka = rsqrt_nr(na2);
    float x0 = p[0] * ka, x1 = p[1] * ka, x2 = p[2] * ka;
    float w0 = fmaf(x1, p[5], -x2 * p[4]), w1 = fmaf(x2, p[3], -x0 * p[5]), w2 = fmaf(x0, p[4], -x1 * p[3]);
    float kz = rsqrt_nr(fmaf(w2, w2, fmaf(w1, w1, w0 * w0)));
    float z0 = w0 * kz, z1 = w1 * kz, z2 = w2 * kz;
    // upstream per column
    float gx0 = g[0], gx1 = g[3], gx2 = g[6];
    float gy0 = g[1], gy1 = g[4], gy2 = g[7];
    float gz0 = g[2], gz1 = g[5], gz2 = g[8];
    // y = z X x :  gz += x X gy ;  gx += gy X z
    gz0 += fmaf(x1, gy2, -x2 * gy1); gz1 += fmaf(x2, gy0, -x0 * gy2); gz2 += fmaf(x0, gy1, -x1 * gy0);
    gx0 += fmaf(gy1, z2, -gy2 * z1); gx1 += fmaf(gy2, z0, -gy0 * z2); gx2 += fmaf(gy0, z1, -gy1 * z0);
    // z = w/|w|
    float dz = fmaf(z2, gz2, fmaf(z1, gz1, z0 * gz0));
    float gw0 = (gz0 - z0 * dz) * kz, gw1 = (gz1 - z1 * dz) * kz, gw2 = (gz2 - z2 * dz) * kz;
    // w = x X b :  gx += b X gw ;  gb = gw X x
    gx0 += fmaf(p[4], gw2, -p[5] * gw1); gx1 += fmaf(p[5], gw0, -p[3] * gw2); gx2 += fmaf(p[3], gw1, -p[4] * gw0);
    dp[3] = fmaf(gw1, x2, -gw2 * x1); dp[4] = fmaf(gw2, x0, -gw0 * x2); dp[5] = fmaf(gw0, x1, -gw1 * x0);
    // x = a/|a|
    float dx = fmaf(x2, gx2, fmaf(x1, gx1, x0 * gx0));
    dp[0] = (gx0 - x0 * dx) * ka; dp[1] = (gx1 - x1 * dx) * ka; dp[2] = (gx2 - x2 * dx) * ka;
}

// Quaternion map, order (w, x, y, z), norm floored at 1e-8
// (reference: compute_rotation_matrix_from_quaternion + normalize_vector, rotation_representation.py:38-50,139-172).
constexpr float kQuatFloor = 1e-8f;
PE_DEV void quat_unit(const float* q, float& w, float& x, float& y, float& z, float& kn, bool& floored) {
    float n2 = fmaf(q[3], q[3], fmaf(q[2], q[2], fmaf(q[1], q[1], q[0] * q[0])));
    floored = !(n2 > kQuatFloor * kQuatFloor);
    kn = floored ? (1.0f / kQuatFloor) : rsqrt_nr(n2);
    w = q[0] * kn; x = q[1] * kn; y = q[2] * kn; z = q[3] * kn;
}
PE_DEV void quat_map(const float* q, float* r) {
    float w, x, y, z, kn; bool fl;
    quat_unit(q, w, x, y, z, kn, fl);
    float xx = x * x, yy = y * y, zz = z * z, xy = x * y, xz = x * z, yz = y * z, xw = x * w, yw = y * w, zw = z * w;
    r[0] = 1.f - 2.f * yy - 2.f * zz; r[1] = 2.f * xy - 2.f * zw;       r[2] = 2.f * xz + 2.f * yw;
    r[3] = 2.f * xy + 2.f * zw;       r[4] = 1.f - 2.f * xx - 2.f * zz; r[5] = 2.f * yz - 2.f * xw;
    r[6] = 2.f * xz - 2.f * yw;       r[7] = 2.f * yz + 2.f * xw;       r[8] = 1.f - 2.f * xx - 2.f * yy;
}
PE_DEV void quat_pullback(const float* q, const float* g, float* dq) {
    float w, x, y, z, kn; bool fl;
    quat_unit(q, w, x, y, z, kn, fl);
    float gw = 2.f * (z * (g[3] - g[1]) + y * (g[2] - g[6]) + x * (g[7] - g[5]));
    float gx = 2.f * (y * (g[1] + g[3]) + z * (g[2] + g[6]) + w * (g[7] - g[5]) - 2.f * x * (g[4] + g[8]));
    float gy = 2.f * (x * (g[1] + g[3]) + z * (g[5] + g[7]) + w * (g[2] - g[6]) - 2.f * y * (g[0] + g[8]));
    float gz = 2.f * (x * (g[2] + g[6]) + y * (g[5] + g[7]) + w * (g[3] - g[1]) - 2.f * z * (g[0] + g[4]));
    // through u = q * kn : kn constant when floored, 1/|q| otherwise
    float d = fl ? 0.f : (w * gw + x * gx + y * gy + z * gz);
    dq[0] = (gw - w * d) * kn; dq[1] = (gx - x * d) * kn; dq[2] = (gy - y * d) * kn; dq[3] = (gz - z * d) * kn;
}

// Euler map (reference: compute_rotation_matrix_from_euler, rotation_representation.py:92-115).  Note the index
// swap: angle "2" comes from e[2] and angle "3" from e[1] (lines 101-104).
PE_DEV void euler_map(const float* e, float* r) {
    float s1, c1, s2, c2, s3, c3;
    sincos_f(e[0], &s1, &c1); sincos_f(e[2], &s2, &c2); sincos_f(e[1], &s3, &c3);
    r[0] = c2 * c3;                 r[1] = -s2;     r[2] = c2 * s3;
    r[3] = c1 * s2 * c3 + s1 * s3;  r[4] = c1 * c2; r[5] = c1 * s2 * s3 - s1 * c3;
    r[6] = s1 * s2 * c3 - c1 * s3;  r[7] = s1 * c2; r[8] = s1 * s2 * s3 + c1 * c3;
}
PE_DEV void euler_pullback(const float* e, const float* g, float* de) {
    float s1, c1, s2, c2, s3, c3;
    sincos_f(e[0], &s1, &c1); sincos_f(e[2], &s2, &c2); sincos_f(e[1], &s3, &c3);
    float r3 = c1 * s2 * c3 + s1 * s3, r4 = c1 * c2, r5 = c1 * s2 * s3 - s1 * c3;
    float r6 = s1 * s2 * c3 - c1 * s3, r7 = s1 * c2, r8 = s1 * s2 * s3 + c1 * c3;
    // d/d angle1: rows 1 and 2 rotate into each other
    de[0] = g[6] * r3 + g[7] * r4 + g[8] * r5 - g[3] * r6 - g[4] * r7 - g[5] * r8;
    // d/d angle2 (= e[2])
    de[2] = -g[0] * s2 * c3 - g[1] * c2 - g[2] * s2 * s3
            + g[3] * c1 * c2 * c3 - g[4] * c1 * s2 + g[5] * c1 * c2 * s3
            + g[6] * s1 * c2 * c3 - g[7] * s1 * s2 + g[8] * s1 * c2 * s3;
    // d/d angle3 (= e[1])
    de[1] = -g[0] * c2 * s3 + g[2] * c2 * c3
            + g[3] * (s1 * c3 - c1 * s2 * s3) + g[5] * (c1 * s2 * c3 + s1 * s3)
            - g[6] * (s1 * s2 * s3 + c1 * c3) + g[8] * (s1 * s2 * c3 - c1 * s3);
}

// ---------------------------------------------------------------------------------------------------------
// 5D map (reference: compute_rotation_matrix_from_ortho5d + stereographic_unproject(axis=0),
// rotation_representation.py:72-90, 120-136).  a[2:5] is scaled by (sqrt2+1, sqrt2+1, sqrt2), unprojected from
// R^3 to the 3-sphere with the new coordinate in FRONT, divided by the norm of its last three coordinates, and
// appended to a[0:2] to form the 6D input.
PE_DEV void ortho5d_to_6d(const float* a, float* b, float* w, float& invD, float& k) {
    const float c0 = 2.41421356237309515f, c2 = 1.41421356237309515f;
    float s0 = a[2] * c0, s1 = a[3] * c0, s2 = a[4] * c2;
    float ss = fmaf(s2, s2, fmaf(s1, s1, s0 * s0));
    invD = 1.0f / (ss + 1.0f);
    w[0] = (ss - 1.0f) * invD;
    w[1] = 2.0f * s0 * invD; w[2] = 2.0f * s1 * invD; w[3] = 2.0f * s2 * invD;
    k = rsqrt_nr(fmaf(w[3], w[3], fmaf(w[2], w[2], w[1] * w[1])));
    b[0] = a[0]; b[1] = a[1];
    b[2] = w[0] * k; b[3] = w[1] * k; b[4] = w[2] * k; b[5] = w[3] * k;
}
PE_DEV void ortho5d(const float* a, float* r) {
    float b[6], w[4], invD, k;
    ortho5d_to_6d(a, b, w, invD, k);
    ortho6d(b, r);
}
PE_DEV void ortho5d_pullback(const float* a, const float* g, float* da) {
    const float c0 = 2.41421356237309515f, c2 = 1.41421356237309515f;
    float b[6], w[4], invD, k, db[6];
    ortho5d_to_6d(a, b, w, invD, k);
    ortho6d_pullback(b, g, db);
    da[0] = db[0]; da[1] = db[1];
    // u = w * k, k = 1/|w[1:]|
    float dk = fmaf(db[5], w[3], fmaf(db[4], w[2], fmaf(db[3], w[1], db[2] * w[0])));
    float dn = -dk * k * k * k;  // d|w[1:]| = -dk/|.|^2, then d w_j += dn * w_j/|.|
    float dw0 = db[2] * k;
    float dw1 = fmaf(dn, w[1], db[3] * k), dw2 = fmaf(dn, w[2], db[4] * k), dw3 = fmaf(dn, w[3], db[5] * k);
    // w0 = 1 - 2/D, w_j = 2 s_j/D, D = |s|^2 + 1
    float s0 = a[2] * c0, s1 = a[3] * c0, s2 = a[4] * c2;
    float dots = fmaf(dw3, s2, fmaf(dw2, s1, dw1 * s0));
    float q = 4.0f * invD * invD * (dw0 - dots);
    da[2] = c0 * fmaf(q, s0, 2.0f * invD * dw1);
    da[3] = c0 * fmaf(q, s1, 2.0f * invD * dw2);
    da[4] = c2 * fmaf(q, s2, 2.0f * invD * dw3);
}

// ---------------------------------------------------------------------------------------------------------
// Exponential map (reference: vec_3d_to_SO3 -> so3_exp_map, rotation_representation.py:245-320).
// theta = sqrt(max(|v|^2, 1e-4)); R = (sin theta/theta) K + ((1-cos theta)/theta^2) K^2 + I, K = hat(v).
// K^2 = v v^T - |v|^2 I uses the UNclamped |v|^2, as the reference's bmm(K, K) does.
constexpr float kExpEps = 1e-4f;
PE_DEV void expmap_factors(const float* v, float eps, float& n2, float& th, float& f1, float& f2, float& sn, float& cs) {
    n2 = fmaf(v[2], v[2], fmaf(v[1], v[1], v[0] * v[0]));
    float c = fmaxf(n2, eps);
    float it = rsqrt_nr(c);
    th = c * it;
    sincos_f(th, &sn, &cs);
    float sh, ch;
    sincos_f(0.5f * th, &sh, &ch);
    f1 = sn * it;
    f2 = 2.0f * sh * sh * it * it;  // (1 - cos) / theta^2 without the cancellation
}
PE_DEV void expmap(const float* v, float* r, float eps = kExpEps) {
    float n2, th, f1, f2, sn, cs;
    expmap_factors(v, eps, n2, th, f1, f2, sn, cs);
    float x = v[0], y = v[1], z = v[2];
    r[0] = fmaf(f2, x * x - n2, 1.0f); r[1] = fmaf(f2, x * y, -f1 * z);   r[2] = fmaf(f2, x * z, f1 * y);
    r[3] = fmaf(f2, x * y, f1 * z);    r[4] = fmaf(f2, y * y - n2, 1.0f); r[5] = fmaf(f2, y * z, -f1 * x);
    r[6] = fmaf(f2, x * z, -f1 * y);   r[7] = fmaf(f2, y * z, f1 * x);    r[8] = fmaf(f2, z * z - n2, 1.0f);
}
PE_DEV void expmap_pullback(const float* v, const float* g, float* dv, float eps = kExpEps) {
    float n2, th, f1, f2, sn, cs;
    expmap_factors(v, eps, n2, th, f1, f2, sn, cs);
    float x = v[0], y = v[1], z = v[2];
    float a0 = g[7] - g[5], a1 = g[2] - g[6], a2 = g[3] - g[1];  // sum_ij G_ij dK_ij/dv
    float tr = g[0] + g[4] + g[8];
    // (G + G^T) v
    float h0 = 2.0f * g[0] * x + (g[1] + g[3]) * y + (g[2] + g[6]) * z;
    float h1 = (g[1] + g[3]) * x + 2.0f * g[4] * y + (g[5] + g[7]) * z;
    float h2 = (g[2] + g[6]) * x + (g[5] + g[7]) * y + 2.0f * g[8] * z;
    float df1 = fmaf(x, a0, fmaf(y, a1, z * a2));
    float df2 = 0.5f * fmaf(x, h0, fmaf(y, h1, z * h2)) - n2 * tr;
    float d0 = fmaf(f2, fmaf(-2.0f * tr, x, h0), f1 * a0);
    float d1 = fmaf(f2, fmaf(-2.0f * tr, y, h1), f1 * a1);
    float d2 = fmaf(f2, fmaf(-2.0f * tr, z, h2), f1 * a2);
    if (n2 >= eps) {  // torch.clamp passes the gradient only where the input is not below the bound
        float t2 = th * th, p1, p2;  // p1 = f1'(theta)/theta, p2 = f2'(theta)/theta
        if (th < 0.25f) {
            p1 = -(1.0f / 3.0f) + t2 * ((1.0f / 30.0f) - t2 * (1.0f / 840.0f));
            p2 = -(1.0f / 12.0f) + t2 * ((1.0f / 180.0f) - t2 * (1.0f / 6720.0f));
        } else {
            float it2 = 1.0f / t2;
            p1 = (cs - f1) * it2;              // (theta cos - sin)/theta^3
            p2 = (f1 - 2.0f * f2) * it2;       // (theta sin - 2(1 - cos))/theta^4
        }
        float sct = fmaf(df1, p1, df2 * p2);
        d0 = fmaf(sct, x, d0); d1 = fmaf(sct, y, d1); d2 = fmaf(sct, z, d2);
    }
    dv[0] = d0; dv[1] = d1; dv[2] = d2;
}

// ---------------------------------------------------------------------------------------------------------
// Cloud-pair rotation sampler, deterministic part (reference: get_sampled_rotation_matrices_by_axisAngle,
// point_cloud/prepare.py:21-49): theta and the un-normalised axis come from the caller's RNG; the axis is divided by
// max(|axis|, 1e-8) (prepare.py:12-18) and the matrix is built from q = (cos theta, axis sin theta) with the unit
// quaternion formula.  NOTE the reference feeds the FULL angle into the quaternion, so the rotation is by 2 theta;
// that is reproduced, not corrected.
PE_DEV void axis_angle_matrix(float theta, const float* axis, float* r) {
    float n2 = fmaf(axis[2], axis[2], fmaf(axis[1], axis[1], axis[0] * axis[0]));
    float k = n2 > kQuatFloor * kQuatFloor ? rsqrt_nr(n2) : (1.0f / kQuatFloor);
    float sn, cs;
    sincos_f(theta, &sn, &cs);
    float w = cs, x = axis[0] * k * sn, y = axis[1] * k * sn, z = axis[2] * k * sn;
    float xx = x * x, yy = y * y, zz = z * z, xy = x * y, xz = x * z, yz = y * z, xw = x * w, yw = y * w, zw = z * w;
    r[0] = 1.f - 2.f * yy - 2.f * zz; r[1] = 2.f * xy - 2.f * zw;       r[2] = 2.f * xz + 2.f * yw;
    r[3] = 2.f * xy + 2.f * zw;       r[4] = 1.f - 2.f * xx - 2.f * zz; r[5] = 2.f * yz - 2.f * xw;
    r[6] = 2.f * xz - 2.f * yw;       r[7] = 2.f * yz + 2.f * xw;       r[8] = 1.f - 2.f * xx - 2.f * yy;
}

// ---------------------------------------------------------------------------------------------------------
// Kabsch backward, per-pair part.  Forward (pe_kabsch.cu): H = sum_k (q_k - qbar)(p_k - pbar)^T, R = svd_orth(H),
// t = qbar - R pbar.  With upstream gR (9) and gt (3):
//   gR' = gR - gt pbar^T ;  G_H = pullback(H, R, gR') ;
//   dq_k = G_H (p_k - pbar) + gt/N           = A_q p_k + b_q ,  A_q = G_H   , b_q = gt/N - G_H pbar
//   dp_k = G_H^T (q_k - qbar) - R^T gt / N   = A_p q_k + b_p ,  A_p = G_H^T , b_p = -R^T gt/N - G_H^T qbar
// (the centroid terms of dH/dq_k vanish because sum_k (p_k - pbar) = 0).  ctx = (H, pbar, qbar); cq, cp = (A, b).
PE_DEV void kabsch_bwd_coeffs(const float* ctx, const float* r, const float* g, float inv_n, float* cq, float* cp) {
    const float* h = ctx;
    const float* pb = ctx + 9;
    const float* qb = ctx + 12;
    const float* gt = g + 9;
    float gr[9], gh[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) gr[3 * i + j] = fmaf(-gt[i], pb[j], g[3 * i + j]);
    svd_orth_pullback(h, r, gr, gh);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 3; ++j) { cq[3 * i + j] = gh[3 * i + j]; cp[3 * i + j] = gh[3 * j + i]; }
        cq[9 + i] = fmaf(gt[i], inv_n, -(gh[3 * i] * pb[0] + gh[3 * i + 1] * pb[1] + gh[3 * i + 2] * pb[2]));
        float rtg = r[i] * gt[0] + r[3 + i] * gt[1] + r[6 + i] * gt[2];
        cp[9 + i] = -fmaf(rtg, inv_n, gh[i] * qb[0] + gh[3 + i] * qb[1] + gh[6 + i] * qb[2]);
    }
}

// ---------------------------------------------------------------------------------------------------------
// CosyPose-style pose update (reference: calculate_T_pred, Iterative/utility.py:90-128; camera constants from
// get_scene_parameters, :73-88).  o = network output (9 rotation entries, then vx, vy, vz); ti = T_init (4x4,
// row-major); to = T_pred.   dR = svd_orth(o[0:9]);  R_new = dR R_k;  z' = vz z_k;
// x' = (vx/fx + x_k/z_k) z';  y' = (vy/fy + y_k/z_k) z';  last row (0, 0, 0, 1)  (the intent of `combine`, :63-71).
PE_DEV void pose_update(const float* o, const float* ti, float inv_fx, float inv_fy, float* to) {
    float dr[9], ds;
    svd_orth_fast(o, dr, ds);
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            to[4 * i + j] = fmaf(dr[3 * i + 2], ti[8 + j], fmaf(dr[3 * i + 1], ti[4 + j], dr[3 * i] * ti[j]));
    float zk = ti[11], zn = o[11] * zk;
    to[3] = (o[9] * inv_fx + ti[3] / zk) * zn;
    to[7] = (o[10] * inv_fy + ti[7] / zk) * zn;
    to[11] = zn;
    to[12] = 0.f; to[13] = 0.f; to[14] = 0.f; to[15] = 1.f;
}
// pullback of pose_update w.r.t. the network output (T_init is a detached input in the reference, Iterative/main.py:93)
PE_DEV void pose_update_pullback(const float* o, const float* ti, const float* g, float inv_fx, float inv_fy,
                                 float* d_o) {
    float dr[9], ds, gd[9];
    svd_orth_fast(o, dr, ds);
    // G_dR = G_Rnew R_k^T
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            gd[3 * i + j] = fmaf(g[4 * i + 2], ti[4 * j + 2], fmaf(g[4 * i + 1], ti[4 * j + 1], g[4 * i] * ti[4 * j]));
    svd_orth_pullback(o, dr, gd, d_o);
    float zk = ti[11], zn = o[11] * zk;
    float ax = o[9] * inv_fx + ti[3] / zk, ay = o[10] * inv_fy + ti[7] / zk;
    d_o[9] = g[3] * zn * inv_fx;
    d_o[10] = g[7] * zn * inv_fy;
    d_o[11] = zk * (g[11] + g[3] * ax + g[7] * ay);
}

// ---------------------------------------------------------------------------------------------------------
// Distances on SO(3). tr(A B^T) = tr(A^T B) = sum_ij A_ij B_ij, so one inner product serves
// compute_geodesic_distance_from_two_matrices (rotation_representation.py:209-227, m1 m2^T),
// angle_error (:230-242, R1^T R2, float64, degrees) and geodesic (point_cloud/main.py:61-73).
PE_DEV float dot9(const float* a, const float* b) {
    float s = a[0] * b[0];
#pragma unroll
    for (int i = 1; i < 9; ++i) s = fmaf(a[i], b[i], s);
    return s;
}
PE_DEV double dot9_f64(const float* a, const float* b) {
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 9; ++i) s = fma((double)a[i], (double)b[i], s);
    return s;
}
// acos in DEGREES of a float64 cosine already clamped to [-1, 1], to ~2e-12 degrees (3e-14 relative near 0; measured
// against libm over 2e7 cosines incl. 1 - 1e-16 .. 1 - 1e-7): what angle_error's `torch.acos(...) * 180/pi` in float64
// needs, at 14 FP64-pipe instructions instead of libdevice acos's 42 (plus its 54 constant moves and two divergent
// paths) - the fused map + angle_error kernels were issue-bound on exactly that.
// Method: psi = acos(|c|)/2 solves sin^2(psi) = t with t = (1 - |c|)/2 EXACT in float64.  psi0 = asin(sqrt(t)) in fp32
// (MUFU.SQRT + a degree-6 polynomial, 1e-7 relative), then ONE Newton step in float64 on sin^2(psi) - t: the sine of
// psi0 (<= pi/4) is a degree-15 Taylor polynomial, the derivative sin(2 psi) = 2 sqrt(t (1 - t)) only scales a
// correction that is itself 1e-7 psi, so it is fp32 (MUFU.RSQ).  Error after the step: cot(2 psi) e0^2 <= 1e-13 psi.
PE_DEV double acos_deg_fast(double c) {
    const double ac = fabs(c);
    const double t = fma(-0.5, ac, 0.5);
    const float t32 = (float)t;
    const float h = sqrt_fast(t32);
    const float u = h * h;
    float q = 0.10931160300970078f;
    q = fmaf(q, u, -0.08920358866453171f);
    q = fmaf(q, u, 0.07470467686653137f);
    q = fmaf(q, u, 0.01712605729699135f);
    q = fmaf(q, u, 0.04638223350048065f);
    q = fmaf(q, u, 0.07489369064569473f);
    q = fmaf(q, u, 0.1666688472032547f);
    const float psi0 = fmaf(h * u, q, h);  // asin(h) = h + h^3 Q(h^2) on [0, 0.7072]
    const double x = (double)psi0;
    const double z = x * x;
    double p = 1.0 / 1307674368000.0;      // 1/15!
    p = fma(-p, z, 1.0 / 6227020800.0);
    p = fma(-p, z, 1.0 / 39916800.0);
    p = fma(-p, z, 1.0 / 362880.0);
    p = fma(-p, z, 1.0 / 5040.0);
    p = fma(-p, z, 1.0 / 120.0);
    p = fma(-p, z, 1.0 / 6.0);
    const double s = fma(-x * z, p, x);    // sin(psi0)
    const float d32 = (float)fma(s, s, -t);
    const float w = fmaf(-t32, t32, t32);  // t (1 - t) = (sin(2 psi) / 2)^2
    const float delta = w > 1e-30f ? d32 * 0.5f * rsqrt_fast(w) : 0.f;
    const double psi = x - (double)delta;
    const double k = 114.59155902616464175359630962821;  // 2 * 180 / pi
    return c >= 0.0 ? k * psi : fma(-k, psi, 180.0);
}
// angle_error of one pair (rotation_representation.py:230-242): float64 trace of R1^T R2 (the nine fp32 products are
// exact in float64), cosine, range flag, clamp, acos, degrees.  `flag` gets bit 0 when the cosine is outside
// [-1.1, 1.1] (false for NaN, exactly like the reference's two comparisons, :237).
#ifndef PE_ANGLE_LIBM
#define PE_ANGLE_LIBM 0   // 1: libdevice acos (the round-1 path, kept for A/B runs)
#endif
#ifndef PE_FUSED_ANGLE_F64
#define PE_FUSED_ANGLE_F64 0   // 1: the fused means finish the angle in float64 as well (A/B runs)
#endif
// tr = tr(R1^T R2) in float64 -> degrees (the second half of angle_error; split so that a kernel can form the trace
// while both matrices are in registers and finish the angle later, when they are dead)
PE_DEV double angle_deg_from_trace(double tr, int& flag) {
    const double c = 0.5 * (tr - 1.0);
    if (c < -1.1 || c > 1.1) flag |= 1;
    const double cc = fmin(fmax(c, -1.0), 1.0);
#if PE_ANGLE_LIBM
    return acos(cc) * 57.295779513082320876798154814105;
#else
    return c == c ? acos_deg_fast(cc) : c;  // NaN in, NaN out as torch.clamp / torch.acos (fmin / fmax would drop it)
#endif
}
// The same angle for kernels that only SUM it over the batch (the fused means): the float64 part stops at
// t = (1 - |c|)/2 = 1/2 - |tr - 1|/4, which is exact, and the arcsine of sqrt(t) is float32 - psi = asin(sqrt t) is
// well conditioned everywhere (relative error 1e-7 from t's rounding, the polynomial and MUFU.SQRT), so the angle is good to
// 1e-7 rad + 180 * 2^-24 degrees: <= 2e-5 degrees per record, without the blow-up near 0 and 180 degrees that a
// float32 TRACE would have; the per-record errors are independent, so a batch mean moves by less than 1e-6 degrees
// (north_star: 1e-4).  12 FP32 + 2 MUFU + 3 FP64 instructions instead of 24 FP32 + 14 FP64 + 4 conversions.
constexpr float kAsinC6 = 0.10931160300970078f, kAsinC5 = -0.08920358866453171f, kAsinC4 = 0.07470467686653137f,
                kAsinC3 = 0.01712605729699135f, kAsinC2 = 0.04638223350048065f, kAsinC1 = 0.07489369064569473f,
                kAsinC0 = 0.1666688472032547f;
constexpr float kTwoRadToDeg = 114.59155902616464f;  // 2 * 180 / pi
PE_DEV float angle_deg_tail32(float t, bool cos_nonneg) {
    t = t < 0.f ? 0.f : t;  // |c| > 1: the reference clamps the cosine; NaN stays NaN
    const float h = sqrt_fast(t);
    const float u = h * h;
    float q = kAsinC6;
    q = fmaf(q, u, kAsinC5);
    q = fmaf(q, u, kAsinC4);
    q = fmaf(q, u, kAsinC3);
    q = fmaf(q, u, kAsinC2);
    q = fmaf(q, u, kAsinC1);
    q = fmaf(q, u, kAsinC0);
    const float psi = fmaf(h * u, q, h);
    return cos_nonneg ? kTwoRadToDeg * psi : fmaf(-kTwoRadToDeg, psi, 180.0f);
}
PE_DEV double angle_deg_sum_term(double tr, int& flag) {
#if PE_FUSED_ANGLE_F64
    return angle_deg_from_trace(tr, flag);
#else
    const double a = tr - 1.0;                       // 2 c
    const float a32 = (float)a;
    if (a32 < -2.2f || a32 > 2.2f) flag |= 1;        // c outside [-1.1, 1.1]
    const float t = (float)fma(-0.25, fabs(a), 0.5); // (1 - |c|) / 2, exact in float64 before the rounding
    return (double)angle_deg_tail32(t, !(a < 0.0));
#endif
}
PE_DEV double angle_deg_f64(const float* a, const float* b, int& flag) {
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 9; ++i) s = fma((double)a[i], (double)b[i], s);
    return angle_deg_from_trace(s, flag);
}

// per-matrix Frobenius distance |A - B|_F (loss_frobenius, Comparison/main.py:12-16, before the batch mean)
PE_DEV float frob_dist(const float* a, const float* b) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 9; ++i) { float d = a[i] - b[i]; s = fmaf(d, d, s); }
    return sqrtf(s);
}

}  // namespace pe
