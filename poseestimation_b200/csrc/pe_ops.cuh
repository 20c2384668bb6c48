// The per-record ops of the path, written against the streaming engine (pe_stream.cuh).
// Each Op names the reference function it replaces; the math is in pe_math.cuh.
#pragma once
#include "pe_lanes.cuh"
#include "pe_stream.cuh"

namespace pe {

// (ring depths: stages_for in pe_stream.cuh; a one-record-per-thread CTA stays near 54 KB = 4 CTAs per SM)
// Output ring of an op that does not store in place: three buffers (one barrier per tile, see kOutBufs in pe_stream.cuh)
// unless an input stage is so large (> 32 KB) that a single CTA with a two-stage ring owns the SM: there the refill has to
// go out as early as possible - before the output tile is written, which is what the two-barrier order does
// (measured: pose_update_bwd, 45 KB per stage, 0.077 ms with two buffers, 0.093 ms with three; every other op is as fast
// or up to 8 % faster with three).
constexpr int outbufs_for(int witot, int wotot) {
    return wotot == 0 ? 0 : (kOutBufs >= 3 && kTile * 4 * witot <= 32 * 1024 ? kOutBufs : 2);
}

// Shared-memory budget of a one-record-per-thread CTA.  54 KB = 4 CTAs/SM.  Measured alternatives (r1m, 2^24 records):
// 36 KB (6 CTAs/SM) gains 5-6 % on euler_bwd / expmap_bwd, 0-1 % on the write-heavy forward maps and loses 1 % on the
// float64 angle-error kernels; a four-buffer output ring (-DPE_OUT_BUFS=4) gains 3.6 % on ortho6d_fwd and nothing
// elsewhere.  Neither is a clear win, so the defaults stay; the forward maps are not limited by stores in flight.
#ifndef PE_LIGHT_BUDGET
#define PE_LIGHT_BUDGET (54 * 1024)
#endif

struct NoExtra {};

#define PE_OP_COMMON_R(NAME, I0, I1, I2, O0, O1, NR, FL, MB, R)                              \
    static constexpr int WI0 = (I0), WI1 = (I1), WI2 = (I2), WO0 = (O0), WO1 = (O1), NRED = (NR); \
    static constexpr bool FLAG = (FL);                                                       \
    static constexpr int RPT = (R);                                                          \
    static constexpr bool kInPlace = (R) == 2 && WO0 == WI0 && WO1 == 0;                     \
    static constexpr int kBudget = (R) == 2 && (MB) < 4 ? 75 * 1024 : PE_LIGHT_BUDGET;           \
    static constexpr int OUTBUFS = outbufs_for((R) * (WI0 + WI1 + WI2), kInPlace ? 0 : (R) * (WO0 + WO1)); \
    static constexpr int STAGES = stages_for((R) * (WI0 + WI1 + WI2), kInPlace ? 0 : (R) * (WO0 + WO1), kBudget, \
                                             OUTBUFS >= 3 ? OUTBUFS : 2);                              \
    static constexpr int MINB = (MB);                                                        \
    static constexpr bool WIDE_OK = (R) == 2 && (I0) == 9 && (I2) == 0;                      \
    static const char* name() { return NAME; }
// one record per thread
#define PE_OP_COMMON(NAME, I0, I1, I2, O0, O1, NR, FL, MB) PE_OP_COMMON_R(NAME, I0, I1, I2, O0, O1, NR, FL, MB, 1)
// two records per thread, packed fp32x2 arithmetic (the SVD ops: issue-bound)
#define PE_OP_COMMON2(NAME, I0, I1, I2, O0, O1, NR, FL, MB) PE_OP_COMMON_R(NAME, I0, I1, I2, O0, O1, NR, FL, MB, 2)

// ---- a1 symmetric_orthogonalization (rotation_representation.py:192-206) ------------------------------------
// Every SVD op has a lane-generic body over lazy sources: `run` (T = float, registers) serves the direct kernel,
// `run2` (T = F2, shared-memory sources read on demand) the TMA kernel.
template <int W>
struct RegSrc {  // scalar twin of PairSrc: the record is already in registers
    const float* v;
    __device__ __forceinline__ void load(float* dst) const {
#pragma unroll
        for (int i = 0; i < W; ++i) dst[i] = v[i];
    }
};
template <class T, class Src>
__device__ __forceinline__ void svd_from(const Src& src, T* r, T& ds) {
    svd_orth_f<T>([&src](T* dst) { src.load(dst); }, r, ds);
}
// decomposition that keeps (V, singular values) for the pullback in the same thread, and that pullback
template <class T, class Src>
__device__ __forceinline__ void svd_from_ctx(const Src& src, T* r, SvdCtx<T>& ctx) {
    T ds;
    svd_orth_f<T>([&src](T* dst) { src.load(dst); }, r, ds, &ctx);
}
template <class T, class Src>
__device__ __forceinline__ void pullback_ctx(const SvdCtx<T>& ctx, const Src& src, const T* r, T z0, T z1, T z2, T* dm) {
    svd_orth_pullback_ctx<T>(ctx, [&src](T* dst) { src.load(dst); }, r, z0, z1, z2, dm);
}

template <bool DET>
struct SvdFwd {
    PE_OP_COMMON2("svd_orth_fwd", 9, 0, 0, 9, DET ? 1 : 0, 0, false, 3)
    using Extra = NoExtra;
    template <class T, class S0>
    __device__ static __forceinline__ void body(const S0& m, T* r, T* d) {
        T ds;
        svd_from<T>(m, r, ds);
        if constexpr (DET) d[0] = ds;
    }
    __device__ static __forceinline__ void run(const float* m, const float*, const float*, float* r, float* d,
                                               double*, int&, const Extra&) { body<float>(RegSrc<9>{m}, r, d); }
    template <class S0>  // PairSrc<9>, or PairSrcWide for the out[:, :9] view of a 12-word-row output
    __device__ static __forceinline__ void run2(const S0& m, const PairSrc<0>&, const PairSrc<0>&, F2* r, F2* d,
                                                double*, int&, bool, bool, const Extra&) { body<F2>(m, r, d); }
};
// pullback with the saved forward output: in = (m, r, g)
struct SvdBwdSaved {
    PE_OP_COMMON2("svd_orth_bwd", 9, 9, 9, 9, 0, 0, false, 2)
    using Extra = NoExtra;
    __device__ static __forceinline__ void run(const float* m, const float* r, const float* g, float* dm, float*,
                                               double*, int&, const Extra&) { svd_orth_pullback_t(m, r, g, dm); }
    __device__ static __forceinline__ void run2(const PairSrc<9>& ms, const PairSrc<9>& rs, const PairSrc<9>& gs, F2* dm,
                                                F2*, double*, int&, bool, bool, const Extra&) {
        F2 m[9], r[9], g[9];
        ms.load(m); rs.load(r); gs.load(g);
        svd_orth_pullback_t(m, r, g, dm);
    }
};
// pullback that recomputes r from m: in = (m, g)
struct SvdBwdRecompute {
    PE_OP_COMMON2("svd_orth_bwd_recompute", 9, 9, 0, 9, 0, 0, false, 3)
    using Extra = NoExtra;
    template <class T, class S0, class S1>
    __device__ static __forceinline__ void body(const S0& ms, const S1& gs, T* dm) {
        T r[9], g[9];
        SvdCtx<T> ctx;
        svd_from_ctx<T>(ms, r, ctx);
        gs.load(g);
        T z0, z1, z2;
        skew_axial_rtg(r, g, z0, z1, z2);
        pullback_ctx<T>(ctx, ms, r, z0, z1, z2, dm);
    }
    __device__ static __forceinline__ void run(const float* m, const float* g, const float*, float* dm, float*,
                                               double*, int&, const Extra&) {
        body<float>(RegSrc<9>{m}, RegSrc<9>{g}, dm);
    }
    template <class S0, class S1>  // S1: PairSrc<9>, or PairSrcLate<9> (split wide ring: the load waits for the tile)
    __device__ static __forceinline__ void run2(const S0& m, const S1& g, const PairSrc<0>&, F2* dm, F2*,
                                                double*, int&, bool, bool, const Extra&) { body<F2>(m, g, dm); }
};

// ---- elementwise maps: forward (x -> 9 words) and pullback ((x, g) -> dx) ---------------------------------------
#define PE_MAP_OPS(Tag, W, FWD, BWD, LABEL)                                                                   \
    struct Tag##Fwd {                                                                                         \
        PE_OP_COMMON(LABEL "_fwd", W, 0, 0, 9, 0, 0, false, 4)                                                \
        using Extra = NoExtra;                                                                                \
        __device__ static __forceinline__ void run(const float* x, const float*, const float*, float* r,     \
                                                   float*, double*, int&, const Extra&) {                    \
            FWD(x, r);                                                                                        \
        }                                                                                                     \
    };                                                                                                        \
    struct Tag##Bwd {                                                                                         \
        PE_OP_COMMON(LABEL "_bwd", W, 9, 0, W, 0, 0, false, 4)                                                \
        using Extra = NoExtra;                                                                                \
        __device__ static __forceinline__ void run(const float* x, const float* g, const float*, float* dx,  \
                                                   float*, double*, int&, const Extra&) {                    \
            BWD(x, g, dx);                                                                                    \
        }                                                                                                     \
    };
PE_MAP_OPS(Ortho6d, 6, ortho6d, ortho6d_pullback, "ortho6d")   // rotation_representation.py:21-36
PE_MAP_OPS(Quat, 4, quat_map, quat_pullback, "quat")           // rotation_representation.py:139-172
PE_MAP_OPS(Euler, 3, euler_map, euler_pullback, "euler")       // rotation_representation.py:92-115
PE_MAP_OPS(Ortho5d, 5, ortho5d, ortho5d_pullback, "ortho5d")   // rotation_representation.py:120-136
PE_MAP_OPS(Expmap, 3, expmap, expmap_pullback, "expmap")       // rotation_representation.py:245-320 (eps = 1e-4)
// so3_exp_map(log_rot, eps) with the caller's clamp bound (rotation_representation.py:245-272 takes any eps)
struct ExpEps {
    float eps;
};
struct ExpmapEpsFwd {
    PE_OP_COMMON("expmap_eps_fwd", 3, 0, 0, 9, 0, 0, false, 4)
    using Extra = ExpEps;
    __device__ static __forceinline__ void run(const float* x, const float*, const float*, float* r, float*, double*,
                                               int&, const Extra& ex) { expmap(x, r, ex.eps); }
};
struct ExpmapEpsBwd {
    PE_OP_COMMON("expmap_eps_bwd", 3, 9, 0, 3, 0, 0, false, 4)
    using Extra = ExpEps;
    __device__ static __forceinline__ void run(const float* x, const float* g, const float*, float* dx, float*, double*,
                                               int&, const Extra& ex) { expmap_pullback(x, g, dx, ex.eps); }
};

// ---- a5 loss_frobenius (Comparison/main.py:12-16) ---------------------------------------------------------------
struct FrobFwd {
    PE_OP_COMMON("frobenius_fwd", 9, 9, 0, 0, 0, 1, false, 4)
    using Extra = NoExtra;
    __device__ static __forceinline__ void run(const float* a, const float* b, const float*, float*, float*,
                                               double* red, int&, const Extra&) {
        red[0] += (double)frob_dist(a, b);
    }
};
struct GradScale {
    const float* gout;  // device scalar: upstream gradient of the 0-d loss
    float mul;          // 1/n (mean) times the sign of the operand
};
// G = gscale (a - b)/|a - b|;  BOTH: out0 = G, out1 = -G;  else out0 = G (mul carries the sign for db-only)
template <bool BOTH>
struct FrobBwd {
    PE_OP_COMMON("frobenius_bwd", 9, 9, 0, 9, BOTH ? 9 : 0, 0, false, 4)
    using Extra = GradScale;
    __device__ static __forceinline__ void run(const float* a, const float* b, const float*, float* da, float* db,
                                               double*, int&, const Extra& ex) {
        float d[9], s = 0.f;
#pragma unroll
        for (int i = 0; i < 9; ++i) { d[i] = a[i] - b[i]; s = fmaf(d[i], d[i], s); }
        float k = s > 0.f ? __ldg(ex.gout) * ex.mul * rsqrt_nr(s) : 0.f;
#pragma unroll
        for (int i = 0; i < 9; ++i) {
            da[i] = k * d[i];
            if constexpr (BOTH) db[i] = -da[i];
        }
    }
};

// ---- a6 / a8 geodesic distance (rotation_representation.py:209-227, point_cloud/main.py:61-73) ---------------------
struct Clamp {
    float lo, hi;
};
template <bool THETA, bool SUM>
struct GeoFwd {
    PE_OP_COMMON("geodesic_fwd", 9, 9, 0, THETA ? 1 : 0, 0, SUM ? 1 : 0, false, 4)
    using Extra = Clamp;
    __device__ static __forceinline__ void run(const float* a, const float* b, const float*, float* th, float*,
                                               double* red, int&, const Extra& ex) {
        float c = 0.5f * (dot9(a, b) - 1.0f);
        float t = acosf(fminf(fmaxf(c, ex.lo), ex.hi));
        if constexpr (THETA) th[0] = t;
        if constexpr (SUM) red[0] += (double)t;
    }
};
struct GeoGrad {
    float lo, hi, gmul;
    const float* gscalar;
};
// MODE 0: out0 = d1 ; 1: out0 = d2 ; 2: out0 = d1, out1 = d2.   GS: upstream gradient is one device scalar.
template <int MODE, bool GS>
struct GeoBwd {
    PE_OP_COMMON("geodesic_bwd", 9, 9, GS ? 0 : 1, 9, MODE == 2 ? 9 : 0, 0, false, 4)
    using Extra = GeoGrad;
    __device__ static __forceinline__ void run(const float* a, const float* b, const float* g, float* o0, float* o1,
                                               double*, int&, const Extra& ex) {
        float c = 0.5f * (dot9(a, b) - 1.0f);
        float up = (GS ? __ldg(ex.gscalar) : g[0]) * ex.gmul;
        // d theta/d c = -1/sqrt(1 - c^2); d c/d a = b/2.  Zero where the clamp is active.
        float k = (c >= ex.lo && c <= ex.hi) ? -0.5f * up * rsqrt_nr(fmaf(-c, c, 1.0f)) : 0.f;
#pragma unroll
        for (int i = 0; i < 9; ++i) {
            if constexpr (MODE == 0) o0[i] = k * b[i];
            if constexpr (MODE == 1) o0[i] = k * a[i];
            if constexpr (MODE == 2) { o0[i] = k * b[i]; o1[i] = k * a[i]; }
        }
    }
};

// ---- a7 angle_error (rotation_representation.py:230-242): float64 arithmetic, degrees, range flag ---------------
template <bool DEG, bool SUM>
struct AngleFwd {
    PE_OP_COMMON("angle_error_fwd", 9, 9, 0, DEG ? 2 : 0, 0, SUM ? 1 : 0, true, 4)
    using Extra = NoExtra;
    __device__ static __forceinline__ void run(const float* a, const float* b, const float*, float* out, float*,
                                               double* red, int& flag, const Extra&) {
        const double t = angle_deg_f64(a, b, flag);
        if constexpr (DEG) {
            out[0] = __int_as_float(__double2loint(t));
            out[1] = __int_as_float(__double2hiint(t));
        }
        if constexpr (SUM) red[0] += t;
    }
};

// ---- fused: loss_frobenius(symmetric_orthogonalization(x), R_true) ------------------------------------------------
// Seed of the pullback for G = gs (R - R_t)/|R - R_t|_F without forming G:  R^T G = k (R^T R - R^T R_t) and R^T R is
// symmetric, so axial(R^T G - G^T R) = -k axial(S - S^T) with S = R^T R_t.  f = |R - R_t|_F, k = gs / f (0 if f = 0).
// axial(S - S^T) = -sum_k r_k x rt_k over the rows, and with D = R - R_t that is sum_k r_k x d_k: the difference is
// formed anyway for f, and the cross products of the small d do not cancel when R is close to R_t.
template <class T>
__device__ __forceinline__ void frob_seed(const T* r, const T* rt, float gs, T& f, T& z0, T& z1, T& z2) {
    T d[9];
    d[0] = t_sub(r[0], rt[0]);
    T s = t_mul(d[0], d[0]);
#pragma unroll
    for (int i = 1; i < 9; ++i) { d[i] = t_sub(r[i], rt[i]); s = t_fma(d[i], d[i], s); }
    T nk = s;
    f = s;
#pragma unroll
    for (int l = 0; l < LaneCount<T>::value; ++l) {
        float sl = lane_get(s, l);
        float inv = sl > 0.f ? rsqrt_nr(sl) : 0.f;
        lane_put(nk, l, -gs * inv);
        lane_put(f, l, sl * inv);
    }
    z0 = t_mul(r[1], d[2]); z1 = t_mul(r[2], d[0]); z2 = t_mul(r[0], d[1]);
    z0 = t_fma(t_neg(r[2]), d[1], z0); z1 = t_fma(t_neg(r[0]), d[2], z1); z2 = t_fma(t_neg(r[1]), d[0], z2);
#pragma unroll
    for (int k = 1; k < 3; ++k) {
        const T a0 = r[3 * k], a1 = r[3 * k + 1], a2 = r[3 * k + 2];
        const T b0 = d[3 * k], b1 = d[3 * k + 1], b2 = d[3 * k + 2];
        z0 = t_fma(a1, b2, z0); z1 = t_fma(a2, b0, z1); z2 = t_fma(a0, b1, z2);
        z0 = t_fma(t_neg(a2), b1, z0); z1 = t_fma(t_neg(a0), b2, z1); z2 = t_fma(t_neg(a1), b0, z2);
    }
    z0 = t_mul(z0, nk); z1 = t_mul(z1, nk); z2 = t_mul(z2, nk);
}

template <bool WRITE_R>
struct SvdFrobFwd {
    PE_OP_COMMON2("svd_orth_frobenius_fwd", 9, 9, 0, WRITE_R ? 9 : 0, 0, 1, false, 3)
    using Extra = NoExtra;
    template <class T, class S0, class S1>
    __device__ static __forceinline__ T body(const S0& ms, const S1& rts, T* r) {
        T rr[9], rt[9], d[9], ds;
        svd_from<T>(ms, rr, ds);
        rts.load(rt);
        T s = diff9_sq(rr, rt, d);
        if constexpr (WRITE_R) {
#pragma unroll
            for (int i = 0; i < 9; ++i) r[i] = rr[i];
        }
        return t_sqrt(s);
    }
    __device__ static __forceinline__ void run(const float* m, const float* rt, const float*, float* r, float*,
                                               double* red, int&, const Extra&) {
        red[0] += (double)body<float>(RegSrc<9>{m}, RegSrc<9>{rt}, r);
    }
    template <class S0, class S1>  // S1: PairSrc<9>, or PairSrcLate<9> (split wide ring: the load waits for the tile)
    __device__ static __forceinline__ void run2(const S0& m, const S1& rt, const PairSrc<0>&, F2* r, F2*,
                                                double* red, int&, bool v0, bool v1, const Extra&) {
        F2 f = body<F2>(m, rt, r);
        red[0] += (double)((v0 ? f.x : 0.f) + (v1 ? f.y : 0.f));
    }
};
struct SvdFrobBwd {
    PE_OP_COMMON2("svd_orth_frobenius_bwd", 9, 9, 0, 9, 0, 0, false, 3)
    using Extra = GradScale;
    template <class T, class S0, class S1>
    __device__ static __forceinline__ void body(const S0& ms, const S1& rts, T* dm, float gs) {
        T r[9], m[9];
        SvdCtx<T> ctx;
        svd_from_ctx<T>(ms, r, ctx);
        rts.load(m);  // R_true
        T f;
        T z0, z1, z2;
        frob_seed(r, m, gs, f, z0, z1, z2);
        pullback_ctx<T>(ctx, ms, r, z0, z1, z2, dm);
    }
    __device__ static __forceinline__ void run(const float* m, const float* rt, const float*, float* dm, float*,
                                               double*, int&, const Extra& ex) {
        body<float>(RegSrc<9>{m}, RegSrc<9>{rt}, dm, __ldg(ex.gout) * ex.mul);
    }
    template <class S0, class S1>  // S1: PairSrc<9>, or PairSrcLate<9> (split wide ring: the load waits for the tile)
    __device__ static __forceinline__ void run2(const S0& m, const S1& rt, const PairSrc<0>&, F2* dm, F2*,
                                                double*, int&, bool, bool, const Extra& ex) {
        body<F2>(m, rt, dm, __ldg(ex.gout) * ex.mul);
    }
};

// ---- 8f.4 cloud-pair rotation sampler, deterministic part (point_cloud/prepare.py:21-49): (theta, axis) -> R ------
struct AxisAngleFwd {
    PE_OP_COMMON("axis_angle_fwd", 1, 3, 0, 9, 0, 0, false, 4)
    using Extra = NoExtra;
    __device__ static __forceinline__ void run(const float* th, const float* ax, const float*, float* r, float*,
                                               double*, int&, const Extra&) { axis_angle_matrix(th[0], ax, r); }
};
// ---- 8f.3 Kabsch backward, per-pair part: (ctx, R, [gR | gt]) -> affine coefficients for dQ and dP -----------------
struct InvCount {
    float inv_n;
};
struct KabschBwdCoeffs {
    PE_OP_COMMON("kabsch_bwd_coeffs", 15, 9, 12, 12, 12, 0, false, 2)
    using Extra = InvCount;
    __device__ static __forceinline__ void run(const float* ctx, const float* r, const float* g, float* cq, float* cp,
                                               double*, int&, const Extra& ex) {
        kabsch_bwd_coeffs(ctx, r, g, ex.inv_n, cq, cp);
    }
};
// weighted Kabsch: the context record carries W = sum of the pair's weights as a 16th word; 1/N becomes 1/W
struct KabschBwdCoeffsW {
    PE_OP_COMMON("kabsch_weighted_bwd_coeffs", 16, 9, 12, 12, 12, 0, false, 2)
    using Extra = NoExtra;
    __device__ static __forceinline__ void run(const float* ctx, const float* r, const float* g, float* cq, float* cp,
                                               double*, int&, const Extra&) {
        kabsch_bwd_coeffs(ctx, r, g, ctx[15] > 0.f ? 1.0f / ctx[15] : 0.f, cq, cp);
    }
};
// ---- 8f.2 CosyPose pose update (Iterative/utility.py:90-128): (net output 12, T_init 16) -> T_pred 16 ---------------
struct Camera {
    float inv_fx, inv_fy;
};
struct PoseUpdateFwd {
    PE_OP_COMMON("pose_update_fwd", 12, 16, 0, 16, 0, 0, false, 2)
    using Extra = Camera;
    __device__ static __forceinline__ void run(const float* o, const float* ti, const float*, float* to, float*,
                                               double*, int&, const Extra& ex) {
        pose_update(o, ti, ex.inv_fx, ex.inv_fy, to);
    }
};
struct PoseUpdateBwd {
    PE_OP_COMMON("pose_update_bwd", 12, 16, 16, 12, 0, 0, false, 2)
    using Extra = Camera;
    __device__ static __forceinline__ void run(const float* o, const float* ti, const float* g, float* d_o, float*,
                                               double*, int&, const Extra& ex) {
        pose_update_pullback(o, ti, g, ex.inv_fx, ex.inv_fy, d_o);
    }
};

// single pass of the training step: the loss AND the gradient for a unit upstream gradient from ONE decomposition
// (reads M, R_true; writes dM = mul * pullback((R - R_true)/|R - R_true|_F); adds |R - R_true|_F to the sum).  108 B.
struct StepScale {
    float mul;  // 1/n of the batch mean
};
// ANGLE: additionally sums angle_error(R, R_true) in degrees (float64, rotation_representation.py:230-242) into a second
// accumulator and ORs its range flag - every training loop of the reference computes BOTH from the same outR each step
// (Comparison/main.py:59-70, 3D-Pose/main.py:60-62,85-91), and the pair [sum |.|_F, sum theta] is what the ranks
// all-reduce.  The trace of R^T R_true is formed in float64 while both matrices are in registers (exact
// products); the acos runs at the end of the record, when only dM is live.
// PE_STEP_ANGLE: how the step kernel evaluates the angle (A/B switch; DESIGN.md 4.4 has the measurements)
//   0  float64 trace (9 DFMA + 18 conversions per lane) and the float64 Newton acos
//   1  float64 trace, float32 arcsine tail (angle_deg_sum_term)
//   2  error-free float32 trace (trace_minus_one_split, packed for both lanes) and the float32 tail: no float64 at all
//      until the angle is added to the accumulator
#ifndef PE_STEP_ANGLE
#define PE_STEP_ANGLE 2
#endif
// sum of the angles of one record (pair) from the error-free float32 trace (pe_lanes.cuh), with the range flag
template <class T>
__device__ __forceinline__ void angle_sum_from_split(T ahi, T lo, double& red, int& flag, bool v0, bool v1) {
    const bool valid[2] = {v0, v1};
    decltype(t_ltz(ahi)) neg;
    T a32;
    const T t = half_versine_from_split(ahi, lo, neg, a32);
    const T deg = angle_deg_tail32_t(t, neg);
#pragma unroll
    for (int l = 0; l < LaneCount<T>::value; ++l) {
        if (valid[l]) {
            const float a = lane_get(a32, l);
            if (a < -2.2f || a > 2.2f) flag |= 1;  // cosine outside [-1.1, 1.1] (rotation_representation.py:237)
            red += (double)lane_get(deg, l);
        }
    }
}
template <class T>
struct AngleSeed {
#if PE_STEP_ANGLE == 2
    T ahi, lo;
#else
    double tr[LaneCount<T>::value];
#endif
};
template <bool ANGLE>
struct SvdFrobStepT {
    PE_OP_COMMON2(ANGLE ? "svd_orth_frobenius_step_metrics" : "svd_orth_frobenius_step", 9, 9, 0, 9, 0, ANGLE ? 2 : 1,
                  ANGLE, 3)
    using Extra = StepScale;
    template <class T, class S0, class S1>
    __device__ static __forceinline__ T body(const S0& ms, const S1& rts, T* dm, float gs, AngleSeed<T>& seed) {
        T r[9], m[9];
        SvdCtx<T> ctx;
        svd_from_ctx<T>(ms, r, ctx);
        rts.load(m);  // R_true
        if constexpr (ANGLE) {
#if PE_STEP_ANGLE == 2
            trace_minus_one_split(r, m, seed.ahi, seed.lo);
#else
#pragma unroll
            for (int l = 0; l < LaneCount<T>::value; ++l) {
                double s = 0.0;
#pragma unroll
                for (int i = 0; i < 9; ++i) s = fma((double)lane_get(r[i], l), (double)lane_get(m[i], l), s);
                seed.tr[l] = s;
            }
#endif
        }
        T f;
        T z0, z1, z2;
        frob_seed(r, m, gs, f, z0, z1, z2);
        pullback_ctx<T>(ctx, ms, r, z0, z1, z2, dm);
        return f;
    }
    // the angle(s) of the record (pair) from what the body kept; runs when only dM is live
    template <class T>
    __device__ static __forceinline__ void finish(const AngleSeed<T>& seed, double* red, int& flag, bool v0, bool v1) {
#if PE_STEP_ANGLE == 2
        angle_sum_from_split(seed.ahi, seed.lo, red[1], flag, v0, v1);
#else
        const bool valid[2] = {v0, v1};
#pragma unroll
        for (int l = 0; l < LaneCount<T>::value; ++l) {
            if (valid[l]) {
#if PE_STEP_ANGLE == 1
                red[1] += angle_deg_sum_term(seed.tr[l], flag);
#else
                red[1] += angle_deg_from_trace(seed.tr[l], flag);
#endif
            }
        }
#endif
    }
    __device__ static __forceinline__ void run(const float* m, const float* rt, const float*, float* dm, float*,
                                               double* red, int& flag, const Extra& ex) {
        AngleSeed<float> seed;
        red[0] += (double)body<float>(RegSrc<9>{m}, RegSrc<9>{rt}, dm, ex.mul, seed);
        if constexpr (ANGLE) finish<float>(seed, red, flag, true, false);
    }
    template <class S0, class S1>  // S1: PairSrc<9>, or PairSrcLate<9> (split wide ring: the load waits for the tile)
    __device__ static __forceinline__ void run2(const S0& m, const S1& rt, const PairSrc<0>&, F2* dm, F2*,
                                                double* red, int& flag, bool v0, bool v1, const Extra& ex) {
        AngleSeed<F2> seed;
        F2 f = body<F2>(m, rt, dm, ex.mul, seed);
        red[0] += (double)((v0 ? f.x : 0.f) + (v1 ? f.y : 0.f));
        if constexpr (ANGLE) finish<F2>(seed, red, flag, v0, v1);
    }
};
using SvdFrobStep = SvdFrobStepT<false>;
using SvdFrobStepMetrics = SvdFrobStepT<true>;

// ---- fused map + metric: metric(map(x), R_true) summed over the batch, R never written ------
// The evaluation loops of the reference compute R = transform_output(out) and then angle_error(R, R_true).mean()
// (UPNA/main.py:93-96, 3D-Pose/main.py:116-120, Comparison/main.py:70-75): 4W + 36 bytes written, then 72 read.
// Fused: 4W + 36 bytes per record in all.  METRIC 0: |R - R_t|_F (loss_frobenius), 1: geodesic angle in radians, fp32
// acos, clamp from Extra (a6 / a8), 2: angle_error in degrees, float64, with its range flag (a7).
template <int METRIC>
__device__ __forceinline__ void metric_accumulate(const float* r, const float* rt, double* red, int& flag,
                                                  const Clamp& ex) {
    if constexpr (METRIC == 0) {
        red[0] += (double)frob_dist(rt, r);
    } else if constexpr (METRIC == 1) {
        float c = 0.5f * (dot9(r, rt) - 1.0f);
        red[0] += (double)acosf(fminf(fmaxf(c, ex.lo), ex.hi));
    } else {
#if PE_STEP_ANGLE == 2
        // error-free float32 trace: the float64 one costs 18 conversions per record, and those saturate the XU pipe
        // (ncu r2c: quat / euler + angle_error at 66 % / 59 % XU, FMA pipe at 24 %)
        float ahi, lo;
        trace_minus_one_split_scalar(r, rt, ahi, lo);
        angle_sum_from_split<float>(ahi, lo, red[0], flag, true, false);
#else
        red[0] += angle_deg_sum_term(dot9_f64(r, rt), flag);
#endif
    }
}
#define PE_MAP_METRIC(Tag, W, FWD, LABEL)                                                                     \
    template <int METRIC>                                                                                     \
    struct Tag##Metric {                                                                                      \
        PE_OP_COMMON(LABEL "_metric", W, 9, 0, 0, 0, 1, METRIC == 2, 4)                                       \
        using Extra = Clamp;                                                                                  \
        __device__ static __forceinline__ void run(const float* x, const float* rt, const float*, float*,    \
                                                   float*, double* red, int& flag, const Extra& ex) {        \
            float r[9];                                                                                       \
            FWD(x, r);                                                                                        \
            metric_accumulate<METRIC>(r, rt, red, flag, ex);                                                  \
        }                                                                                                     \
    };
PE_MAP_METRIC(Ortho6d, 6, ortho6d, "ortho6d")
PE_MAP_METRIC(Quat, 4, quat_map, "quat")
PE_MAP_METRIC(Euler, 3, euler_map, "euler")
PE_MAP_METRIC(Ortho5d, 5, ortho5d, "ortho5d")
PE_MAP_METRIC(Expmap, 3, expmap, "expmap")
template <int METRIC>
struct SvdMetric {
    PE_OP_COMMON2("svd_orth_metric", 9, 9, 0, 0, 0, 1, METRIC == 2, 3)
    using Extra = Clamp;
    __device__ static __forceinline__ void run(const float* m, const float* rt, const float*, float*, float*,
                                               double* red, int& flag, const Extra& ex) {
        float r[9], ds;
        svd_from<float>(RegSrc<9>{m}, r, ds);
        metric_accumulate<METRIC>(r, rt, red, flag, ex);
    }
    template <class S0, class S1>  // S1: PairSrc<9>, or PairSrcLate<9> (split wide ring: the load waits for the tile)
    __device__ static __forceinline__ void run2(const S0& m, const S1& rts, const PairSrc<0>&, F2*, F2*,
                                                double* red, int& flag, bool v0, bool v1, const Extra& ex) {
        F2 r[9], rt[9], ds;
        svd_from<F2>(m, r, ds);
        rts.load(rt);
#if PE_STEP_ANGLE == 2
        if constexpr (METRIC == 2) {  // both lanes at once: error-free float32 trace, float32 arcsine
            F2 ahi, lo;
            trace_minus_one_split(r, rt, ahi, lo);
            angle_sum_from_split(ahi, lo, red[0], flag, v0, v1);
            return;
        }
#endif
        float a[9], b[9];
        if (v0) {
#pragma unroll
            for (int i = 0; i < 9; ++i) { a[i] = r[i].x; b[i] = rt[i].x; }
            metric_accumulate<METRIC>(a, b, red, flag, ex);
        }
        if (v1) {
#pragma unroll
            for (int i = 0; i < 9; ++i) { a[i] = r[i].y; b[i] = rt[i].y; }
            metric_accumulate<METRIC>(a, b, red, flag, ex);
        }
    }
};

}  // namespace pe
