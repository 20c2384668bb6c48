"""The kernels' per-record formulas (csrc/pe_math.cuh, compiled for the host) against the oracle in float64.
CPU only: this is the arithmetic the GPU runs, minus MUFU.RSQ's last bits (the host uses 1/sqrtf)."""
import numpy as np
import pytest
import torch

from oracle import reference_port as O
import hostmath

T = lambda a: torch.from_numpy(np.asarray(a))


@pytest.fixture(scope="module")
def hm():
    return hostmath.load()


def _autograd(fn, x, g):
    x = T(x).double().requires_grad_(True)
    y = fn(x)
    (dx,) = torch.autograd.grad(y, x, T(g).double().reshape(y.shape))
    return y.detach().numpy().reshape(len(g), 9), dx.numpy()


def _rel(a, b):
    a, b = a.reshape(len(a), -1), b.reshape(len(b), -1)
    return np.abs(a - b).max(1) / (np.abs(b).max(1) + 1e-30)


MAPS = [
    ("ortho6d", 6, O.compute_rotation_matrix_from_ortho6d, "randn"),
    ("quat", 4, O.compute_rotation_matrix_from_quaternion, "randn"),
    ("euler", 3, O.compute_rotation_matrix_from_euler, "angle"),
    ("ortho5d", 5, O.compute_rotation_matrix_from_ortho5d, "randn"),
    ("expmap", 3, O.vec_3d_to_SO3, "randn"),
]


@pytest.mark.parametrize("name,w,fn,dist", MAPS)
def test_map_and_pullback_vs_oracle(hm, name, w, fn, dist):
    rng = np.random.default_rng(7)
    n = 20000
    x = rng.standard_normal((n, w)).astype(np.float32)
    if dist == "angle":
        x = rng.uniform(-np.pi, np.pi, (n, w)).astype(np.float32)
    g = rng.standard_normal((n, 9)).astype(np.float32)
    r = hostmath.call(hm, f"hm_{name}", x, (n, 9))
    dx = hostmath.call(hm, f"hm_{name}_bwd", x, g, (n, w))
    r_ref, dx_ref = _autograd(fn, x, g)
    assert np.abs(r - r_ref).max() <= 1e-5          # north_star: outputs <= 1e-5 in fp32
    # gradients: 1e-4 relative, away from the map's own singularities (tiny |a|, collinear 6D input).  A record whose
    # gradient nearly cancels is measured against a tenth of the typical gradient rather than its own tiny one.
    scale = np.abs(dx_ref).max(1)
    err = np.abs(dx - dx_ref).max(1) / np.maximum(scale, 0.1 * np.median(scale))
    ok = scale < 50 * np.median(scale)
    assert ok.mean() > 0.95
    assert np.quantile(err, 0.999) <= 1e-5
    assert err[ok].max() <= 5e-4             # the few near-collinear 6D/5D inputs lose digits in any fp32 method


def test_expmap_small_angles(hm):
    rng = np.random.default_rng(3)
    n = 4000
    x = (rng.standard_normal((n, 3)) * np.logspace(-6, 0, n)[:, None]).astype(np.float32)
    g = rng.standard_normal((n, 9)).astype(np.float32)
    r = hostmath.call(hm, "hm_expmap", x, (n, 9))
    dx = hostmath.call(hm, "hm_expmap_bwd", x, g, (n, 3))
    r_ref, dx_ref = _autograd(O.vec_3d_to_SO3, x, g)
    assert np.abs(r - r_ref).max() <= 1e-6
    assert _rel(dx, dx_ref).max() <= 1e-4


def test_svd_orth_vs_oracle(hm):
    torch.manual_seed(0)
    n = 1 << 16
    x = torch.randn(n, 9)
    ref = O.symmetric_orthogonalization(x.double()).numpy().reshape(n, 9)
    cond = O.svd_orth_conditioning(x).numpy()
    dref = O.svd_orth_det_sign(x.double()).numpy()
    r, d = hostmath.call(hm, "hm_svd_orth", x.numpy(), (n, 9), (n,))
    err = np.abs(r - ref).max(1)
    # the error of ANY fp32 method scales like eps/cond; torch's own fp32 path reaches 1.9e-6
    assert (err * cond).max() <= 1e-6
    assert err[cond > 0.1].max() <= 1e-5
    assert np.array_equal(np.sign(d), np.sign(dref))            # determinant sign: exact agreement
    rr = r.reshape(n, 3, 3).astype(np.float64)
    assert np.abs(rr @ rr.transpose(0, 2, 1) - np.eye(3)).max() <= 2e-6
    assert np.abs(np.linalg.det(rr) - 1).max() <= 2e-6


def test_svd_orth_pullback_vs_oracle(hm):
    torch.manual_seed(1)
    n = 1 << 15
    x, g = torch.randn(n, 9), torch.randn(n, 9)
    r, _ = hostmath.call(hm, "hm_svd_orth", x.numpy(), (n, 9), (n,))
    dm = hostmath.call(hm, "hm_svd_orth_bwd", x.numpy(), r, g.numpy(), (n, 9))
    ref = O.svd_orth_pullback(x, g).numpy().reshape(n, 9)
    cond = O.svd_orth_conditioning(x).numpy()
    rel = _rel(dm, ref)
    assert rel[cond > 0.1].max() <= 1e-4                        # north_star: 1e-4 away from degenerate spectra
    assert (rel * cond).max() <= 2e-5
    assert np.isfinite(dm).all()


def test_svd_orth_edge_cases(hm):
    eye = np.eye(3, dtype=np.float32).reshape(1, 9)
    cases = np.concatenate([
        eye, -eye, np.zeros((1, 9), np.float32),
        np.diag([1, 1, -1]).astype(np.float32).reshape(1, 9),
        np.diag([3, 0, 0]).astype(np.float32).reshape(1, 9),              # rank 1
        np.diag([2, 1, 0]).astype(np.float32).reshape(1, 9),              # rank 2
        (1e-30 * np.arange(1, 10, dtype=np.float32)).reshape(1, 9),       # tiny
        (1e30 * np.array([[1, 2, 3, 4, 5, 6, 7, 8, 10]], np.float32)),    # huge
    ])
    r, d = hostmath.call(hm, "hm_svd_orth", cases, (len(cases), 9), (len(cases),))
    rr = r.reshape(-1, 3, 3).astype(np.float64)
    assert np.isfinite(rr).all()
    assert np.abs(rr @ rr.transpose(0, 2, 1) - np.eye(3)).max() <= 1e-5
    assert np.abs(np.linalg.det(rr) - 1).max() <= 1e-5
    assert np.allclose(rr[0], np.eye(3), atol=1e-6)
    # rank-2 input is still uniquely determined: compare with the oracle
    ref = O.symmetric_orthogonalization(T(cases[5:6]).double()).numpy()
    assert np.abs(rr[5] - ref[0]).max() <= 1e-6
    # the two fixed matrices of 3D-Pose/loss.py:28-34 are rotations: a rotation maps to itself
    c, s = np.cos(0.3), np.sin(0.3)
    rot = np.array([[c, -s, 0, s, c, 0, 0, 0, 1]], np.float32)
    r2, _ = hostmath.call(hm, "hm_svd_orth", rot, (1, 9), (1,))
    assert np.abs(r2 - rot).max() <= 1e-6


def test_packed_two_lane_instantiation_matches_scalar(hm):
    """svd_orth_t<F2> / svd_orth_pullback_t<F2> (what the kernels run, two matrices per thread) vs the scalar code."""
    torch.manual_seed(4)
    n = 1 << 14
    x, g = torch.randn(n, 9), torch.randn(n, 9)
    x[5] = 0                                                   # degenerate lanes take the scalar fix-up path
    x[n // 2 + 7] = torch.tensor([3.0, 0, 0, 0, 0, 0, 0, 0, 0])
    r1, d1 = hostmath.call(hm, "hm_svd_orth", x.numpy(), (n, 9), (n,))
    r2, d2 = hostmath.call(hm, "hm_svd_orth_x2", x.numpy(), (n, 9), (n,))
    assert np.array_equal(d1, d2)
    assert np.abs(r1 - r2).max() <= 2e-6
    ref = O.symmetric_orthogonalization(x.double()).numpy().reshape(n, 9)
    cond = O.svd_orth_conditioning(x).numpy()
    ok = np.isfinite(cond) & (cond > 0)
    assert (np.abs(r2 - ref).max(1) * cond)[ok].max() <= 1e-6
    b1 = hostmath.call(hm, "hm_svd_orth_bwd", x.numpy(), r1, g.numpy(), (n, 9))
    b2 = hostmath.call(hm, "hm_svd_orth_bwd_x2", x.numpy(), r1, g.numpy(), (n, 9))
    assert np.isfinite(b2).all()
    scale = np.abs(b1).max(1)
    assert (np.abs(b1 - b2).max(1) / np.maximum(scale, 0.1 * np.median(scale))).max() <= 1e-5
    # the pullback the recompute / fused kernels run: from the decomposition's own V and singular values (no second
    # read of M, adjugate form of the same regularised solve); degenerate lanes fall back to the generic solve
    b3 = hostmath.call(hm, "hm_svd_orth_bwd_ctx_x2", x.numpy(), g.numpy(), (n, 9))
    assert np.isfinite(b3).all()
    e3 = np.abs(b1 - b3).max(1) / np.maximum(scale, 0.1 * np.median(scale))
    assert np.quantile(e3, 0.999) <= 1e-5
    ref_b = O.svd_orth_pullback(x, g).numpy().reshape(n, 9)
    rel3, rel1 = _rel(b3, ref_b), _rel(b1, ref_b)
    assert rel3[ok & (cond > 0.1)].max() <= 1e-4                # north_star: 1e-4 away from degenerate spectra
    assert (rel3 * cond)[ok].max() <= max(2e-5, 1.5 * (rel1 * cond)[ok].max())


# ---- SURVEY 8(f) rows ------------------------------------------------------------------------------------------------
def test_sampler_map_vs_golden(hm, golden):
    z = golden("sampler")
    r = hostmath.call(hm, "hm_axis_angle", z["theta"], z["axis"], (len(z["theta"]), 9))
    assert np.abs(r.reshape(-1, 3, 3) - z["matrix"]).max() <= 2e-6


def test_pose_update_and_pullback_vs_golden_and_oracle(hm, golden):
    z = golden("iterative")
    n = len(z["out"])
    tp = hostmath.call(hm, "hm_pose_update", z["out"], z["T_init"].reshape(n, 16), (n, 16))
    assert np.abs(tp.reshape(n, 4, 4) - z["T_pred_ref_fp32"]).max() <= 2e-5          # the reference itself, fp32
    o = T(z["out"]).double().requires_grad_(True)
    t64 = O.calculate_T_pred(o, T(z["T_init"]).double())
    assert np.abs(tp.reshape(n, 4, 4) - t64.detach().numpy()).max() <= 1e-5
    (g,) = torch.autograd.grad(t64, o, T(z["g_T"]))
    d = hostmath.call(hm, "hm_pose_update_bwd", z["out"], z["T_init"].reshape(n, 16),
                      z["g_T"].astype(np.float32).reshape(n, 16), (n, 12))
    assert _rel(d, g.numpy()).max() <= 1e-4


def test_kabsch_backward_coefficients_vs_oracle_autograd(hm):
    """dL/dP, dL/dQ of L = sum(gR * R) + sum(gt * t) through the oracle's Kabsch (float64 autograd through torch.svd)
    against the affine coefficients the kernel computes per pair."""
    torch.manual_seed(11)
    b, npts = 64, 40
    p = (torch.rand(b, npts, 3) - 0.5).float()
    rot = O.sample_rotations_axis_angle(b, torch.Generator().manual_seed(3)).float()
    q = (p @ rot.transpose(1, 2) + 0.05 * torch.randn(b, npts, 3) + torch.randn(b, 1, 3)).float()
    g_r, g_t = torch.randn(b, 3, 3).double(), torch.randn(b, 3).double()
    p64, q64 = p.double().requires_grad_(True), q.double().requires_grad_(True)
    r, t = O.kabsch(p64, q64)
    dp_ref, dq_ref = torch.autograd.grad((r * g_r).sum() + (t * g_t).sum(), (p64, q64))
    # context as the forward kernel leaves it: H = sum (q - qbar)(p - pbar)^T, pbar, qbar
    pb, qb = p.double().mean(1), q.double().mean(1)
    h = (q.double() - qb[:, None]).transpose(1, 2) @ (p.double() - pb[:, None])
    ctx = torch.cat([h.reshape(b, 9), pb, qb], 1).float().numpy()
    g = torch.cat([g_r.reshape(b, 9), g_t], 1).float().numpy()
    inv_n = np.full((b, 1), 1.0 / npts, np.float32)
    cq, cp = hostmath.call(hm, "hm_kabsch_bwd_coeffs", ctx, r.detach().float().numpy().reshape(b, 9), g, inv_n,
                           (b, 12), (b, 12))
    A_q, b_q = T(cq[:, :9]).double().reshape(b, 3, 3), T(cq[:, 9:]).double()
    A_p, b_p = T(cp[:, :9]).double().reshape(b, 3, 3), T(cp[:, 9:]).double()
    dq = p.double() @ A_q.transpose(1, 2) + b_q[:, None]
    dp = q.double() @ A_p.transpose(1, 2) + b_p[:, None]
    assert _rel(dq.numpy(), dq_ref.numpy()).max() <= 1e-4
    assert _rel(dp.numpy(), dp_ref.numpy()).max() <= 1e-4


def test_newton_acos_in_degrees_vs_libm(hm):
    """acos_deg_fast (pe_math.cuh): 14 FP64 instructions instead of libdevice acos; must agree with float64 acos to
    far below the 1e-9 degree tolerance of the angle_error tests, including next to +-1 and at 0."""
    import ctypes
    rng = np.random.default_rng(0)
    c = np.concatenate([rng.uniform(-1, 1, 400000), 1 - 10.0 ** rng.uniform(-16, -1, 100000),
                        -1 + 10.0 ** rng.uniform(-16, -1, 100000), rng.uniform(-1e-3, 1e-3, 10000),
                        np.array([1.0, -1.0, 0.0, 0.5, -0.5])])
    c = np.clip(c, -1.0, 1.0)
    out = np.empty_like(c)
    dp = ctypes.POINTER(ctypes.c_double)
    hm.hm_acos_deg(c.ctypes.data_as(dp), out.ctypes.data_as(dp), ctypes.c_long(len(c)))
    ref = np.degrees(np.arccos(c))
    assert np.abs(out - ref).max() <= 5e-12
    nz = ref > 0
    assert (np.abs(out - ref)[nz] / ref[nz]).max() <= 1e-12
    assert out[-5] == 0.0 and out[-4] == 180.0


def test_angle_error_record_vs_oracle(hm):
    """angle_deg_f64 = float64 trace of the fp32 entries + range flag + clamp + Newton acos, against the oracle's
    angle_error (rotation_representation.py:230-242) on the same fp32 inputs; NaN propagates, the flag does not fire on it."""
    import ctypes
    g = torch.Generator().manual_seed(3)
    r1 = O.symmetric_orthogonalization(torch.randn(20000, 9, generator=g).double()).float()
    r2 = O.symmetric_orthogonalization(torch.randn(20000, 9, generator=g).double()).float()
    r2[:100] = r1[:100]                                  # identical pairs: the acos floor near zero angle
    a, b = np.ascontiguousarray(r1.reshape(-1, 9).numpy()), np.ascontiguousarray(r2.reshape(-1, 9).numpy())
    deg, flag = np.empty(len(a)), np.empty(len(a), np.int32)
    fp, dp, ip = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int)
    hm.hm_angle_deg(a.ctypes.data_as(fp), b.ctypes.data_as(fp), deg.ctypes.data_as(dp), flag.ctypes.data_as(ip),
                    ctypes.c_long(len(a)))
    ref = O.angle_error(r1, r2).numpy()
    assert np.abs(deg - ref).max() <= 1e-10 and not flag.any()
    bad = np.ascontiguousarray(np.stack([3 * np.eye(3).reshape(9), np.full(9, np.nan)]).astype(np.float32))
    eye = np.ascontiguousarray(np.stack([np.eye(3).reshape(9)] * 2).astype(np.float32))
    d2, f2 = np.empty(2), np.empty(2, np.int32)
    hm.hm_angle_deg(bad.ctypes.data_as(fp), eye.ctypes.data_as(fp), d2.ctypes.data_as(dp), f2.ctypes.data_as(ip),
                    ctypes.c_long(2))
    assert f2[0] == 1 and d2[0] == 0.0                   # cos = 4 -> flagged; clamped to 1 -> 0 degrees
    assert f2[1] == 0 and np.isnan(d2[1])                # NaN: no flag (both comparisons false, :237), NaN angle


def test_expmap_with_other_clamp_bounds_vs_golden(hm, golden):
    z = golden("round2")
    v = z["exp_v"]
    for tag, eps in (("1em2", 1e-2), ("1em6", 1e-6)):
        e = np.array([eps], np.float32)
        r = hostmath.call(hm, "hm_expmap_eps", v, e, (len(v), 9))
        assert np.abs(r.reshape(-1, 3, 3) - z[f"exp_r_{tag}"]).max() <= 2e-6
        dv = hostmath.call(hm, "hm_expmap_eps_bwd", v, z["exp_w"].reshape(-1, 9), e, (len(v), 3))
        scale = np.abs(z[f"exp_g_{tag}"]).max(1)
        assert (np.abs(dv - z[f"exp_g_{tag}"]).max(1) / np.maximum(scale, 1.0)).max() <= 2e-5


def test_fused_sum_angle_forms_vs_oracle(hm):
    """The angle the fused-mean kernels add up: (a) float64 trace + float32 arcsine tail, (b) error-free float32 trace
    (big-accumulator FMA + recovered rounding errors) + the same tail, two lanes at once.  Against the oracle's float64
    angle_error on the same fp32 matrices: <= 2e-5 degrees per record everywhere - also for identical pairs, where a
    float32 trace would be off by 0.03 degrees - and the error-free trace equals the float64 one to 2e-11."""
    import ctypes
    g = torch.Generator().manual_seed(4)
    n = 40000
    r1 = O.symmetric_orthogonalization(torch.randn(n, 9, generator=g).double()).float()
    r2 = O.symmetric_orthogonalization(torch.randn(n, 9, generator=g).double()).float()
    r2[:2000] = r1[:2000]                                                  # zero angle: the reference's acos floor
    small = O.so3_exp_map(1e-3 * torch.randn(2000, 3, generator=g).double())
    r2[2000:4000] = torch.bmm(r1[2000:4000].double(), small).float()       # ~0.06 degree apart
    flip = torch.diag(torch.tensor([1.0, -1.0, -1.0]))
    r2[4000:6000] = torch.matmul(r1[4000:6000], flip)                      # 180 degrees
    a, b = np.ascontiguousarray(r1.reshape(-1, 9).numpy()), np.ascontiguousarray(r2.reshape(-1, 9).numpy())
    ref = O.angle_error(r1, r2).numpy()
    fp, dp, ip = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int)
    deg, flag = np.empty(n), np.empty(n, np.int32)
    hm.hm_angle_sum_term(a.ctypes.data_as(fp), b.ctypes.data_as(fp), deg.ctypes.data_as(dp), flag.ctypes.data_as(ip),
                         ctypes.c_long(n))
    assert np.abs(deg - ref).max() <= 2e-5 and not flag.any()
    deg2, pair = np.empty(n), np.empty((n, 2), np.float32)
    hm.hm_angle_split_x2(a.ctypes.data_as(fp), b.ctypes.data_as(fp), deg2.ctypes.data_as(dp), pair.ctypes.data_as(fp),
                         ctypes.c_long(n))
    tr = (r1.double().reshape(n, 9) * r2.double().reshape(n, 9)).sum(1).numpy()
    assert np.abs(pair[:, 0].astype(np.float64) + pair[:, 1].astype(np.float64) - (tr - 1.0)).max() <= 2e-11
    assert np.abs(deg2 - ref).max() <= 2e-5
    assert abs(deg2.mean() - ref.mean()) <= 1e-6 and abs(deg.mean() - ref.mean()) <= 1e-6
    # the one-record-per-thread form (elements 0-4 / 5-8 in the two halves of packed registers)
    deg3, pair3 = np.empty(n), np.empty((n, 2), np.float32)
    hm.hm_angle_split_scalar(a.ctypes.data_as(fp), b.ctypes.data_as(fp), deg3.ctypes.data_as(dp),
                             pair3.ctypes.data_as(fp), ctypes.c_long(n))
    assert np.abs(pair3[:, 0].astype(np.float64) + pair3[:, 1].astype(np.float64) - (tr - 1.0)).max() <= 2e-11
    assert np.abs(deg3 - ref).max() <= 2e-5 and abs(deg3.mean() - ref.mean()) <= 1e-6


def test_rotation_is_invariant_under_power_of_two_scales(hm):
    """The exact power-of-two prescale of svd_orth_f: the rotation must not depend on the magnitude of the input, from
    2^-40 to 2^40, in the packed two-lane form, also when the two lanes of a pair differ by 2^30.  (Written for a
    variant that skipped the prescale inside a safe window - measured slower, see DESIGN.md 8 - and kept as a test.)"""
    g = torch.Generator().manual_seed(11)
    n = 4000
    x = torch.randn(n, 9, generator=g).numpy()
    r0, d0 = hostmath.call(hm, "hm_svd_orth_x2", x, (n, 9), (n,))
    for k in list(range(-14, 13)) + [-40, -25, 25, 40]:
        xs = (x * np.float32(2.0 ** k)).astype(np.float32)
        r, d = hostmath.call(hm, "hm_svd_orth_x2", xs, (n, 9), (n,))
        assert np.abs(r - r0).max() <= 5e-7 and np.array_equal(d, d0), k
    # one lane of a pair inside the window, its partner outside (record i rides with record i + n/2)
    mix = x.copy()
    mix[: n // 2] *= np.float32(2.0 ** 30)
    r, d = hostmath.call(hm, "hm_svd_orth_x2", mix, (n, 9), (n,))
    assert np.abs(r - r0).max() <= 5e-7 and np.array_equal(d, d0)
