"""Out-of-bounds WRITE check for every entry point of the C ABI (compute-sanitizer is not available on the GPU pool).

Every output array of every call sits between two guard bands filled with a sentinel bit pattern; after the call the
bands must be bit-for-bit intact and the payload must have been written completely (no sentinel left in it).  Runs each
call with 16-byte aligned arrays (TMA-staged kernels, 128-bit stores) and with arrays offset by one float (the direct
kernels), at one record, ragged tails, and batches that give a CTA several tiles.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

SENT = 0x7FC0DEAD            # a quiet NaN with a recognisable payload
GUARD = 1024                 # floats on either side


@pytest.fixture(scope="module")
def abi():
    from poseestimation_b200 import _lib
    return _lib


class Guarded:
    """`n` 4-byte words with guard bands; `off` shifts the payload by that many words (alignment of the base)."""

    def __init__(self, n, off=0, fill=None):
        self.n, self.off = n, off
        self.buf = torch.full((n + 2 * GUARD + 4,), SENT, dtype=torch.int32, device="cuda")
        self.lo = GUARD + off
        if fill is not None:
            self.words().view(torch.float32).copy_(fill.reshape(-1))

    @property
    def ptr(self):
        return self.buf.data_ptr() + 4 * self.lo

    def words(self):
        return self.buf[self.lo:self.lo + self.n]

    def f32(self, *shape):
        return self.words().view(torch.float32).reshape(*shape)

    def check(self, name, written=True):
        torch.cuda.synchronize()
        assert bool((self.buf[:self.lo] == SENT).all()), f"{name}: wrote below the array"
        assert bool((self.buf[self.lo + self.n:] == SENT).all()), f"{name}: wrote past the array"
        if written and self.n:
            assert not bool((self.words() == SENT).any()), f"{name}: part of the output was never written"


def _inp(t, off):
    return Guarded(t.numel(), off, fill=t)


def _call(abi, name, *args):
    rc = getattr(abi.lib, name)(*args)
    assert rc == 0, f"{name}: rc={rc} {abi.lib.pe_last_error()}"


@pytest.mark.parametrize("off", [0, 1])
@pytest.mark.parametrize("n", [1, 255, 513, 70001, 200003])
def test_record_kernels_stay_inside_their_outputs(abi, n, off):
    torch.manual_seed(n)
    st = torch.cuda.current_stream().cuda_stream
    rnd = lambda *s: torch.randn(*s, device="cuda")
    outs = []

    def out(words, f64=False):            # float64 outputs (sums, degrees) must be 8-byte aligned: no odd offset
        g = Guarded(words, 0 if f64 else off)
        outs.append(g)
        return g

    def done(name):
        for g in outs:
            g.check(name)
        outs.clear()

    m, g9 = _inp(rnd(n, 9), off), _inp(rnd(n, 9), off)
    r, ds = out(9 * n), out(n)
    _call(abi, "pe_svd_orth_fwd", m.ptr, 9, r.ptr, ds.ptr, n, st)
    rot = _inp(r.f32(n, 9).clone(), off)
    done("pe_svd_orth_fwd")
    for saved in (rot.ptr, None):
        dm = out(9 * n)
        _call(abi, "pe_svd_orth_bwd", m.ptr, 9, saved, g9.ptr, dm.ptr, n, st)
        done("pe_svd_orth_bwd")
    # strided rows read in place
    wide = _inp(rnd(n, 12), off)
    r = out(9 * n)
    _call(abi, "pe_svd_orth_fwd", wide.ptr, 12, r.ptr, None, n, st)
    done("pe_svd_orth_fwd(strided)")
    for name, w in (("ortho6d", 6), ("quat", 4), ("euler", 3), ("ortho5d", 5), ("expmap", 3)):
        x = _inp(rnd(n, w), off)
        r = out(9 * n)
        _call(abi, f"pe_{name}_fwd", x.ptr, w, r.ptr, n, st)
        done(f"pe_{name}_fwd")
        dx = out(w * n)
        _call(abi, f"pe_{name}_bwd", x.ptr, w, g9.ptr, dx.ptr, n, st)
        done(f"pe_{name}_bwd")
    rot2 = _inp(rot.f32(n, 9).flip(0).contiguous(), off)
    one = _inp(torch.full((1,), 0.5, device="cuda"), 0)
    s = out(2, f64=True)
    _call(abi, "pe_frobenius_fwd", rot.ptr, rot2.ptr, s.ptr, n, st)
    done("pe_frobenius_fwd")
    s, mean = out(2, f64=True), out(1)
    _call(abi, "pe_frobenius_mean_fwd", rot.ptr, rot2.ptr, s.ptr, mean.ptr, n, st)
    torch.cuda.synchronize()
    ref = float(s.words().view(torch.float64)[0]) / n
    assert abs(float(mean.f32(1)[0]) - ref) <= 1.2e-7 * abs(ref)           # (float)(sum * (1/n)): one rounding
    done("pe_frobenius_mean_fwd")
    da, db = out(9 * n), out(9 * n)
    _call(abi, "pe_frobenius_bwd", rot.ptr, rot2.ptr, one.ptr, 1.0 / n, da.ptr, db.ptr, n, st)
    done("pe_frobenius_bwd")
    th, s = out(n), out(2, f64=True)
    _call(abi, "pe_geodesic_fwd", rot.ptr, rot2.ptr, -1.0, 1.0, th.ptr, s.ptr, n, st)
    theta = _inp(th.f32(n).clone(), off)
    done("pe_geodesic_fwd")
    s2, mean = out(2, f64=True), out(1)
    _call(abi, "pe_geodesic_reduce_fwd", rot.ptr, rot2.ptr, -1.0, 1.0, s2.ptr, mean.ptr, 1.0 / n, n, st)
    torch.cuda.synchronize()
    ref = float(s2.words().view(torch.float64)[0]) / n
    assert abs(float(mean.f32(1)[0]) - ref) <= 1.2e-7 * abs(ref)
    assert abs(float(s2.words().view(torch.float64)[0]) - float(s.words().view(torch.float64)[0])) <= 1e-9 * n
    done("pe_geodesic_reduce_fwd")
    for scalar in (0, 1):
        d1, d2 = out(9 * n), out(9 * n)
        gsrc = one if scalar else theta
        _call(abi, "pe_geodesic_bwd", rot.ptr, rot2.ptr, -1.0, 1.0, gsrc.ptr, scalar, 1.0, d1.ptr, d2.ptr, n, st)
        done("pe_geodesic_bwd")
    # float64 outputs need 8-byte alignment: odd word offsets do not apply to them
    deg, s, flag = Guarded(2 * n, 0), Guarded(2, 0), Guarded(1, 0)
    outs.extend([deg, s, flag])
    _call(abi, "pe_angle_error_fwd", rot.ptr, rot2.ptr, deg.ptr, s.ptr, flag.ptr, n, st)
    done("pe_angle_error_fwd")
    r, s = out(9 * n), out(2, f64=True)
    _call(abi, "pe_svd_orth_frobenius_fwd", m.ptr, 9, rot2.ptr, r.ptr, s.ptr, n, st)
    done("pe_svd_orth_frobenius_fwd")
    dm = out(9 * n)
    _call(abi, "pe_svd_orth_frobenius_bwd", m.ptr, 9, rot2.ptr, one.ptr, 1.0 / n, dm.ptr, n, st)
    done("pe_svd_orth_frobenius_bwd")
    dm, s = out(9 * n), out(2, f64=True)
    _call(abi, "pe_svd_orth_frobenius_step", m.ptr, 9, rot2.ptr, 1.0 / n, dm.ptr, s.ptr, n, st)
    done("pe_svd_orth_frobenius_step")
    dm2, s2, mean = out(9 * n), out(2, f64=True), out(1)
    _call(abi, "pe_svd_orth_frobenius_step_mean", m.ptr, 9, rot2.ptr, 1.0 / n, dm2.ptr, s2.ptr, mean.ptr, n, st)
    torch.cuda.synchronize()
    ref = float(s2.words().view(torch.float64)[0]) / n
    assert abs(float(mean.f32(1)[0]) - ref) <= 1.2e-7 * abs(ref)
    assert abs(float(s2.words().view(torch.float64)[0]) - float(s.words().view(torch.float64)[0])) <= 1e-9 * n
    assert torch.equal(dm2.words(), dm.words())
    done("pe_svd_orth_frobenius_step_mean")
    # the step with its metric: [sum |.|_F, sum theta, flag] + float32 / float64 means + int flag, dense and 12-word rows
    for src, stride in ((m, 9), (wide, 12)):
        dm3, s3 = out(9 * n), Guarded(6, 0)
        mean, amean, flag = out(1), Guarded(2, 0), Guarded(1, 0)
        outs.extend([s3, amean, flag])
        _call(abi, "pe_svd_orth_frobenius_step_metrics", src.ptr, stride, rot2.ptr, 1.0 / n, dm3.ptr, s3.ptr, mean.ptr,
              amean.ptr, flag.ptr, None, n, st)
        torch.cuda.synchronize()
        if stride == 9:
            assert torch.equal(dm3.words(), dm.words())
        v = s3.words().view(torch.float64)
        assert abs(float(amean.words().view(torch.float64)[0]) - float(v[1]) / n) <= 1e-12 and float(v[2]) == 0.0
        done("pe_svd_orth_frobenius_step_metrics")
    # the other SVD-family kernels on 12-word rows (wide TMA stage when aligned, direct kernel otherwise)
    dmw = out(9 * n)
    _call(abi, "pe_svd_orth_bwd", wide.ptr, 12, None, g9.ptr, dmw.ptr, n, st)
    done("pe_svd_orth_bwd(strided)")
    rw, sw = out(9 * n), out(2, f64=True)
    _call(abi, "pe_svd_orth_frobenius_fwd", wide.ptr, 12, rot2.ptr, rw.ptr, sw.ptr, n, st)
    done("pe_svd_orth_frobenius_fwd(strided)")
    dmw = out(9 * n)
    _call(abi, "pe_svd_orth_frobenius_bwd", wide.ptr, 12, rot2.ptr, one.ptr, 1.0 / n, dmw.ptr, n, st)
    done("pe_svd_orth_frobenius_bwd(strided)")
    x3 = _inp(rnd(n, 3), off)
    r3, d3 = out(9 * n), out(3 * n)
    _call(abi, "pe_expmap_eps_fwd", x3.ptr, 3, 1e-2, r3.ptr, n, st)
    _call(abi, "pe_expmap_eps_bwd", x3.ptr, 3, g9.ptr, 1e-2, d3.ptr, n, st)
    done("pe_expmap_eps")
    if off == 0:                                                    # documented: 16-byte aligned arrays only
        y = out(9 * n)
        _call(abi, "pe_scale_by_scalar", m.ptr, one.ptr, y.ptr, 9 * n, st)
        done("pe_scale_by_scalar")
        y = Guarded(9 * n, 0, fill=rnd(n, 9))
        outs.append(y)
        _call(abi, "pe_scale_by_scalar_inplace", y.ptr, one.ptr, 9 * n, st)
        done("pe_scale_by_scalar_inplace")
    th1, ax = _inp(rnd(n), off), _inp(rnd(n, 3), off)
    r = out(9 * n)
    _call(abi, "pe_axis_angle_fwd", th1.ptr, ax.ptr, r.ptr, n, st)
    done("pe_axis_angle_fwd")
    t_init = torch.eye(4, device="cuda").repeat(n, 1, 1) + 0.1 * rnd(n, 4, 4)
    t_init[:, 2, 3] += 2.0
    o12, ti, gt = _inp(rnd(n, 12), off), _inp(t_init, off), _inp(rnd(n, 16), off)
    tp = out(16 * n)
    _call(abi, "pe_pose_update_fwd", o12.ptr, 12, ti.ptr, 444.0, 444.0, tp.ptr, n, st)
    done("pe_pose_update_fwd")
    do = out(12 * n)
    _call(abi, "pe_pose_update_bwd", o12.ptr, 12, ti.ptr, gt.ptr, 444.0, 444.0, do.ptr, n, st)
    done("pe_pose_update_bwd")


@pytest.mark.parametrize("off", [0, 1])
@pytest.mark.parametrize("b,npts", [(1, 1), (3, 5), (5, 1024), (33, 1021), (150, 260), (700, 64)])
def test_cloud_kernels_stay_inside_their_outputs(abi, b, npts, off):
    torch.manual_seed(b * npts)
    st = torch.cuda.current_stream().cuda_stream
    rnd = lambda *s: torch.randn(*s, device="cuda")
    outs = []

    def out(words, f64=False):            # float64 outputs (sums, degrees) must be 8-byte aligned: no odd offset
        g = Guarded(words, 0 if f64 else off)
        outs.append(g)
        return g

    def done(name):
        for g in outs:
            g.check(name)
        outs.clear()

    p, q = _inp(rnd(b, npts, 3), off), _inp(rnd(b, npts, 3), off)
    w = _inp(torch.rand(b, npts, device="cuda"), off)
    g12 = _inp(rnd(b, 12), off)
    r, t = out(9 * b), out(3 * b)
    _call(abi, "pe_kabsch_fwd", p.ptr, q.ptr, b, npts, r.ptr, t.ptr, st)
    done("pe_kabsch_fwd")
    r, t, c = out(9 * b), out(3 * b), out(15 * b)
    _call(abi, "pe_kabsch_fwd_ctx", p.ptr, q.ptr, b, npts, r.ptr, t.ptr, c.ptr, st)
    rk, ck = _inp(r.f32(b, 9).clone(), off), _inp(c.f32(b, 15).clone(), off)
    done("pe_kabsch_fwd_ctx")
    cq, cp = out(12 * b), out(12 * b)
    _call(abi, "pe_kabsch_bwd_coeffs", ck.ptr, rk.ptr, g12.ptr, npts, cq.ptr, cp.ptr, b, st)
    cqi, cpi = _inp(cq.f32(b, 12).clone(), off), _inp(cp.f32(b, 12).clone(), off)
    done("pe_kabsch_bwd_coeffs")
    r, t, c = out(9 * b), out(3 * b), out(16 * b)
    _call(abi, "pe_kabsch_weighted_fwd", p.ptr, q.ptr, w.ptr, b, npts, r.ptr, t.ptr, c.ptr, st)
    rk, cw = _inp(r.f32(b, 9).clone(), off), _inp(c.f32(b, 16).clone(), off)
    done("pe_kabsch_weighted_fwd")
    cq, cp = out(12 * b), out(12 * b)
    _call(abi, "pe_kabsch_weighted_bwd_coeffs", cw.ptr, rk.ptr, g12.ptr, cq.ptr, cp.ptr, b, st)
    done("pe_kabsch_weighted_bwd_coeffs")
    dp, dq, dw = out(3 * b * npts), out(3 * b * npts), out(b * npts)
    _call(abi, "pe_kabsch_weighted_bwd_points", p.ptr, q.ptr, w.ptr, cqi.ptr, cpi.ptr, cw.ptr, b, npts, dp.ptr, dq.ptr,
          dw.ptr, st)
    done("pe_kabsch_weighted_bwd_points")
    dw = out(b * npts)
    _call(abi, "pe_kabsch_weighted_bwd_points", p.ptr, q.ptr, w.ptr, cqi.ptr, cpi.ptr, cw.ptr, b, npts, None, None,
          dw.ptr, st)
    done("pe_kabsch_weighted_bwd_points(dw)")
    m, s = out(9 * b), out(3 * b)
    _call(abi, "pe_cloud_outer_sum", q.ptr, p.ptr, b, npts, m.ptr, s.ptr, st)
    done("pe_cloud_outer_sum")
    o = out(3 * b * npts)
    _call(abi, "pe_affine_points", cqi.ptr, 12, 3, 0, cqi.ptr + 36, 12, 1, p.ptr, b, npts, o.ptr, st)
    done("pe_affine_points")
    poses = torch.eye(4, device="cuda").repeat(b, 1, 1) + 0.1 * rnd(b, 4, 4)
    tg, tp = _inp(poses, off), _inp(poses + 0.05 * rnd(b, 4, 4), off)
    o = out(3 * b * npts)
    _call(abi, "pe_affine_points", tg.ptr, 16, 4, 1, tg.ptr + 12, 16, 4, p.ptr, b, npts, o.ptr, st)
    done("pe_affine_points(pose, transposed)")
    for dis in (0, 1):
        item, gr = out(b), out(12 * b)
        s = Guarded(2, 0)
        outs.append(s)
        _call(abi, "pe_add_l1", tg.ptr, tp.ptr, p.ptr, b, npts, dis, item.ptr, s.ptr, gr.ptr, st)
        done("pe_add_l1")


def test_guard_harness_detects_an_overrun(abi):
    """The harness itself: a deliberately over-long call must trip it."""
    st = torch.cuda.current_stream().cuda_stream
    x = _inp(torch.randn(300, 3, device="cuda"), 0)
    r = Guarded(9 * 299, 0)
    _call(abi, "pe_euler_fwd", x.ptr, 3, r.ptr, 300, st)               # one record more than the array holds
    with pytest.raises(AssertionError):
        r.check("overrun")
