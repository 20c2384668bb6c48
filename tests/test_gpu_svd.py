"""GPU parity for symmetric orthogonalisation (forward, analytic backward, fused loss step) - BASELINE configs 1, 2."""
import numpy as np
import pytest
import torch

from oracle import reference_port as O

pytestmark = pytest.mark.gpu

T = lambda a: torch.from_numpy(np.asarray(a))


@pytest.fixture(scope="module")
def rr():
    import poseestimation_b200.rotation_representation as m
    return m


@pytest.fixture(scope="module")
def ops():
    from poseestimation_b200 import ops as o
    return o


def _relrows(a, b):
    """per-record gradient error relative to the record's own gradient, or to a tenth of the typical gradient when
    the record's gradient nearly cancels (relative error of a near-zero vector is meaningless)."""
    a, b = np.asarray(a, np.float64).reshape(len(a), -1), np.asarray(b, np.float64).reshape(len(b), -1)
    scale = np.abs(b).max(1)
    return np.abs(a - b).max(1) / (np.maximum(scale, 0.1 * np.median(scale)) + 1e-30)


def test_svd_orth_matches_reference_golden(rr, ops, golden):
    z = golden("svd_orth")
    n = int(z["n_random"])
    x = T(z["x"]).cuda().requires_grad_(True)
    r = rr.symmetric_orthogonalization(x)
    r.backward(T(z["g"]).cuda())
    got = r.detach().cpu().numpy()
    s = z["s"]
    d = np.sign(z["det"])
    cond = np.minimum(s[:, 1] + d * s[:, 2], s[:, 0] + d * s[:, 2]) / s[:, 0]
    err = np.abs(got - z["r"]).reshape(len(s), -1).max(1)
    # random part: the reference's own fp32 result is the yardstick for what fp32 can do on the same input
    ref32 = np.abs(z["r_fp32"].astype(np.float64) - z["r"]).reshape(len(s), -1).max(1)
    assert err[:n][cond[:n] > 0.05].max() <= 1e-5
    assert (err[:n] * cond[:n]).max() <= 1e-6
    assert np.median(err[:n]) <= 2 * np.median(ref32[:n]) + 1e-7
    # known answers: identity, diag(1,1,-1) -> det +1, rotations map to themselves
    assert np.abs(got[n] - np.eye(3)).max() <= 1e-6
    assert abs(np.linalg.det(got[n + 1].astype(np.float64)) - 1) <= 1e-5
    assert np.abs(got[n + 3:n + 67] - z["x"][n + 3:n + 67].reshape(-1, 3, 3)).max() <= 2e-6
    # determinant sign: exact agreement with the reference wherever det(M) is not at rounding level
    _, ds = ops.svd_orth_det(T(z["x"]).cuda())
    m64 = z["x"].reshape(-1, 3, 3).astype(np.float64)
    clear = np.abs(np.linalg.det(m64)) > 1e-6 * np.abs(m64).max((1, 2)) ** 3
    assert np.array_equal(np.sign(ds.cpu().numpy())[clear], d[clear])
    # gradient vs the reference's autograd, away from degenerate spectra
    gap = np.minimum(s[:, 0] - s[:, 1], s[:, 1] - s[:, 2]) / s[:, 0]
    ok = (cond > 0.1) & (gap > 1e-3)
    ok[n:] = False
    assert ok.sum() > 300
    # (the appended known-answer cases are degenerate: the reference's autograd returns NaN for some of them)
    assert _relrows(x.grad.cpu().numpy()[ok], z["dx"][ok]).max() <= 1e-4


@pytest.mark.parametrize("n", [1, 2, 255, 256, 257, 4096, (1 << 20) + 3])
def test_svd_orth_fwd_bwd_vs_oracle(rr, ops, n):
    g = torch.Generator().manual_seed(n)
    x, up = torch.randn(n, 9, generator=g), torch.randn(n, 3, 3, generator=g)
    xd = x.cuda().requires_grad_(True)
    r = rr.symmetric_orthogonalization(xd)
    r.backward(up.cuda())
    ref = O.symmetric_orthogonalization(x.double()).numpy()
    cond = O.svd_orth_conditioning(x).numpy()
    err = np.abs(r.detach().cpu().numpy() - ref).reshape(n, -1).max(1)
    assert (err * cond).max() <= 1e-6
    assert err[cond > 0.05].max(initial=0) <= 1e-5
    _, ds = ops.svd_orth_det(x.cuda())
    assert np.array_equal(np.sign(ds.cpu().numpy()), np.sign(O.svd_orth_det_sign(x.double()).numpy()))
    dref = O.svd_orth_pullback(x, up).numpy()
    rel = _relrows(xd.grad.cpu().numpy(), dref)
    assert rel[cond > 0.1].max(initial=0) <= 1e-4
    assert (rel * cond).max() <= 5e-5
    # recompute variant of the backward gives the same gradient
    d2 = ops.svd_orth_bwd_recompute(x.cuda(), up.cuda())
    assert _relrows(d2.cpu().numpy(), dref)[cond > 0.1].max(initial=0) <= 1e-4


def test_svd_orth_near_rotation_inputs(rr):
    """What a trained network emits: R + 0.1 * noise (second distribution)."""
    g = torch.Generator().manual_seed(3)
    n = 1 << 16
    rot = O.sample_rotations_axis_angle(n, g).float()
    x = (rot + 0.1 * torch.randn(n, 3, 3, generator=g)).reshape(n, 9)
    r = rr.symmetric_orthogonalization(x.cuda()).cpu().numpy()
    ref = O.symmetric_orthogonalization(x.double()).numpy()
    assert np.abs(r - ref).max() <= 1e-5            # well conditioned: the plain 1e-5 bound holds for every matrix


def test_strided_rows_are_read_in_place(rr):
    """out[:, :9] of a (B, 12) network output (3D-Pose/main2.py:94, Iterative/utility.py:105)."""
    torch.manual_seed(2)
    out = torch.randn(3001, 12).cuda().requires_grad_(True)
    r = rr.symmetric_orthogonalization(out[:, :9])
    ref = O.symmetric_orthogonalization(out.detach().cpu()[:, :9].double())
    cond = O.svd_orth_conditioning(out.detach().cpu()[:, :9].contiguous()).numpy()
    assert (np.abs(r.detach().cpu().numpy() - ref.numpy()).reshape(3001, -1).max(1) * cond).max() <= 1e-6
    r.sum().backward()
    assert out.grad.shape == (3001, 12) and float(out.grad[:, 9:].abs().max()) == 0.0
    # 3-D input too
    r3 = rr.symmetric_orthogonalization(out.detach()[:, :9].reshape(-1, 3, 3))
    assert torch.equal(r3, r.detach())


def test_full_size_properties(rr):
    """BASELINE config 2 size (2^24): size-independent properties instead of an elementwise oracle."""
    n = 1 << 24
    torch.manual_seed(0)
    x = torch.randn(n, 9, device="cuda")
    r = rr.symmetric_orthogonalization(x)
    rtr = torch.bmm(r.transpose(1, 2), r) - torch.eye(3, device="cuda")
    assert float(rtr.abs().max()) <= 5e-6                      # orthogonal
    det = torch.linalg.det(r.double())
    assert float((det - 1).abs().max()) <= 1e-5                # proper
    r2 = rr.symmetric_orthogonalization(r.reshape(n, 9))
    assert float((r2 - r).abs().max()) <= 5e-6                 # idempotent on rotations
    # equivariance: svd_orth(Q M) = Q svd_orth(M) for a rotation Q (holds to conditioning-limited accuracy)
    q = r[:1].expand(1 << 20, 3, 3)
    m = x[1:(1 << 20) + 1].reshape(-1, 3, 3)
    lhs = rr.symmetric_orthogonalization(torch.bmm(q, m).reshape(-1, 9))
    rhs = torch.bmm(q, r[1:(1 << 20) + 1])
    assert float((lhs - rhs).abs().median()) <= 1e-6
    # the tr(R^T M) objective is maximal: no worse than the float64 oracle on a slice
    sl = slice(0, 1 << 16)
    ref = O.symmetric_orthogonalization(x[sl].cpu().double())
    obj = (r[sl].cpu().double() * x[sl].cpu().double().reshape(-1, 3, 3)).sum((1, 2))
    obj_ref = (ref * x[sl].cpu().double().reshape(-1, 3, 3)).sum((1, 2))
    assert float((obj_ref - obj).max()) <= 1e-5


def test_full_size_tiles_are_each_processed_once(rr):
    """The SVD kernels draw their tiles from a global counter: a full-size, ragged batch must give bit for bit what
    the same records give when they are processed in two separate launches (no tile dropped, none done twice), on
    repeated launches (the scheduler slot is re-armed), and the fused step must agree with the separate kernels."""
    n = (1 << 24) + 77
    torch.manual_seed(1)
    x = torch.randn(n, 9, device="cuda")
    r = rr.symmetric_orthogonalization(x)
    h = (1 << 23) + 5
    assert torch.equal(r[:h], rr.symmetric_orthogonalization(x[:h]))
    assert torch.equal(r[h:], rr.symmetric_orthogonalization(x[h:]))
    for _ in range(3):
        assert torch.equal(r, rr.symmetric_orthogonalization(x))
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda"))
    loss, g = rr.svd_frobenius_loss_and_grad(x, rt)
    loss_sep = rr.loss_frobenius(r, rt)
    assert abs(loss.item() - loss_sep.item()) <= 1e-6
    x2 = x.clone().requires_grad_(True)
    rr.svd_frobenius_loss(x2, rt, single_pass=False).backward()
    scale = x2.grad.abs().amax(dim=1, keepdim=True)
    assert float(((g - x2.grad).abs() / scale).max()) <= 1e-4
    # halves of the fused step: sums of the two half-batches reproduce the whole (loss is a mean over n)
    l1, g1 = rr.svd_frobenius_loss_and_grad(x[:h], rt[:h])
    l2, g2 = rr.svd_frobenius_loss_and_grad(x[h:], rt[h:])
    assert abs((l1.item() * h + l2.item() * (n - h)) / n - loss.item()) <= 1e-6
    # (the 1/n factor is folded into the seed ahead of the 3x3 solve, so the roundings differ on ill-conditioned rows)
    for part, whole, w in ((g1, g[:h], h / n), (g2, g[h:], (n - h) / n)):
        rel = (part * w - whole).abs() / whole.abs().amax(dim=1, keepdim=True).clamp_min(1e-30)
        assert float(rel.max()) <= 1e-4


def test_full_size_directional_derivative(rr):
    """2^24 matrices: <dL/dx, V> against a central difference of the loss along V = the row-normalised gradient (a
    random V gives a derivative that cancels to ~n^-1/2, below the float32 resolution of the returned loss)."""
    n = 1 << 24
    torch.manual_seed(2)
    x = torch.randn(n, 9, device="cuda")
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda"))
    _, g = rr.svd_frobenius_loss_and_grad(x, rt)
    v = g / g.norm(dim=1, keepdim=True).clamp_min(1e-30)
    lhs = (g.double() * v.double()).sum().item()
    eps = 1e-3
    lp = rr.svd_frobenius_loss(x + eps * v, rt).item()      # means accumulated in float64 inside the kernel
    lm = rr.svd_frobenius_loss(x - eps * v, rt).item()
    rhs = (lp - lm) / (2 * eps)
    assert abs(lhs - rhs) <= 2e-2 * max(abs(rhs), 1e-3), (lhs, rhs)


def test_config1_svd_plus_frobenius(rr):
    """BASELINE config 1: 4096 random matrices, forward + Frobenius loss, vs the reference formulas on the CPU."""
    torch.manual_seed(0)
    x = torch.randn(4096, 9)
    r_true = O.symmetric_orthogonalization(torch.randn(4096, 9).double()).float()
    xr = x.double().requires_grad_(True)
    loss_ref = O.loss_frobenius(O.symmetric_orthogonalization(xr), r_true.double())
    loss_ref.backward()
    xd = x.cuda().requires_grad_(True)
    loss = rr.loss_frobenius(rr.symmetric_orthogonalization(xd), r_true.cuda())
    loss.backward()
    assert loss.dtype == torch.float32 and loss.dim() == 0
    assert abs(loss.item() - loss_ref.item()) <= 1e-6
    cond = O.svd_orth_conditioning(x).numpy()
    rel = _relrows(xd.grad.cpu().numpy(), xr.grad.numpy())
    assert rel[cond > 0.1].max() <= 1e-4
    # fused op: same loss, same gradient
    xf = x.cuda().requires_grad_(True)
    lf = rr.svd_frobenius_loss(xf, r_true.cuda())
    lf.backward()
    assert abs(lf.item() - loss_ref.item()) <= 1e-6
    assert _relrows(xf.grad.cpu().numpy(), xr.grad.numpy())[cond > 0.1].max() <= 1e-4


@pytest.mark.parametrize("n", [1, 63, 511, 512, 513, 4096 + 77])
def test_single_pass_step_matches_two_pass_and_reference(rr, n):
    """svd_frobenius_loss in its default single-pass form (one SVD per matrix, backward = one scaling pass), against
    the recompute form and the float64 reference; upstream gradient != 1 to exercise the device-scalar scaling."""
    torch.manual_seed(n)
    x = torch.randn(n, 9)
    r_true = O.symmetric_orthogonalization(torch.randn(n, 9).double()).float()
    xr = x.double().requires_grad_(True)
    loss_ref = O.loss_frobenius(O.symmetric_orthogonalization(xr), r_true.double())
    (loss_ref * 2.5).backward()
    grads = []
    for single in (True, False):
        xd = x.cuda().requires_grad_(True)
        loss = rr.svd_frobenius_loss(xd, r_true.cuda(), single_pass=single)
        assert loss.dtype == torch.float32 and loss.dim() == 0 and loss.requires_grad
        (loss * 2.5).backward()
        assert abs(loss.item() - loss_ref.item()) <= 2e-6
        grads.append(xd.grad.cpu().numpy())
    cond = O.svd_orth_conditioning(x).numpy()
    ok = cond > 0.1
    if ok.any():
        assert _relrows(grads[0], xr.grad.numpy())[ok].max() <= 1e-4
        assert _relrows(grads[1], xr.grad.numpy())[ok].max() <= 1e-4
    # the two forms run the same per-matrix arithmetic
    assert np.abs(grads[0] - grads[1]).max() <= 1e-6 * max(1.0, np.abs(grads[1]).max())
    # explicit loss-and-gradient call, no autograd
    l2, g2 = rr.svd_frobenius_loss_and_grad(x.cuda(), r_true.cuda())
    assert abs(l2.item() - loss_ref.item()) <= 2e-6 and g2.shape == (n, 9) and not g2.requires_grad
    assert np.abs(g2.cpu().numpy() * 2.5 - grads[0]).max() <= 1e-6 * max(1.0, np.abs(grads[0]).max())
    # no gradient requested: the forward-only kernel serves the call
    assert not rr.svd_frobenius_loss(x.cuda(), r_true.cuda()).requires_grad


def test_single_pass_step_backward_details(rr, ops):
    torch.manual_seed(5)
    x = torch.randn(1000, 9, device="cuda")
    rt = rr.symmetric_orthogonalization(torch.randn(1000, 9, device="cuda"))
    # upstream gradient exactly 1 (loss.backward()): the stored gradient is handed over untouched
    x1 = x.clone().requires_grad_(True)
    rr.svd_frobenius_loss(x1, rt).backward()
    _, g = rr.svd_frobenius_loss_and_grad(x, rt)
    assert torch.equal(x1.grad, g)
    # gradient accumulation over two steps adds up
    rr.svd_frobenius_loss(x1, rt).backward()
    assert torch.allclose(x1.grad, 2 * g, rtol=1e-6, atol=0)
    # a second traversal of the same graph finds the stored gradient handed over and takes the recompute kernel
    x2 = x.clone().requires_grad_(True)
    loss = rr.svd_frobenius_loss(x2, rt)
    loss.backward(retain_graph=True)
    loss.backward()
    assert torch.allclose(x2.grad, 2 * g, rtol=1e-5, atol=1e-9)
    with pytest.raises(RuntimeError):                      # ... and without retain_graph torch refuses, as for any graph
        loss.backward()
    # the registered (dispatcher) op is functional: its dx output is never written again, two traversals are exact
    x4 = x.clone().requires_grad_(True)
    l4, dx4 = ops.svd_orth_frobenius_step.op(x4, rt)
    keep = dx4.clone()
    (l4 * 3.0).backward(retain_graph=True)
    (l4 * 3.0).backward()
    assert torch.equal(dx4, keep) and torch.allclose(x4.grad, 6 * g, rtol=1e-6, atol=0)
    x3 = x.clone().requires_grad_(True)
    loss = rr.svd_frobenius_loss(x3, rt, single_pass=False)
    loss.backward(retain_graph=True)
    loss.backward()
    assert torch.allclose(x3.grad, 2 * g, rtol=1e-5, atol=1e-9)
    # the in-place scaling op itself, ragged length
    v = torch.randn(1003, device="cuda")
    w = v.clone()
    ops.scale_by_scalar_(w, torch.tensor(1.0, device="cuda"))
    assert torch.equal(w, v)
    ops.scale_by_scalar_(w, torch.tensor(-0.5, device="cuda"))
    assert torch.equal(w, v * -0.5)


@pytest.mark.parametrize("k", [-60, -17, -16, -15, -3, 29, 30, 31, 60])
def test_scale_sweep_across_the_magnitude_window(rr, k):
    """R must not depend on a power-of-two scale of the input and dM must scale with its inverse, from 2^-60 to 2^60
    (the exact power-of-two prescale of the kernels), also inside a batch that mixes scales."""
    torch.manual_seed(3)
    n = 4096 + 5
    x = torch.randn(n, 9, device="cuda")
    g = torch.randn(n, 3, 3, device="cuda")
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda"))
    sc = 2.0 ** k
    x0, xs = x.clone().requires_grad_(True), (x * sc).requires_grad_(True)
    r0, rs = rr.symmetric_orthogonalization(x0), rr.symmetric_orthogonalization(xs)
    assert (r0 - rs).abs().max().item() <= 1e-6
    r0.backward(g); rs.backward(g)
    scale = x0.grad.abs().amax(dim=1, keepdim=True)
    assert ((xs.grad * sc - x0.grad).abs() / scale).max().item() <= 1e-5
    # the fused step too (loss identical, gradient scaled); mixed batch: half the rows scaled
    mix = x.clone()
    mix[::2] *= sc
    l0, g0 = rr.svd_frobenius_loss_and_grad(x, rt)
    l1, g1 = rr.svd_frobenius_loss_and_grad(mix, rt)
    assert abs(l0.item() - l1.item()) <= 1e-6
    g1 = g1.clone()
    g1[::2] *= sc
    assert ((g1 - g0).abs() / g0.abs().amax(dim=1, keepdim=True)).max().item() <= 1e-5


def test_degenerate_inputs_are_finite_rotations(rr):
    cases = torch.tensor([
        [0.0] * 9, [1, 0, 0, 0, 1, 0, 0, 0, 1], [-1, 0, 0, 0, -1, 0, 0, 0, -1], [1, 0, 0, 0, 1, 0, 0, 0, -1],
        [3, 0, 0, 0, 0, 0, 0, 0, 0], [2, 0, 0, 0, 1, 0, 0, 0, 0], [1, 1, 1, 1, 1, 1, 1, 1, 1],
        [1e-30 * k for k in range(1, 10)], [1e30, 2e30, 3e30, 4e30, 5e30, 6e30, 7e30, 8e30, 1e31]], dtype=torch.float32)
    x = cases.cuda().requires_grad_(True)
    r = rr.symmetric_orthogonalization(x)
    r.sum().backward()
    rd = r.detach().cpu().double()
    assert torch.isfinite(rd).all() and torch.isfinite(x.grad).all()
    assert (rd @ rd.transpose(1, 2) - torch.eye(3, dtype=torch.float64)).abs().max() <= 1e-5
    assert (torch.linalg.det(rd) - 1).abs().max() <= 1e-5


def test_opcheck(ops):
    x = torch.randn(33, 9, device="cuda", requires_grad=True)
    torch.library.opcheck(ops.svd_orth.op, (x,), test_utils=("test_schema", "test_faketensor", "test_autograd_registration"))
    a = torch.randn(33, 3, 3, device="cuda", requires_grad=True)
    b = torch.randn(33, 3, 3, device="cuda")
    torch.library.opcheck(ops.frobenius_mean.op, (a, b),
                          test_utils=("test_schema", "test_faketensor", "test_autograd_registration"))
    torch.library.opcheck(ops.svd_orth_frobenius_step.op, (x, b),
                          test_utils=("test_schema", "test_faketensor", "test_autograd_registration"))
    torch.library.opcheck(ops.ortho6d.op, (torch.randn(33, 6, device="cuda", requires_grad=True),),
                          test_utils=("test_schema", "test_faketensor", "test_autograd_registration"))


def test_eager_fast_path_equals_dispatcher_route(ops, rr):
    """ops.X(...) calls the op body directly (plain autograd.Function) for ordinary CUDA tensors; ops.X.op(...) is the
    registered pe_b200::X through the dispatcher.  Same kernels, same autograd formulas: results are bit-identical."""
    g = torch.Generator(device="cuda").manual_seed(7)
    rt = rr.symmetric_orthogonalization(torch.randn(257, 9, device="cuda", generator=g)).detach()
    up = torch.randn(257, 3, 3, device="cuda", generator=g)
    for name, w in (("svd_orth", 9), ("ortho6d", 6), ("quat", 4), ("euler", 3)):
        x = torch.randn(257, w, device="cuda", generator=g)
        res = []
        for route in (getattr(ops, name), getattr(ops, name).op):
            xi = x.clone().requires_grad_(True)
            r = route(xi)
            r.backward(up)
            res.append((r.detach(), xi.grad))
        assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])
        assert type(getattr(ops, name)(x).grad_fn) is type(None)
        assert "pe_b200_" + name in type(getattr(ops, name)(x.clone().requires_grad_(True)).grad_fn).__name__
    x = torch.randn(257, 9, device="cuda", generator=g)
    out = []
    for route in (ops.svd_orth_frobenius_step, ops.svd_orth_frobenius_step.op):
        xi = x.clone().requires_grad_(True)
        loss, _ = route(xi, rt)
        (loss * 3.0).backward()
        out.append((loss.detach(), xi.grad))
    # the loss sum is accumulated with float64 atomics in an order that varies from launch to launch
    assert abs(float(out[0][0]) - float(out[1][0])) <= 1e-6 and torch.equal(out[0][1], out[1][1])
    a = torch.randn(64, 3, 3, device="cuda", generator=g)
    ga = []
    for route in (ops.frobenius_mean, ops.frobenius_mean.op):
        ai = a.clone().requires_grad_(True)
        route(ai, rt[:64]).backward()
        ga.append(ai.grad)
    assert torch.equal(ga[0], ga[1])
    # double backward is refused on the fast path as on the registered one
    xi = x.clone().requires_grad_(True)
    (gx,) = torch.autograd.grad(ops.svd_orth(xi).sum(), xi, create_graph=True)
    with pytest.raises(RuntimeError):
        gx.sum().backward()


def test_registered_ops_trace_under_torch_compile(rr):
    """torch.compile sees pe_b200::* through register_fake / register_autograd (the wrapper steps aside while
    compiling); aot_eager runs the traced graph with the same kernels."""
    g = torch.Generator(device="cuda").manual_seed(11)
    x = torch.randn(130, 9, device="cuda", generator=g)
    rt = rr.symmetric_orthogonalization(torch.randn(130, 9, device="cuda", generator=g)).detach()

    def step(x, rt):
        return rr.loss_frobenius(rr.symmetric_orthogonalization(x), rt)

    xe = x.clone().requires_grad_(True)
    le = step(xe, rt)
    le.backward()
    xc = x.clone().requires_grad_(True)
    lc = torch.compile(step, backend="aot_eager", fullgraph=True)(xc, rt)
    lc.backward()
    # eager: loss_frobenius recognises the svd_orth output and runs the single-pass step; traced: the two registered ops
    # (svd_orth + frobenius_mean with their own backward kernels).  Same mathematics, different kernels.
    assert abs(float(le.detach()) - float(lc.detach())) <= 1e-6
    assert (xe.grad - xc.grad).abs().max().item() <= 2e-6 * xe.grad.abs().max().item()


@pytest.mark.parametrize("n", [1, 2, 3, 512, 1001, 70001])
def test_strided_view_runs_the_packed_direct_kernels(ops, rr, n):
    """A stride-12 view goes through the non-TMA kernels, which keep two matrices per thread (records i and
    i + ceil(n/2)): forward, recompute backward and the single-pass step must equal the dense TMA route bit for bit,
    for odd and even batch sizes."""
    g = torch.Generator(device="cuda").manual_seed(n)
    wide = torch.randn(n, 12, device="cuda", generator=g)
    view, dense = wide[:, :9], wide[:, :9].contiguous()
    up = torch.randn(n, 3, 3, device="cuda", generator=g)
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda", generator=g))
    r_v, d_v = ops.svd_orth_det(view)
    r_d, d_d = ops.svd_orth_det(dense)
    assert torch.equal(r_v, r_d) and torch.equal(d_v, d_d)
    assert torch.equal(ops.svd_orth(view), r_d)
    assert torch.equal(ops.svd_orth_bwd_recompute(view, up), ops.svd_orth_bwd_recompute(dense, up))
    l_v, g_v = ops.svd_orth_frobenius_step(view, rt)
    l_d, g_d = ops.svd_orth_frobenius_step(dense, rt)
    assert torch.equal(g_v, g_d) and abs(float(l_v) - float(l_d)) <= 1e-6
    assert abs(float(rr.map_metric("SVD", view, rt)) - float(rr.map_metric("SVD", dense, rt))) <= 1e-9 * 180
