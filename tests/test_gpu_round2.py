"""Round-2 additions on the GPU: the step kernel with its fused metric, the unchanged call surface routed onto it, the
launch-slot ownership rules (streams, CUDA graphs, caller scratch), the wide TMA stage for 12-word rows and the small
gaps against the reference surface.  Tolerances are BASELINE.json's: outputs 1e-5, mean angle error 1e-4 degrees
(asserted far tighter where the arithmetic allows), gradients 1e-4 relative."""
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest
import torch

from oracle import reference_port as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
T = lambda a: torch.from_numpy(np.asarray(a))


@pytest.fixture(scope="module")
def rr():
    import poseestimation_b200.rotation_representation as m
    return m


@pytest.fixture(scope="module")
def ops():
    from poseestimation_b200 import ops as m
    return m


@pytest.fixture(scope="module")
def abi():
    from poseestimation_b200 import _lib
    return _lib


# ---------------------------------------------------------------------------------------------------------------------
# the training step with its metric
def test_step_with_angle_vs_reference_golden(rr, golden):
    """(loss, mean angle error) of ONE kernel against the unmodified reference's step + metric on 4096 matrices
    (Comparison/main.py:59-70): the float64 evaluation and the scripts' own fp32 pipeline."""
    z = golden("round2")
    x, rt = T(z["step_x"]).cuda(), T(z["step_r_true"]).cuda()
    xd = x.clone().requires_grad_(True)
    loss, ang = rr.svd_frobenius_loss(xd, rt, with_angle=True)
    assert ang.dtype == torch.float64 and ang.dim() == 0 and loss.dtype == torch.float32
    assert abs(loss.item() - float(z["step_loss_f64"])) <= 1e-6
    assert abs(ang.item() - z["step_angle_f64"].mean()) <= 1e-4                # north_star
    assert abs(ang.item() - z["step_angle_fp32_pipeline"].mean()) <= 1e-4
    # the two-call route on the same device: the step forms the same fp32 R in registers, its trace error-free in
    # float32, and finishes with a float32 arcsine: <= 2e-5 degrees per record, independent errors
    unfused = rr.mean_angle_error(rr.symmetric_orthogonalization(x), rt)
    assert abs(ang.item() - unfused.item()) <= 2e-6
    loss.backward()
    xe = x.clone().requires_grad_(True)
    rr.svd_frobenius_loss(xe, rt).backward()
    assert torch.equal(xd.grad, xe.grad)                                        # the metric does not touch the gradient


@pytest.mark.parametrize("n", [1, 2, 511, 513, 1 << 20])
def test_step_with_angle_vs_oracle_and_unfused(rr, ops, n):
    g = torch.Generator().manual_seed(n)
    x = torch.randn(n, 9, generator=g)
    rt = O.symmetric_orthogonalization(torch.randn(n, 9, generator=g).double()).float()
    xd, rtd = x.cuda(), rt.cuda()
    loss, ang, flag, dx = ops.svd_orth_frobenius_step_metrics(xd, rtd)
    assert int(flag.item()) == 0
    want_ang = O.angle_error(O.symmetric_orthogonalization(x.double()).float(), rt).mean().item()
    assert abs(ang.item() - want_ang) <= 1e-4
    l2, dx2 = ops.svd_orth_frobenius_step(xd, rtd)
    assert torch.equal(dx, dx2) and abs(loss.item() - l2.item()) <= 1e-6
    assert abs(ang.item() - rr.mean_angle_error(rr.symmetric_orthogonalization(xd), rtd).item()) <= 2e-5 / max(1.0, n ** 0.5 / 8)
    # strided (B, 12) rows and an unaligned base: the wide TMA stage and the direct kernel give the same numbers
    wide = torch.zeros(n, 12, device="cuda")
    wide[:, :9] = xd
    _, a_w, _, dx_w = ops.svd_orth_frobenius_step_metrics(wide[:, :9], rtd)
    assert torch.equal(dx_w, dx) and abs(a_w.item() - ang.item()) <= 1e-9
    odd = torch.zeros(n * 9 + 1, device="cuda")[1:].view(n, 9)
    odd.copy_(xd)
    _, a_o, _, dx_o = ops.svd_orth_frobenius_step_metrics(odd, rtd)
    assert torch.equal(dx_o, dx) and abs(a_o.item() - ang.item()) <= 1e-9


def test_step_with_angle_raises_like_angle_error(rr):
    x = torch.randn(64, 9, device="cuda")
    bad = 3.0 * torch.eye(3, device="cuda").repeat(64, 1, 1)                    # cos = (tr - 1)/2 far outside [-1.1, 1.1]
    with pytest.raises(ValueError, match="angle out of range"):
        rr.svd_frobenius_loss(x, bad, with_angle=True)
    loss, ang = rr.svd_frobenius_loss(x, bad, with_angle=True, check=False)
    assert torch.isfinite(loss) and torch.isfinite(ang)
    e = torch.zeros(0, 9, device="cuda")
    loss, ang = rr.svd_frobenius_loss(e, e, with_angle=True)
    assert torch.isnan(loss) and torch.isnan(ang)


def test_c_abi_step_metrics_vector_and_caller_scratch(abi, rr):
    """sums = [sum |.|_F, sum theta, flag] (the all-reduced vector); with caller-owned scratch the launch uses
    no library memory and leaves the scratch zero."""
    n = 70001
    x = torch.randn(n, 9, device="cuda")
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda")).reshape(n, 9)
    st = torch.cuda.current_stream().cuda_stream
    dx = torch.empty(n, 9, device="cuda")
    out = []
    for scratch in (None, torch.zeros(8, dtype=torch.int32, device="cuda")):
        sums = torch.full((3,), -1.0, dtype=torch.float64, device="cuda")
        mean32 = torch.empty((), device="cuda")
        mean64 = torch.empty((), dtype=torch.float64, device="cuda")
        flag = torch.full((1,), 7, dtype=torch.int32, device="cuda")
        for _ in range(3):                                                      # re-armed after every launch
            abi.check(abi.lib.pe_svd_orth_frobenius_step_metrics(
                x.data_ptr(), 9, rt.data_ptr(), 1.0 / n, dx.data_ptr(), sums.data_ptr(), mean32.data_ptr(),
                mean64.data_ptr(), flag.data_ptr(), None if scratch is None else scratch.data_ptr(), n, st), "step")
        torch.cuda.synchronize()
        if scratch is not None:
            assert int(scratch.abs().sum()) == 0
        assert int(flag.item()) == 0 and sums[2].item() == 0.0
        assert abs(mean32.item() - sums[0].item() / n) <= 1e-6 and abs(mean64.item() - sums[1].item() / n) <= 1e-12
        out.append(sums.clone())
    assert torch.allclose(out[0], out[1], rtol=1e-12, atol=0)
    assert abs(out[0][1].item() / n - rr.mean_angle_error(rr.symmetric_orthogonalization(x), rt).item()) <= 1e-6
    # bad targets: the flag reaches both forms
    sums = torch.zeros(3, dtype=torch.float64, device="cuda")
    bad = (3.0 * torch.eye(3, device="cuda")).repeat(n, 1, 1).reshape(n, 9).contiguous()
    abi.check(abi.lib.pe_svd_orth_frobenius_step_metrics(x.data_ptr(), 9, bad.data_ptr(), 1.0 / n, dx.data_ptr(),
                                                         sums.data_ptr(), None, None, None, None, n, st), "step")
    assert sums[2].item() == 1.0


# ---------------------------------------------------------------------------------------------------------------------
# the unchanged call surface: loss_frobenius(symmetric_orthogonalization(x), R)
def _fused(loss):
    """the loss came from the single-pass step: Python node `_FusedFromSvdOutputBackward` or C++ node `FusedFromSvdOutput`"""
    return "FusedFromSvdOutput" in repr(loss)


@pytest.fixture(params=["cpp_nodes", "python_nodes"])
def route(request, ops, monkeypatch):
    """the unchanged surface runs through C++ autograd nodes (csrc/pe_torch.cpp) when _pe_torch.so is built and through
    the Python nodes of ops.py otherwise: every drop-in test runs on both"""
    if request.param == "cpp_nodes":
        if ops._C is None:
            pytest.skip("_pe_torch.so not built")
    else:
        monkeypatch.setattr(ops, "_C", None)
    return request.param


def _dropin(rr, x, rt, swap=False):
    xd = x.clone().requires_grad_(True)
    r = rr.symmetric_orthogonalization(xd)
    loss = rr.loss_frobenius(rt, r) if swap else rr.loss_frobenius(r, rt)
    return xd, r, loss


@pytest.mark.parametrize("swap", [False, True])
def test_dropin_route_is_the_single_pass_step(rr, abi, swap, route):
    """Three launches (svd_orth fwd, step, the device-side no-op scaling) instead of four heavy ones, and the gradient is
    bit for bit that of svd_frobenius_loss.  UPNA/main.py:57 passes (R, out): both argument orders."""
    n = 5000
    x = torch.randn(n, 9, device="cuda")
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda"))
    torch.cuda.synchronize()
    before = abi.launch_count()
    xd, r, loss = _dropin(rr, x, rt, swap)
    assert _fused(loss)
    loss.backward()
    assert abi.launch_count() - before == 3
    xe = x.clone().requires_grad_(True)
    le = rr.svd_frobenius_loss(xe, rt)
    le.backward()
    assert torch.equal(xd.grad, xe.grad) and abs(loss.item() - le.item()) <= 1e-7
    # and it is the right gradient: float64 autograd through the oracle
    xr = x.cpu().double().requires_grad_(True)
    O.loss_frobenius(O.symmetric_orthogonalization(xr), rt.cpu().double()).backward()
    cond = O.svd_orth_conditioning(x.cpu()).numpy()
    gr = xr.grad.numpy()
    scale = np.abs(gr).max(1)
    rel = np.abs(xd.grad.cpu().double().numpy() - gr).max(1) / np.maximum(scale, 0.1 * np.median(scale))
    assert rel[cond > 0.1].max() <= 1e-4


def test_dropin_route_upstream_gradient_and_second_consumer(rr, route):
    """loss weighting (upstream != 1) and a second differentiable consumer of R: gradients add up exactly as on the
    unfused route."""
    n = 3001
    x = torch.randn(n, 9, device="cuda")
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda"))
    w = torch.randn(n, 3, 3, device="cuda")

    def run(fused):
        xd = x.clone().requires_grad_(True)
        r = rr.symmetric_orthogonalization(xd)
        if not fused:
            r = r * 1.0                                     # a new tensor: not the tagged svd_orth output any more
        total = 0.25 * rr.loss_frobenius(r, rt) + (r * w).sum() * 1e-3 + 2.0 * rr.loss_frobenius(r, rt.roll(1, 0))
        total.backward()
        return total.item(), xd.grad

    lf, gf = run(True)
    lu, gu = run(False)
    assert abs(lf - lu) <= 1e-6 * max(1.0, abs(lu))
    assert (gf - gu).abs().max().item() <= 1e-6 * gu.abs().max().item() + 1e-9


def test_dropin_route_retain_graph_and_fallbacks(rr, abi, route):
    n = 2000
    x = torch.randn(n, 9, device="cuda")
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda"))
    # retain_graph: the second traversal takes the unfused kernels and gives the same gradient
    xd, r, loss = _dropin(rr, x, rt)
    loss.backward(retain_graph=True)
    g1 = xd.grad.clone()
    xd.grad = None
    loss.backward()
    assert (xd.grad - g1).abs().max().item() <= 2e-6 * g1.abs().max().item()
    # not fusable: R modified in place, R retained, hooks on R, a view of R, no_grad, a target that needs a gradient
    xd = x.clone().requires_grad_(True)
    r = rr.symmetric_orthogonalization(xd)
    r.retain_grad()
    loss = rr.loss_frobenius(r, rt)
    assert not _fused(loss)
    loss.backward()
    assert r.grad is not None and (xd.grad - g1).abs().max().item() <= 2e-6 * g1.abs().max().item()
    xd = x.clone().requires_grad_(True)
    r = rr.symmetric_orthogonalization(xd)
    seen = []
    r.register_hook(lambda g: seen.append(g.clone()))
    loss = rr.loss_frobenius(r, rt)
    loss.backward()
    assert len(seen) == 1 and seen[0].abs().max().item() > 0       # the hook saw the true dL/dR, not the zero marker
    xd = x.clone().requires_grad_(True)
    loss = rr.loss_frobenius(rr.symmetric_orthogonalization(xd).view(-1, 3, 3), rt)
    assert not _fused(loss)
    with torch.no_grad():
        assert rr.loss_frobenius(rr.symmetric_orthogonalization(x), rt).grad_fn is None
    xd = x.clone().requires_grad_(True)
    rtg = rt.clone().requires_grad_(True)
    loss = rr.loss_frobenius(rr.symmetric_orthogonalization(xd), rtg)
    loss.backward()
    assert rtg.grad is not None and (xd.grad - g1).abs().max().item() <= 2e-6 * g1.abs().max().item()
    # the loss is never backpropagated, R is: the parked gradient must not leak into that backward
    xd = x.clone().requires_grad_(True)
    r = rr.symmetric_orthogonalization(xd)
    _ = rr.loss_frobenius(r, rt)
    (r * rt).sum().backward()
    xe = x.clone().requires_grad_(True)
    (rr.symmetric_orthogonalization(xe) * rt).sum().backward()
    assert torch.equal(xd.grad, xe.grad)


def test_star_import_like_the_reference_scripts():
    """`sys.path` first, `from rotation_representation import *` (Comparison/models.py:11, UPNA/main.py:23), one training
    step and its metric with the reference's names - in a fresh interpreter, the way a script picks the module up."""
    code = textwrap.dedent(f"""
        import sys
        sys.path.insert(0, {os.path.join(ROOT, 'poseestimation_b200')!r})
        from rotation_representation import *
        assert torch.__name__ == 'torch' and np.__name__ == 'numpy' and nn is torch.nn     # the star-import hands these on
        dim, fn = transform_output['SVD']
        out = torch.randn(64, dim, device='cuda', requires_grad=True)
        R = symmetric_orthogonalization(torch.randn(64, 9, device='cuda'))
        outR = fn(out)
        loss = loss_frobenius(outR, R)
        loss.backward()
        mae = angle_error(t_R1=outR.detach(), t_R2=R).mean().item()
        assert out.grad.shape == (64, 9) and 0 < mae < 180 and loss.item() > 0
        assert compute_rotation_matrix_from_ortho6d(torch.randn(5, 6, device='cuda')).shape == (5, 3, 3)
        print('star-import ok', 'FusedFromSvdOutput' in repr(loss))
    """)
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600, cwd="/")
    assert res.returncode == 0, res.stdout + res.stderr
    assert "star-import ok True" in res.stdout


def test_cpp_autograd_node_equals_the_python_node(rr, ops):
    """csrc/pe_torch.cpp (the host-latency route of svd_frobenius_loss) against ops._FusedStep: the same C-ABI calls, so
    loss, mean angle, flag and gradient are bit-identical - dense, on the (B, 12) view, with an upstream gradient other
    than 1, over a second traversal (retain_graph), and for an empty batch."""
    if ops._C is None:
        pytest.skip("_pe_torch.so not built")
    g = torch.Generator(device="cuda").manual_seed(21)
    n = 3001
    wide = torch.randn(n, 12, device="cuda", generator=g)
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda", generator=g))
    for strided in (False, True):
        for with_angle in (False, True):
            # leaves: a dense (n, 9) tensor, or the (n, 12) output whose [:, :9] view is read in place (wide TMA stage)
            lc, lp = ((wide if strided else wide[:, :9].contiguous()).detach().clone().requires_grad_(True) for _ in range(2))
            xc, xp = (lc[:, :9], lp[:, :9]) if strided else (lc, lp)
            assert xc.stride(0) == (12 if strided else 9)
            out_c = ops._C.fused_step_loss(xc, rt, with_angle)
            metric = [] if with_angle else None
            loss_p = ops._FusedStep.apply(xp, rt, metric)
            assert out_c[0].item() == loss_p.item() and out_c[0].dtype == torch.float32 and out_c[0].dim() == 0
            if with_angle:
                assert out_c[1].item() == metric[0].item() and out_c[1].dtype == torch.float64
                assert int(out_c[2].item()) == int(metric[1].item()) == 0 and not out_c[1].requires_grad
            (out_c[0] * 2.5).backward(retain_graph=True)
            (loss_p * 2.5).backward(retain_graph=True)
            assert torch.equal(lc.grad, lp.grad) and lc.grad.abs().max().item() > 0
            out_c[0].backward()                              # second traversal: the recompute kernel on both routes
            loss_p.backward()
            assert torch.equal(lc.grad, lp.grad)
            with pytest.raises(RuntimeError):
                out_c[0].backward()
    e = torch.zeros(0, 9, device="cuda", requires_grad=True)
    out = ops._C.fused_step_loss(e, torch.zeros(0, 3, 3, device="cuda"), True)
    assert torch.isnan(out[0]) and torch.isnan(out[1])
    out[0].backward()
    assert e.grad.shape == (0, 9)
    with pytest.raises(RuntimeError, match="batch mismatch"):
        ops._C.fused_step_loss(wide[:, :9].contiguous().requires_grad_(True), rt[:5], False)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops._C.fused_step_loss(torch.randn(4, 9, requires_grad=True), rt[:4], False)
    # and it is the route the public function takes
    x = wide[:, :9].contiguous().requires_grad_(True)
    loss = rr.svd_frobenius_loss(x, rt)
    assert type(loss.grad_fn).__name__ == "CppFunction" and "FusedStep" in repr(loss)


def _graph_names(t):
    """names of the autograd nodes behind a tensor"""
    seen, todo, names = set(), [t.grad_fn], set()
    while todo:
        fn = todo.pop()
        if fn is None or fn in seen:
            continue
        seen.add(fn)
        names.add(fn.name())
        todo.extend(f for f, _ in fn.next_functions)
    return " ".join(sorted(names))


@pytest.mark.parametrize("key,width", [("6D", 6), ("Quat", 4), ("Euler", 3), ("5D", 5), ("3D", 3), ("3D-eps", 3)])
def test_cpp_map_and_loss_nodes_equal_the_python_route(rr, ops, monkeypatch, key, width):
    """The maps, loss_frobenius on a generic pair and geodesic() as C++ autograd nodes (csrc/pe_torch.cpp) against the
    registered Python ops: the same entry points of the C ABI, so values and gradients are bit-identical - for every
    reduction, a strided input, an upstream gradient other than one, gradients towards either argument, an empty batch."""
    if ops._C is None:
        pytest.skip("_pe_torch.so not built")
    fns = {"6D": rr.compute_rotation_matrix_from_ortho6d, "Quat": rr.compute_rotation_matrix_from_quaternion,
           "Euler": rr.compute_rotation_matrix_from_euler, "5D": rr.compute_rotation_matrix_from_ortho5d,
           "3D": rr.so3_exp_map, "3D-eps": lambda v: rr.so3_exp_map(v, 1e-2)}
    fn = fns[key]
    g = torch.Generator(device="cuda").manual_seed(width)
    n = 2049
    base = torch.randn(n, width + 3, device="cuda", generator=g)
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda", generator=g))
    w = torch.randn(n, device="cuda", generator=g)

    def run(strided, loss_kind, cpp):
        if not cpp:
            monkeypatch.setattr(ops, "_C", None)
        leaf = (base if strided else base[:, :width].contiguous()).detach().clone().requires_grad_(True)
        x = leaf[:, :width] if strided else leaf
        target = rt.detach().clone().requires_grad_(True)      # gradients towards BOTH arguments
        r = fn(x)
        if loss_kind == "frobenius":
            loss = rr.loss_frobenius(r, target)
        elif loss_kind == "none":
            loss = (rr.geodesic(r, target, "none") * w).sum()
        elif loss_kind == "a6":
            loss = (rr.compute_geodesic_distance_from_two_matrices(r, target) * w).sum()
        else:
            loss = rr.geodesic(r, target, loss_kind)
        (loss * 1.75).backward()
        monkeypatch.undo()
        return r.detach(), loss.detach(), leaf.grad, target.grad, _graph_names(r), _graph_names(loss)

    for strided in (False, True):
        for kind in ("frobenius", "mean", "sum", "none", "a6"):
            rc, lc, gc, tc, rep_r, rep_l = run(strided, kind, True)
            rp, lp, gp, tp, rep_rp, _ = run(strided, kind, False)
            assert "MapNode" in rep_r and "MapNode" not in rep_rp
            if kind in ("frobenius", "mean", "sum"):
                assert ("FrobeniusMean" if kind == "frobenius" else "Geodesic") in rep_l
                assert lc.dim() == 0 and lc.dtype == torch.float32
            assert torch.equal(rc, rp) and lc.item() == lp.item()
            # bitwise (int32 views): the clamp of a6 is +-1 exactly, where acos' has no finite derivative - a record that
            # lands there gives inf / NaN on both routes, as it does in the reference
            bits = lambda t: t.contiguous().view(torch.int32)
            assert torch.equal(bits(gc), bits(gp)) and torch.equal(bits(tc), bits(tp)) and gc.nan_to_num().abs().max().item() > 0
    # no gradient asked for: the registered op (no node), same numbers
    with_grad = run(True, "mean", True)[0]
    with torch.no_grad():
        assert torch.equal(fn(base[:, :width]), with_grad)
    assert fn(base[:, :width].contiguous()).grad_fn is None
    # an empty batch
    e = torch.zeros(0, width, device="cuda", requires_grad=True)
    re_ = fn(e)
    assert re_.shape == (0, 3, 3)
    t0 = torch.zeros(0, 3, 3, device="cuda")
    assert torch.isnan(rr.loss_frobenius(re_, t0)) and torch.isnan(rr.geodesic(re_, t0, "mean"))
    s0 = rr.geodesic(re_, t0, "sum")
    assert s0.item() == 0.0
    s0.backward()
    assert e.grad.shape == (0, width)
    # argument errors are the Python route's (ValueError / TypeError), whichever nodes are built
    x = base[:, :width].contiguous().requires_grad_(True)
    with pytest.raises(ValueError, match="batch mismatch"):
        rr.loss_frobenius(fn(x), rt[:7])
    with pytest.raises(TypeError):
        fn(x.double())


def test_autograd_routes_on_a_side_stream_and_from_threads(rr, route):
    """forward on a side stream, backward on the autograd engine's thread (which restores the forward's stream); and four
    host threads stepping at once, each on its own stream: same loss and gradient as the plain call, bit for bit"""
    import threading
    g = torch.Generator(device="cuda").manual_seed(4)
    n = 70001
    x = torch.randn(n, 9, device="cuda", generator=g)
    x6 = torch.randn(n, 6, device="cuda", generator=g)
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda", generator=g)).detach()

    def step():
        a, b = x.clone().requires_grad_(True), x6.clone().requires_grad_(True)
        loss = rr.loss_frobenius(rr.symmetric_orthogonalization(a), rt) + \
            rr.geodesic(rr.compute_rotation_matrix_from_ortho6d(b), rt, "mean")
        loss.backward()
        return loss.detach(), a.grad, b.grad

    want = step()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        got = step()
    side.synchronize()
    assert got[0].item() == want[0].item() and torch.equal(got[1], want[1]) and torch.equal(got[2], want[2])
    results, errors = [None] * 4, []

    def worker(i):
        try:
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                for _ in range(20):
                    results[i] = step()
            s.synchronize()
        except Exception as e:                                 # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for r in results:
        assert r[0].item() == want[0].item() and torch.equal(r[1], want[1]) and torch.equal(r[2], want[2])


def test_autograd_routes_do_not_accumulate_memory(rr, route):
    """300 training steps through every differentiable route (the SVD two-call surface, the fused step, a map with each
    loss, a loss whose backward never runs): the caching allocator's live bytes return to where they were - the nodes,
    the parked gradients and the saved inputs are all released."""
    g = torch.Generator(device="cuda").manual_seed(2)
    n = 257
    x9 = torch.randn(n, 12, device="cuda", generator=g, requires_grad=True)
    x6 = torch.randn(n, 6, device="cuda", generator=g, requires_grad=True)
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda", generator=g)).detach()

    def steps():
        for _ in range(300):
            x9.grad = x6.grad = None
            rr.loss_frobenius(rr.symmetric_orthogonalization(x9[:, :9]), rt).backward()
            rr.svd_frobenius_loss(x9[:, :9], rt, with_angle=True, check=False)[0].backward()
            r6 = rr.compute_rotation_matrix_from_ortho6d(x6)
            (rr.loss_frobenius(r6, rt) + rr.geodesic(r6, rt, "mean")).backward()
            rr.loss_frobenius(rr.symmetric_orthogonalization(x9[:, :9]), rt)      # graph dropped without a backward
            rr.geodesic(rr.compute_rotation_matrix_from_quaternion(x6[:, :4]), rt, "none")
        x9.grad = x6.grad = None

    steps()                                   # warm the allocator and the per-device state
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    steps()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before


# ---------------------------------------------------------------------------------------------------------------------
# launch slots: streams, graphs
def test_graph_replay_next_to_eager_launches_on_another_stream(abi, rr):
    """A captured step replays 5000 times while another stream issues 5000 eager steps: captured launches own their
    slots, an eager stream owns its slot, so tile hand-out and sums never mix (round 1 handed slots out of one ring)."""
    n = 40000                                                  # 79 tiles: several per CTA wave, dynamic scheduler busy
    g = torch.Generator(device="cuda").manual_seed(5)
    xa, xb = torch.randn(n, 9, device="cuda", generator=g), torch.randn(n, 9, device="cuda", generator=g)
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda", generator=g)).reshape(n, 9)
    lib, check = abi.lib, abi.check
    abi.init_device(torch.cuda.current_device())

    def launch(x, dx, sums, stream):
        check(lib.pe_svd_orth_frobenius_step_metrics(x.data_ptr(), 9, rt.data_ptr(), 1.0 / n, dx.data_ptr(),
                                                     sums.data_ptr(), None, None, None, None, n, stream.cuda_stream),
              "step")

    dxa, dxb = torch.empty(n, 9, device="cuda"), torch.empty(n, 9, device="cuda")
    sa, sb = (torch.zeros(3, dtype=torch.float64, device="cuda") for _ in range(2))
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    launch(xa, dxa, sa, s1); launch(xb, dxb, sb, s2)
    torch.cuda.synchronize()
    ref_a, ref_b, ref_dxa, ref_dxb = sa.clone(), sb.clone(), dxa.clone(), dxb.clone()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=s1):
        launch(xa, dxa, sa, s1)
    worst_a = worst_b = 0.0
    for rep in range(50):
        for _ in range(100):
            with torch.cuda.stream(s1):
                graph.replay()
            launch(xb, dxb, sb, s2)
        torch.cuda.synchronize()
        worst_a = max(worst_a, ((sa - ref_a).abs() / ref_a.abs().clamp_min(1)).max().item())
        worst_b = max(worst_b, ((sb - ref_b).abs() / ref_b.abs().clamp_min(1)).max().item())
        assert torch.equal(dxa, ref_dxa) and torch.equal(dxb, ref_dxb)
    assert worst_a <= 1e-12 and worst_b <= 1e-12                # same addends, different order: float64 round-off only


def test_capture_before_init_is_refused_not_allocated():
    """A fresh process that captures its first launch without pe_init gets an error code, not a cudaMalloc inside the
    capture; after pe_init the same capture works."""
    code = textwrap.dedent(f"""
        import ctypes, sys, torch
        sys.path.insert(0, {ROOT!r})
        lib = ctypes.CDLL({os.path.join(ROOT, 'poseestimation_b200', 'libpe_rotation.so')!r})
        lib.pe_last_error.restype = ctypes.c_char_p
        vp, i64 = ctypes.c_void_p, ctypes.c_int64
        lib.pe_svd_orth_fwd.argtypes = [vp, i64, vp, vp, i64, vp]
        x = torch.randn(4096, 9, device='cuda'); r = torch.empty_like(x)
        s = torch.cuda.Stream()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        rc = None
        try:
            with torch.cuda.graph(g, stream=s):
                rc = lib.pe_svd_orth_fwd(x.data_ptr(), 9, r.data_ptr(), None, 4096, s.cuda_stream)
        except Exception:
            pass
        assert rc == -1 and b'pe_init' in lib.pe_last_error(), (rc, lib.pe_last_error())
        assert lib.pe_init(-1) == 0 and lib.pe_init(0) == 0
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            rc = lib.pe_svd_orth_fwd(x.data_ptr(), 9, r.data_ptr(), None, 4096, s.cuda_stream)
        assert rc == 0
        g.replay(); torch.cuda.synchronize()
        d = torch.linalg.det(r.view(-1, 3, 3))
        assert (d - 1).abs().max().item() < 1e-4
        print('capture ok')
    """)
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert res.returncode == 0 and "capture ok" in res.stdout, res.stdout + res.stderr


def test_many_streams_and_threads_share_nothing(abi, rr):
    """Eight streams driven from four host threads, each with its own inputs: every result equals its single-stream one."""
    import threading
    n = 30011
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda")).reshape(n, 9)
    xs = [torch.randn(n, 9, device="cuda") for _ in range(8)]
    want = []
    for x in xs:
        want.append(rr.svd_frobenius_loss_and_grad(x, rt))
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream() for _ in range(8)]
    got = [None] * 8
    errs = []

    def work(tid):
        try:
            for rep in range(200):
                for k in (tid, tid + 4):
                    with torch.cuda.stream(streams[k]):
                        got[k] = rr.svd_frobenius_loss_and_grad(xs[k], rt)
        except Exception as e:  # pragma: no cover
            errs.append(e)

    th = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    [t.start() for t in th]
    [t.join() for t in th]
    torch.cuda.synchronize()
    assert not errs
    for (lw, gw), (lg, gg) in zip(want, got):
        assert torch.equal(gw, gg) and abs(lw.item() - lg.item()) <= 1e-6


def test_repeated_launches_are_bitwise_stable(abi, rr):
    """compute-sanitizer is closed on the GPU pool (profiles/r2_sanitizer_refusal.txt), so this is the race detector for
    the shared-memory rings and the tile hand-over: 200 launches of every ring kernel family on ONE input - small grid
    per SM count, many tiles per CTA, ragged tail - must reproduce the first launch's per-record output bit for bit
    (a record read before its bulk copy landed, or overwritten before its bulk store left, flips bits), and the sums to
    float64 round-off."""
    lib, check = abi.lib, abi.check
    st = torch.cuda.current_stream().cuda_stream
    n = (1 << 20) + 77
    g = torch.Generator(device="cuda").manual_seed(9)
    x = torch.randn(n, 9, device="cuda", generator=g)
    wide = torch.randn(n, 12, device="cuda", generator=g)
    up = torch.randn(n, 9, device="cuda", generator=g)
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda", generator=g)).reshape(n, 9)
    x6 = torch.randn(n, 6, device="cuda", generator=g)
    P = lambda t: t.data_ptr()
    out = torch.empty(n, 9, device="cuda")
    out6 = torch.empty(n, 6, device="cuda")
    sums = torch.zeros(3, dtype=torch.float64, device="cuda")
    deg = torch.empty(n, dtype=torch.float64, device="cuda")
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    rt2 = rt.flip(0).contiguous()
    cases = {
        "svd_fwd": (lambda: lib.pe_svd_orth_fwd(P(x), 9, P(out), None, n, st), out),
        "svd_fwd_wide": (lambda: lib.pe_svd_orth_fwd(P(wide), 12, P(out), None, n, st), out),
        "svd_bwd": (lambda: lib.pe_svd_orth_bwd(P(x), 9, None, P(up), P(out), n, st), out),
        "svd_bwd_wide": (lambda: lib.pe_svd_orth_bwd(P(wide), 12, None, P(up), P(out), n, st), out),
        "step_metrics": (lambda: lib.pe_svd_orth_frobenius_step_metrics(P(x), 9, P(rt), 1.0 / n, P(out), P(sums), None,
                                                                        None, None, None, n, st), out),
        "step_metrics_wide": (lambda: lib.pe_svd_orth_frobenius_step_metrics(P(wide), 12, P(rt), 1.0 / n, P(out),
                                                                             P(sums), None, None, None, None, n, st), out),
        # the split wide ring without an output (in[1] re-armed behind barrier (A)) and with the R output in place
        "svd_angle_metric_wide": (lambda: lib.pe_map_metric_fwd(0, 2, P(wide), 12, P(rt), -1.0, 1.0, P(sums), P(flag), n, st),
                                  sums),
        "svd_frobenius_fwd_wide": (lambda: lib.pe_svd_orth_frobenius_fwd(P(wide), 12, P(rt), P(out), P(sums), n, st), out),
        "ortho6d_fwd": (lambda: lib.pe_ortho6d_fwd(P(x6), 6, P(out), n, st), out),
        "ortho6d_bwd": (lambda: lib.pe_ortho6d_bwd(P(x6), 6, P(up), P(out6), n, st), out6),
        "angle_error": (lambda: lib.pe_angle_error_fwd(P(rt), P(rt2), P(deg), P(sums), P(flag), n, st), deg),
    }
    for name, (call, res) in cases.items():
        check(call(), name)
        first, first_sums = res.clone(), sums.clone()
        bad = 0
        for rep in range(200):
            check(call(), name)
            if rep % 20 == 19:
                bad += int((res.view(torch.int32) != first.view(torch.int32)).sum()) if res.dtype == torch.float32 else \
                    int((res.view(torch.int64) != first.view(torch.int64)).sum())
                assert torch.allclose(sums, first_sums, rtol=1e-12, atol=0), name
        assert bad == 0, f"{name}: {bad} words changed between launches on the same input"


# ---------------------------------------------------------------------------------------------------------------------
# wide TMA stage (12-word rows)
@pytest.mark.parametrize("n", [512, 1024, 1537, 200000])
def test_wide_rows_whose_storage_ends_with_the_last_matrix(ops, rr, n):
    """A stride-12 view whose storage stops right after the ninth word of the last row (as_strided): the wide stage must
    not need the three words behind it, whether or not the last tile is full (n % 512 == 0)."""
    base = torch.randn(n * 12 - 3, device="cuda")
    view = base.as_strided((n, 9), (12, 1))
    dense = view.contiguous()
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda"))
    up = torch.randn(n, 3, 3, device="cuda")
    assert torch.equal(ops.svd_orth(view), ops.svd_orth(dense))
    assert torch.equal(ops.svd_orth_bwd_recompute(view, up), ops.svd_orth_bwd_recompute(dense, up))
    lv, gv = ops.svd_orth_frobenius_step(view, rt)
    ld, gd = ops.svd_orth_frobenius_step(dense, rt)
    assert torch.equal(gv, gd) and abs(lv.item() - ld.item()) <= 1e-6
    sv, _ = ops.svd_orth_frobenius_sum(view, rt, False)
    sd, _ = ops.svd_orth_frobenius_sum(dense, rt, False)
    assert abs(sv.item() - sd.item()) <= 1e-9 * abs(sd.item())
    assert torch.equal(ops.svd_orth_frobenius_bwd(view, rt, torch.ones((), device="cuda")),
                       ops.svd_orth_frobenius_bwd(dense, rt, torch.ones((), device="cuda")))
    assert abs(float(rr.map_metric("SVD", view, rt)) - float(rr.map_metric("SVD", dense, rt))) <= 1e-9 * 180


@pytest.mark.parametrize("stride,offset", [(9, 1), (9, 3), (10, 0), (11, 2), (13, 1), (16, 0), (31, 5), (12, 1), (12, 4)])
@pytest.mark.parametrize("n", [1, 511, 512, 513, 4097, 150001])
def test_rows_of_any_stride_at_any_base(ops, rr, n, stride, offset):
    """ANY row stride and any 4-byte aligned base of the input matrix array: dense rows that start `offset` words into
    their storage, rows of 10..31 words (direct kernel), 12-word rows (wide stage when 16-byte aligned, else direct).
    Every SVD-family op must return what it returns on a contiguous copy, bit for bit, whatever the route."""
    g = torch.Generator(device="cuda").manual_seed(n * 64 + stride)
    base = torch.randn(offset + n * stride, device="cuda", generator=g)
    view = base.as_strided((n, 9), (stride, 1), offset)
    dense = view.contiguous()
    assert view.data_ptr() % 16 == (4 * offset) % 16
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device="cuda", generator=g))
    up = torch.randn(n, 3, 3, device="cuda", generator=g)
    r_v, d_v = ops.svd_orth_det(view)
    r_d, d_d = ops.svd_orth_det(dense)
    assert torch.equal(r_v, r_d) and torch.equal(d_v, d_d)
    assert torch.equal(ops.svd_orth(view), r_d)
    assert torch.equal(ops.svd_orth_bwd_recompute(view, up), ops.svd_orth_bwd_recompute(dense, up))
    lv, gv = ops.svd_orth_frobenius_step(view, rt)
    ld, gd = ops.svd_orth_frobenius_step(dense, rt)
    assert torch.equal(gv, gd) and abs(lv.item() - ld.item()) <= 1e-6
    sv, rv = ops.svd_orth_frobenius_sum(view, rt, True)
    sd, rd = ops.svd_orth_frobenius_sum(dense, rt, True)
    # the two lanes of a thread meet in one float32 add before the float64 accumulation, and the routes pair the records
    # differently (t, t + 256 of a tile / i, i + ceil(n/2)): the sums agree to float32 rounding of those pair sums
    assert torch.equal(rv, rd) and abs(sv.item() - sd.item()) <= 1e-8 * abs(sd.item())
    one = torch.ones((), device="cuda")
    assert torch.equal(ops.svd_orth_frobenius_bwd(view, rt, one), ops.svd_orth_frobenius_bwd(dense, rt, one))
    for metric in ("frobenius", "geodesic", "angle"):
        a = float(rr.map_metric("SVD", view, rt, metric)) if metric != "angle" else float(rr.map_metric("SVD", view, rt))
        b = float(rr.map_metric("SVD", dense, rt, metric)) if metric != "angle" else float(rr.map_metric("SVD", dense, rt))
        assert abs(a - b) <= 1e-9 * 180


def test_wide_route_is_taken_and_can_be_switched_off(rr):
    """PE_NO_WIDE=1 (A/B hook) sends the same view through the direct kernel: same numbers either way."""
    code = textwrap.dedent(f"""
        import sys, torch
        sys.path.insert(0, {ROOT!r})
        import poseestimation_b200.rotation_representation as rr
        torch.manual_seed(0)
        wide = torch.randn(100000, 12, device='cuda')
        r = rr.symmetric_orthogonalization(wide[:, :9])
        print('sum %.10e' % r.double().sum().item())
    """)
    outs = []
    for env in ({}, {"PE_NO_WIDE": "1"}):
        res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600,
                             env={**os.environ, **env})
        assert res.returncode == 0, res.stderr
        outs.append(res.stdout.strip().splitlines()[-1])
    assert outs[0] == outs[1]


# ---------------------------------------------------------------------------------------------------------------------
# small gaps against the reference surface
def test_transform_pts_4d_vs_reference_golden(golden):
    import poseestimation_b200.iterative as it
    z = golden("round2")
    got = it.transform_pts(T(z["tp4_T"]).cuda(), T(z["tp4_pts"]).cuda())
    assert got.shape == (6, 3, 50, 3) and np.abs(got.cpu().double().numpy() - z["tp4_out"]).max() <= 1e-5
    with pytest.raises(ValueError):
        it.transform_pts(torch.zeros(6, 4, device="cuda"), T(z["tp4_pts"]).cuda())


def test_so3_exp_map_with_any_eps_vs_reference_golden(rr, golden):
    z = golden("round2")
    v = T(z["exp_v"]).cuda()
    w = T(z["exp_w"]).float().cuda()
    for tag, eps in (("1em2", 1e-2), ("1em6", 1e-6)):
        vd = v.clone().requires_grad_(True)
        r = rr.so3_exp_map(vd, eps)
        (r * w).sum().backward()
        assert np.abs(r.detach().cpu().double().numpy() - z[f"exp_r_{tag}"]).max() <= 1e-5
        scale = np.abs(z[f"exp_g_{tag}"]).max(1)
        rel = np.abs(vd.grad.cpu().double().numpy() - z[f"exp_g_{tag}"]).max(1) / np.maximum(scale, 1.0)
        assert rel.max() <= 1e-4
    assert torch.equal(rr.so3_exp_map(v), rr.vec_3d_to_SO3(v)) and torch.equal(rr.so3_exp_map(v, 0.0001), rr.so3_exp_map(v))


def test_calculate_t_pred_ignores_rot_repr_and_angle_error_keywords(rr):
    import poseestimation_b200.iterative as it
    out = torch.randn(8, 12, device="cuda")
    t_init = torch.eye(4, device="cuda").repeat(8, 1, 1)
    t_init[:, 2, 3] = 2.0
    a = it.calculate_T_pred(out, t_init, "cuda")
    b = it.calculate_T_pred(out, t_init, "cuda", rot_repr="6D")        # Iterative/utility.py:90-107 never reads it
    assert torch.equal(a, b)
    r1 = rr.symmetric_orthogonalization(torch.randn(8, 9, device="cuda"))
    r2 = rr.symmetric_orthogonalization(torch.randn(8, 9, device="cuda"))
    assert torch.equal(rr.angle_error(t_R1=r1, t_R2=r2), rr.angle_error(r1, r2))


def test_mismatched_batches_raise_instead_of_reading_past_the_end(rr, ops):
    """the two-matrix ops took n from the first argument only."""
    a = rr.symmetric_orthogonalization(torch.randn(100, 9, device="cuda"))
    b = rr.symmetric_orthogonalization(torch.randn(60, 9, device="cuda"))
    for fn in (rr.angle_error, rr.mean_angle_error, rr.compute_geodesic_distance_from_two_matrices, rr.geodesic,
               rr.loss_frobenius):
        with pytest.raises(ValueError, match="batch mismatch"):
            fn(a, b)
    with pytest.raises(ValueError):
        ops.geodesic_bwd(a, a, -1.0, 1.0, torch.ones(60, device="cuda"), 1.0, True, False)
    with pytest.raises(ValueError, match="batch mismatch"):
        ops.svd_orth_bwd_recompute(torch.randn(100, 9, device="cuda"), b)
    with pytest.raises(ValueError, match="batch mismatch"):
        ops.ortho6d_bwd(torch.randn(100, 6, device="cuda"), b)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_second_device_and_tensors_on_two_devices(rr, ops):
    """Per-device set-up (launch slots, kernel attributes) on a device other than the current one, and arguments that
    live on different devices: an error, not a kernel handed a foreign pointer."""
    g = torch.Generator().manual_seed(3)
    x = torch.randn(5000, 9, generator=g)
    rt = O.symmetric_orthogonalization(torch.randn(5000, 9, generator=g).double()).float()
    outs = []
    for dev in ("cuda:0", "cuda:1"):
        xd = x.to(dev).requires_grad_(True)
        loss, ang = rr.svd_frobenius_loss(xd, rt.to(dev), with_angle=True)      # current device stays cuda:0
        loss.backward()
        l2 = rr.loss_frobenius(rr.symmetric_orthogonalization(xd), rt.to(dev))
        outs.append((loss.item(), ang.item(), l2.item(), xd.grad.cpu()))
    assert outs[0][:3] == outs[1][:3] and torch.equal(outs[0][3], outs[1][3])
    a = rr.symmetric_orthogonalization(x.to("cuda:0"))
    b = rt.to("cuda:1")
    for fn in (rr.loss_frobenius, rr.angle_error, rr.mean_angle_error, rr.geodesic,
               rr.compute_geodesic_distance_from_two_matrices):
        with pytest.raises(ValueError, match="different devices"):
            fn(a, b)
    with pytest.raises((ValueError, RuntimeError), match="different devices"):
        rr.svd_frobenius_loss(x.to("cuda:0").requires_grad_(True), b)
    with pytest.raises(ValueError, match="different devices"):
        rr.svd_frobenius_loss(x.to("cuda:0"), b)
    with pytest.raises(ValueError, match="different devices"):
        rr.map_metric("SVD", x.to("cuda:0"), b)
    with pytest.raises(ValueError, match="different devices"):
        rr.kabsch(torch.rand(4, 64, 3, device="cuda:0"), torch.rand(4, 64, 3, device="cuda:1"))
