"""GPU parity for the SURVEY 8(f) rows: cloud-pair construction, differentiable Kabsch, CosyPose pose update, ADD-L1."""
import numpy as np
import pytest
import torch

from oracle import reference_port as O

pytestmark = pytest.mark.gpu

T = lambda a: torch.from_numpy(np.asarray(a))


@pytest.fixture(scope="module")
def it():
    import poseestimation_b200.iterative as m
    return m


@pytest.fixture(scope="module")
def rr():
    import poseestimation_b200.rotation_representation as m
    return m


@pytest.fixture(scope="module")
def pc():
    import poseestimation_b200.point_cloud as m
    return m


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    a, b = a.reshape(len(a), -1), b.reshape(len(b), -1)
    return np.abs(a - b).max(1) / (np.abs(b).max(1) + 1e-30)


# ---- 8f.4 cloud-pair construction ------------------------------------------------------------------------------------
def test_sampler_map_matches_reference_golden(pc, golden):
    z = golden("sampler")
    m = pc.rotation_matrices_from_axis_angle(T(z["theta"]).cuda(), T(z["axis"]).cuda())
    assert m.shape == (256, 3, 3) and m.dtype == torch.float32
    assert np.abs(m.cpu().numpy() - z["matrix"]).max() <= 2e-6
    r = pc.get_sampled_rotation_matrices_by_axisAngle(1000).cpu().double()
    assert (r @ r.transpose(1, 2) - torch.eye(3, dtype=torch.float64)).abs().max() <= 1e-5
    assert (torch.linalg.det(r) - 1).abs().max() <= 1e-5
    # rotation angle is 2 * theta (the reference's quirk): uniform theta -> trace spread over [-1, 3]
    tr = r.diagonal(dim1=1, dim2=2).sum(1)
    assert tr.min() < -0.9 and tr.max() > 2.9


def test_rotate_cloud_matches_reference_construction(pc, golden):
    z = golden("kabsch_kat")                      # pc2 = bmm(expanded gt_rmat, pc1), point_cloud/main.py:173-181
    out = pc.rotate_cloud(T(z["gt_rmat"]).cuda(), T(z["pc1"]).cuda())
    assert np.abs(out.cpu().numpy() - z["pc2"]).max() <= 1e-6


@pytest.mark.parametrize("b,n", [(1, 4), (3, 5), (7, 1024), (33, 1021), (2, 4096)])
def test_affine_points_forward_backward(b, n):
    from poseestimation_b200 import ops
    torch.manual_seed(b * 1000 + n)
    a, t, p, g = torch.randn(b, 3, 3), torch.randn(b, 3), torch.randn(b, n, 3), torch.randn(b, n, 3)
    for transpose in (False, True):
        ad, td, pd = (v.cuda().requires_grad_(True) for v in (a, t, p))
        out = ops.affine_points(ad, td, pd, transpose)
        out.backward(g.cuda())
        a64, t64, p64 = (v.double().requires_grad_(True) for v in (a, t, p))
        m = a64.transpose(1, 2) if transpose else a64
        ref = p64 @ m.transpose(1, 2) + t64[:, None]
        ref.backward(g.double())
        assert np.abs(out.detach().cpu().numpy() - ref.detach().numpy()).max() <= 1e-5
        assert _rel(pd.grad.cpu(), p64.grad).max() <= 1e-5
        assert _rel(ad.grad.cpu(), a64.grad).max() <= 1e-4      # fp32 sums over n points
        assert _rel(td.grad.cpu(), t64.grad).max() <= 1e-4


def test_transform_pts_reads_the_pose_in_place(it, golden):
    z = golden("iterative")
    out = it.transform_pts(T(z["T_gt"]).cuda(), T(z["verts"]).cuda())
    assert np.abs(out.cpu().numpy() - z["transform_pts"]).max() <= 1e-6


# ---- 8f.3 differentiable Kabsch --------------------------------------------------------------------------------------
@pytest.mark.parametrize("b,n,noise", [(64, 40, 0.05), (9, 1024, 0.01), (40, 257, 0.1)])
def test_kabsch_backward_matches_oracle_autograd(pc, b, n, noise):
    torch.manual_seed(n)
    p = (torch.rand(b, n, 3) - 0.5).float()
    rot = O.sample_rotations_axis_angle(b, torch.Generator().manual_seed(3)).float()
    q = (p @ rot.transpose(1, 2) + noise * torch.randn(b, n, 3) + torch.randn(b, 1, 3)).float()
    g_r, g_t = torch.randn(b, 3, 3), torch.randn(b, 3)
    p64, q64 = p.double().requires_grad_(True), q.double().requires_grad_(True)
    r64, t64 = O.kabsch(p64, q64)
    ((r64 * g_r.double()).sum() + (t64 * g_t.double()).sum()).backward()
    pd, qd = p.cuda().requires_grad_(True), q.cuda().requires_grad_(True)
    r, t = pc.kabsch(pd, qd)
    ((r * g_r.cuda()).sum() + (t * g_t.cuda()).sum()).backward()
    assert np.abs(r.detach().cpu().numpy() - r64.detach().numpy()).max() <= 2e-5
    assert np.abs(t.detach().cpu().numpy() - t64.detach().numpy()).max() <= 2e-5
    # per cloud: gradient error relative to the largest gradient entry of that cloud
    assert _rel(pd.grad.cpu(), p64.grad).max() <= 2e-4
    assert _rel(qd.grad.cpu(), q64.grad).max() <= 2e-4
    # only R used, only one cloud needs a gradient
    pd2 = p.cuda().requires_grad_(True)
    r2, _ = pc.kabsch(pd2, q.cuda())
    (r2 * g_r.cuda()).sum().backward()
    p64b = p.double().requires_grad_(True)
    (O.kabsch(p64b, q.double())[0] * g_r.double()).sum().backward()
    assert _rel(pd2.grad.cpu(), p64b.grad).max() <= 2e-4


@pytest.mark.parametrize("b,n,noise", [(64, 40, 0.05), (9, 1024, 0.01), (40, 257, 0.1), (3, 2048 + 8, 0.02)])
def test_weighted_kabsch_forward_backward_matches_oracle_autograd(pc, b, n, noise):
    torch.manual_seed(n + 1)
    p = (torch.rand(b, n, 3) - 0.5).float()
    rot = O.sample_rotations_axis_angle(b, torch.Generator().manual_seed(5)).float()
    q = (p @ rot.transpose(1, 2) + noise * torch.randn(b, n, 3) + torch.randn(b, 1, 3)).float()
    w = torch.rand(b, n).float()
    w[:, 1::7] = 0.0                                                      # some correspondences switched off
    g_r, g_t = torch.randn(b, 3, 3), torch.randn(b, 3)
    p64, q64, w64 = (x.double().requires_grad_(True) for x in (p, q, w))
    r64, t64 = O.kabsch_weighted(p64, q64, w64)
    ((r64 * g_r.double()).sum() + (t64 * g_t.double()).sum()).backward()
    pd, qd, wd = (x.cuda().requires_grad_(True) for x in (p, q, w))
    r, t = pc.kabsch(pd, qd, weights=wd)
    ((r * g_r.cuda()).sum() + (t * g_t.cuda()).sum()).backward()
    assert np.abs(r.detach().cpu().numpy() - r64.detach().numpy()).max() <= 2e-5
    assert np.abs(t.detach().cpu().numpy() - t64.detach().numpy()).max() <= 2e-5
    assert _rel(pd.grad.cpu(), p64.grad).max() <= 2e-4
    assert _rel(qd.grad.cpu(), q64.grad).max() <= 2e-4
    assert _rel(wd.grad.cpu(), w64.grad).max() <= 2e-4
    # only the weights need a gradient (the network-predicted-correspondence case): same dw, nothing else written
    wd2 = w.cuda().requires_grad_(True)
    r2, t2 = pc.kabsch(p.cuda(), q.cuda(), weights=wd2)
    ((r2 * g_r.cuda()).sum() + (t2 * g_t.cuda()).sum()).backward()
    assert torch.equal(wd2.grad, wd.grad)


def test_weighted_kabsch_properties(pc, rr):
    b, n = 50, 512
    g = torch.Generator(device="cuda").manual_seed(8)
    p = torch.rand(b, n, 3, device="cuda", generator=g) - 0.5
    rot, q = pc.make_cloud_pair(p, generator=g)
    q = q + torch.randn(b, 1, 3, device="cuda", generator=g)
    # unit weights == the plain kernel; a common factor on the weights changes nothing
    r0, t0 = pc.kabsch(p, q)
    r1, t1 = pc.kabsch(p, q, weights=torch.ones(b, n, device="cuda"))
    assert float((r1 - r0).abs().max()) <= 2e-6 and float((t1 - t0).abs().max()) <= 2e-6
    w = torch.rand(b, n, device="cuda", generator=g)
    ra, ta = pc.kabsch(p, q, weights=w)
    rb, tb = pc.kabsch(p, q, weights=w * 8.0)
    assert float((ra - rb).abs().max()) <= 2e-6 and float((ta - tb).abs().max()) <= 2e-6
    assert float((ra - rot).abs().max()) <= 2e-5
    # gross outliers with zero weight are ignored
    qo = q.clone()
    qo[:, ::4] += 50.0
    wo = torch.ones(b, n, device="cuda")
    wo[:, ::4] = 0.0
    ro, to = rr.kabsch(p, qo, weights=wo)
    assert float((ro - rot).abs().max()) <= 2e-5 and float((to - t0).abs().max()) <= 5e-5
    # all-zero weights: a finite rotation, not NaN
    rz, tz = pc.kabsch(p[:3], q[:3], weights=torch.zeros(3, n, device="cuda"))
    assert bool(torch.isfinite(rz).all()) and bool(torch.isfinite(tz).all())
    assert float((torch.bmm(rz, rz.transpose(1, 2)) - torch.eye(3, device="cuda")).abs().max()) <= 1e-5
    # direct (non-TMA) kernels give the same bits as the staged ones
    import os
    os.environ["PE_FORCE_DIRECT"] = "1"
    try:
        rd, td = pc.kabsch(p, q, weights=w)
    finally:
        del os.environ["PE_FORCE_DIRECT"]
    assert float((rd - ra).abs().max()) <= 2e-6 and float((td - ta).abs().max()) <= 2e-6
    with pytest.raises((ValueError, RuntimeError)):
        pc.kabsch(p, q, weights=torch.ones(b, n + 1, device="cuda"))


# ---- 8f.2 CosyPose pose update and ADD-L1 ----------------------------------------------------------------------------
def test_pose_update_matches_reference_golden(it, golden):
    z = golden("iterative")
    n = len(z["out"])
    wide = torch.zeros(n, 16)
    wide[:, :12] = T(z["out"])                      # the network emits more columns than the update reads
    for src in (T(z["out"]).cuda(), wide.cuda()):
        o = src.requires_grad_(True)
        tp = it.calculate_T_pred(o, T(z["T_init"]).cuda(), "cuda")
        assert tp.shape == (n, 4, 4) and tp.dtype == torch.float32
        assert np.abs(tp.detach().cpu().numpy() - z["T_pred_ref_fp32"]).max() <= 2e-5     # the reference, fp32
        o64 = T(z["out"]).double().requires_grad_(True)
        t64 = O.calculate_T_pred(o64, T(z["T_init"]).double())
        assert np.abs(tp.detach().cpu().numpy() - t64.detach().numpy()).max() <= 1e-5
        tp.backward(T(z["g_T"]).float().cuda())
        (g64,) = torch.autograd.grad(t64, o64, T(z["g_T"]))
        assert _rel(o.grad.cpu()[:, :12], g64).max() <= 1e-4
        assert o.grad.shape == src.shape and (o.grad[:, 12:] == 0).all()
    # rot_repr is accepted and ignored, exactly as Iterative/utility.py:90-107 does
    assert torch.equal(it.calculate_T_pred(T(z["out"]).cuda(), T(z["T_init"]).cuda(), "cuda", rot_repr="6D"),
                       it.calculate_T_pred(T(z["out"]).cuda(), T(z["T_init"]).cuda(), "cuda"))


def test_parametrization_svd_is_the_pose_update_kernel(rr, it, golden):
    """rotation_representation.py:346-373 (cannot run in the reference): same formulas as Iterative/utility.py:90-128,
    so with the 'SVD' key it must reproduce the golden poses; the '6D' key goes through the native 6D map."""
    z = golden("iterative")
    out, ti = T(z["out"]).cuda(), T(z["T_init"]).cuda()
    fx, fy = it.get_scene_parameters()
    tp = rr.parametrization(out, ti, fx, fy, 160.0, 160.0, "SVD")
    assert np.abs(tp.cpu().numpy() - z["T_pred_ref_fp32"]).max() <= 2e-5
    x6 = torch.randn(len(out), 9, device="cuda")
    t6 = rr.parametrization(x6, ti, fx, fy, 160.0, 160.0, "6D")
    r6 = rr.compute_rotation_matrix_from_ortho6d(x6[:, :6])
    assert float((t6[:, :3, :3] - r6 @ ti[:, :3, :3]).abs().max()) <= 1e-6
    assert float((t6[:, 2, 3] - x6[:, 8] * ti[:, 2, 3]).abs().max()) <= 1e-6
    assert (t6[:, 3] == torch.tensor([0.0, 0.0, 0.0, 1.0], device="cuda")).all()


def test_add_l1_losses_match_reference_golden(it, golden):
    z = golden("iterative")
    tg, v = T(z["T_gt"]).cuda(), T(z["verts"]).cuda()
    tp = T(z["T_pred_ref_fp32"]).cuda().requires_grad_(True)
    add = it.compute_ADD_L1_loss(tg, tp, v)
    assert add.dim() == 0 and add.dtype == torch.float32
    add.backward()
    assert abs(add.item() - float(z["add_l1"])) <= 1e-6
    assert np.abs(tp.grad.cpu().numpy() - z["add_l1_grad_pred"]).max() <= 1e-6
    items = it.compute_ADD_L1_loss(tg, tp.detach(), v, use_batch_mean=False)
    assert np.abs(items.cpu().numpy() - z["add_l1_items"]).max() <= 1e-6
    tp2 = T(z["T_pred_ref_fp32"]).cuda().requires_grad_(True)
    dis = it.compute_disentangled_ADD_L1_loss(tp2, tg, v)
    dis.backward()
    assert abs(dis.item() - float(z["disentangled"])) <= 1e-6
    assert np.abs(tp2.grad.cpu().numpy() - z["disentangled_grad_pred"]).max() <= 1e-6
    # per-item form with a non-uniform upstream gradient
    tp3 = T(z["T_pred_ref_fp32"]).cuda().requires_grad_(True)
    w = torch.randn(len(z["out"]))
    (it.compute_ADD_L1_loss(tg, tp3, v, use_batch_mean=False) * w.cuda()).sum().backward()
    a64 = T(z["T_pred_ref_fp32"]).double().requires_grad_(True)
    (O.compute_ADD_L1_loss(T(z["T_gt"]).double(), a64, T(z["verts"]).double(), use_batch_mean=False) * w.double()).sum().backward()
    assert np.abs(tp3.grad.cpu().numpy() - a64.grad.numpy()).max() <= 1e-5


# (the last three are small batches of large meshes: each item is shared by 4 / 16 / 8 warps and added atomically)
@pytest.mark.parametrize("b,n", [(1, 1), (5, 3), (19, 512), (8, 1500), (150, 2048 + 4), (3, 5000), (5, 32768 + 4),
                                 (64, 8192)])
def test_add_l1_shapes_and_ragged_sizes(it, b, n):
    torch.manual_seed(b + n)
    def poses(k):
        t = torch.zeros(k, 4, 4)
        t[:, :3, :3] = O.symmetric_orthogonalization(torch.randn(k, 9).double()).float()
        t[:, :3, 3] = torch.randn(k, 3)
        t[:, 3, 3] = 1
        return t
    tg, tp, v = poses(b), poses(b), torch.randn(b, n, 3)
    for fn_gpu, fn_ref in ((lambda a, c, d: it.compute_ADD_L1_loss(a, c, d), lambda a, c, d: O.compute_ADD_L1_loss(a, c, d)),
                           (lambda a, c, d: it.compute_disentangled_ADD_L1_loss(c, a, d),
                            lambda a, c, d: O.compute_disentangled_ADD_L1_loss(c, a, d))):
        x = tp.cuda().requires_grad_(True)
        loss = fn_gpu(tg.cuda(), x, v.cuda())
        loss.backward()
        x64 = tp.double().requires_grad_(True)
        ref = fn_ref(tg.double(), x64, v.double())
        ref.backward()
        assert abs(loss.item() - ref.item()) <= 1e-5 * max(1.0, abs(ref.item()))
        assert np.abs(x.grad.cpu().numpy() - x64.grad.numpy()).max() <= 1e-5 * max(1.0, x64.grad.abs().max().item())


def test_refiner_step_end_to_end(it, golden):
    """Iterative/main.py:85-94: network output -> calculate_T_pred -> disentangled ADD-L1 -> backward."""
    z = golden("iterative")
    o = T(z["out"]).cuda().requires_grad_(True)
    loss = it.compute_disentangled_ADD_L1_loss(it.calculate_T_pred(o, T(z["T_init"]).cuda(), "cuda"),
                                               T(z["T_gt"]).cuda(), T(z["verts"]).cuda())
    loss.backward()
    o64 = T(z["out"]).double().requires_grad_(True)
    ref = O.compute_disentangled_ADD_L1_loss(O.calculate_T_pred(o64, T(z["T_init"]).double()), T(z["T_gt"]).double(),
                                             T(z["verts"]).double())
    ref.backward()
    assert abs(loss.item() - ref.item()) <= 1e-5
    # sign(.) flips of individual vertices between fp32 and float64 move the L1 gradient by O(1/N) per flip
    assert _rel(o.grad.cpu(), o64.grad).max() <= 5e-3
    ang = it.angle_error(it.calculate_T_pred(o.detach(), T(z["T_init"]).cuda(), "cuda")[:, :3, :3].contiguous(),
                         T(z["T_gt"])[:, :3, :3].contiguous().cuda())
    assert ang.dtype == torch.float64 and ang.shape == (len(z["out"]),)


def test_empty_and_tiny_batches(it, pc):
    from poseestimation_b200 import ops
    z3 = torch.empty(0, 16, 3, device="cuda")
    assert pc.rotate_cloud(torch.empty(0, 3, 3, device="cuda"), z3).shape == (0, 16, 3)
    r, t = pc.kabsch(z3, z3)
    assert r.shape == (0, 3, 3) and t.shape == (0, 3)
    assert pc.rotation_matrices_from_axis_angle(torch.empty(0, device="cuda"), torch.empty(0, 3, device="cuda")).shape == (0, 3, 3)
    assert it.calculate_T_pred(torch.empty(0, 12, device="cuda"), torch.empty(0, 4, 4, device="cuda")).shape == (0, 4, 4)
    item, tot, g = ops.add_l1(torch.empty(0, 4, 4, device="cuda"), torch.empty(0, 4, 4, device="cuda"), z3, True)
    assert item.shape == (0,) and tot.item() == 0.0 and g.shape == (0, 3, 4)
    # one item, one vertex, identical poses: zero loss and (sign(0) = 0) zero gradient, as torch's abs() gives
    T1 = torch.eye(4, device="cuda").reshape(1, 4, 4)
    tp = T1.clone().requires_grad_(True)
    loss = it.compute_disentangled_ADD_L1_loss(tp, T1, torch.ones(1, 1, 3, device="cuda"))
    loss.backward()
    assert loss.item() == 0.0 and tp.grad.abs().max().item() == 0.0
