"""The oracle (oracle/reference_port.py) against outputs of the UNMODIFIED reference (tests/golden/*.npz,
made by tests/golden/make_golden.py in a container that has /root/reference).  CPU only."""
import numpy as np
import pytest
import torch

from oracle import reference_port as O

T = lambda a: torch.from_numpy(np.asarray(a))


def _fwd_bwd(fn, x, g):
    x = T(x).double().requires_grad_(True)
    y = fn(x)
    (dx,) = torch.autograd.grad(y, x, T(g).double())
    return y.detach().numpy(), dx.numpy()


@pytest.mark.parametrize("name,fn,tol", [
    ("ortho6d", O.compute_rotation_matrix_from_ortho6d, 1e-12),
    ("quaternion", O.compute_rotation_matrix_from_quaternion, 1e-12),
    ("euler", O.compute_rotation_matrix_from_euler, 1e-12),
    ("expmap", O.vec_3d_to_SO3, 1e-12),
    # the reference's stereographic_unproject round-trips through a float32 buffer (rotation_representation.py:84)
    ("ortho5d", O.compute_rotation_matrix_from_ortho5d, 5e-6),
])
def test_maps_match_reference(golden, name, fn, tol):
    z = golden(name)
    r, dx = _fwd_bwd(fn, z["x"], z["g"])
    assert np.abs(r - z["r"]).max() <= tol
    scale = np.abs(z["dx"]).max(axis=-1, keepdims=True) + 1e-300
    assert (np.abs(dx - z["dx"]) / scale).max() <= max(tol, 1e-10) * 10


def test_svd_orth_matches_reference(golden):
    z = golden("svd_orth")
    r, dx = _fwd_bwd(O.symmetric_orthogonalization, z["x"], z["g"])
    assert np.abs(r - z["r"]).max() <= 1e-12
    det = O.svd_orth_det_sign(T(z["x"]).double()).numpy()
    assert np.array_equal(np.sign(det), np.sign(z["det"]))
    # autograd through torch.svd is noisy near repeated singular values: compare where it is not
    s = z["s"]
    ok = (np.minimum(s[:, 0] - s[:, 1], s[:, 1] - s[:, 2]) / s[:, 0] > 1e-3)
    rel = np.abs(dx - z["dx"]).reshape(len(s), -1).max(1) / (np.abs(z["dx"]).reshape(len(s), -1).max(1) + 1e-300)
    assert rel[ok].max() <= 1e-8


def test_svd_orth_closed_form_pullback_matches_reference_autograd(golden):
    z = golden("svd_orth")
    x, g = T(z["x"]), T(z["g"])
    dm = O.svd_orth_pullback(x, g).numpy()
    s = z["s"]
    d = np.sign(z["det"])
    cond = O.svd_orth_conditioning(x).numpy()
    gap = np.minimum(s[:, 0] - s[:, 1], s[:, 1] - s[:, 2]) / s[:, 0]
    ok = (cond > 1e-3) & (gap > 1e-3)
    ref = z["dx"].reshape(-1, 3, 3)
    rel = np.abs(dm - ref).reshape(len(s), -1).max(1) / (np.abs(ref).reshape(len(s), -1).max(1) + 1e-300)
    assert ok.sum() > 500
    assert rel[ok].max() <= 1e-7
    assert d.min() == -1 and d.max() == 1          # both determinant signs are exercised


def test_svd_orth_known_answers(golden):
    z = golden("svd_orth")
    n = int(z["n_random"])
    x = T(z["x"]).double()
    r = O.symmetric_orthogonalization(x).numpy()
    assert np.allclose(r[n], np.eye(3), atol=1e-14)                      # identity -> identity
    assert abs(np.linalg.det(r[n + 1]) - 1) < 1e-12                      # diag(1,1,-1) -> proper rotation
    rots = z["x"][n + 3:n + 3 + 64].reshape(-1, 3, 3)                    # rotations map to themselves
    assert np.abs(r[n + 3:n + 3 + 64] - rots).max() < 1e-6               # (inputs are fp32-rounded rotations)
    assert np.allclose(np.linalg.det(r), 1.0, atol=1e-9)
    assert np.abs(r @ r.transpose(0, 2, 1) - np.eye(3)).max() < 1e-9


def test_euler_matches_getR(golden):
    z = golden("euler_getR")
    r = O.compute_rotation_matrix_from_euler(T(z["x"]).double()).numpy()
    assert np.abs(r - z["r"]).max() < 1e-14


def test_distances_match_reference(golden):
    z = golden("distances")
    r1, r2, direct = T(z["r1"]), T(z["r2"]), T(z["direct"])
    assert np.abs(O.compute_geodesic_distance_from_two_matrices(r1.double(), r2.double()).numpy() - z["geo_rr"]).max() < 1e-12
    assert np.abs(O.compute_geodesic_distance_from_two_matrices(direct.double(), r2.double()).numpy()
                  - z["geo_rr_direct"]).max() < 1e-12
    ae = O.angle_error(r1, r2)
    assert ae.dtype == torch.float64
    assert np.abs(ae.numpy() - z["angle_error"]).max() < 1e-10
    assert np.abs(O.angle_error(direct, r2).numpy() - z["angle_error_direct"]).max() < 1e-10
    for red in ("none", "mean", "sum"):
        got = O.geodesic(r1.double(), r2.double(), red).numpy()
        assert np.abs(got - z[f"geo_pc_{red}"]).max() < 1e-9
    a = direct.double().requires_grad_(True)
    b = r2.double().requires_grad_(True)
    v = O.loss_frobenius(a, b)
    ga, gb = torch.autograd.grad(v, (a, b))
    assert abs(v.item() - float(z["frob"])) < 1e-14
    assert np.abs(ga.numpy() - z["frob_grad_pred"]).max() < 1e-14
    assert np.abs(gb.numpy() - z["frob_grad_true"]).max() < 1e-14
    assert bool(z["angle_error_raises_bad"])
    with pytest.raises(ValueError):
        O.angle_error(T(z["bad"]), r2[:4])


def test_kabsch_recovers_reference_rotation(golden):
    z = golden("kabsch_kat")
    r, t = O.kabsch(T(z["pc1"]), T(z["pc2"]))
    assert np.abs(r.numpy() - z["gt_rmat"]).max() < 1e-5                 # inputs are fp32 products
    assert np.abs(t.numpy()).max() < 1e-5
    # with a translation and an offset far from the origin
    shift = torch.tensor([10.0, -7.0, 3.0])
    r2, t2 = O.kabsch(T(z["pc1"]).double() + 5.0, T(z["pc2"]).double() + shift)
    assert np.abs(r2.numpy() - z["gt_rmat"]).max() < 1e-5


def test_weighted_kabsch_restatement(golden):
    """No reference function: unit weights reduce to the plain alignment, integer weights to repeated
    points, zero weights remove outliers, and a scaling of all weights changes nothing."""
    z = golden("kabsch_kat")
    p, q = T(z["pc1"]).double(), T(z["pc2"]).double()
    b, n = p.shape[0], p.shape[1]
    r0, t0 = O.kabsch(p, q)
    r1, t1 = O.kabsch_weighted(p, q, torch.ones(b, n))
    assert (r1 - r0).abs().max() < 1e-12 and (t1 - t0).abs().max() < 1e-12
    g = torch.Generator().manual_seed(4)
    qn = q + 0.05 * torch.randn(q.shape, generator=g, dtype=torch.float64)
    k = torch.randint(0, 4, (b, n), generator=g)
    k[:, 0] = 1
    rw, tw = O.kabsch_weighted(p, qn, k.double())
    for i in range(min(b, 4)):                                            # repeat point j k[i, j] times
        idx = torch.repeat_interleave(torch.arange(n), k[i])
        ri, ti = O.kabsch(p[i:i + 1, idx], qn[i:i + 1, idx])
        assert (ri - rw[i:i + 1]).abs().max() < 1e-10 and (ti - tw[i:i + 1]).abs().max() < 1e-10
    rs, ts = O.kabsch_weighted(p, qn, 7.5 * k.double())
    assert (rs - rw).abs().max() < 1e-12 and (ts - tw).abs().max() < 1e-12
    # gross outliers with zero weight do not move the answer
    qo = q.clone()
    qo[:, ::5] += 100.0
    w = torch.ones(b, n, dtype=torch.float64)
    w[:, ::5] = 0
    ro, to = O.kabsch_weighted(p, qo, w)
    assert np.abs(ro.numpy() - z["gt_rmat"]).max() < 1e-5 and to.abs().max() < 1e-5


def test_sampler_gives_rotations():
    g = torch.Generator().manual_seed(0)
    r = O.sample_rotations_axis_angle(64, g).numpy()
    assert np.abs(r @ r.transpose(0, 2, 1) - np.eye(3)).max() < 1e-12
    assert np.allclose(np.linalg.det(r), 1.0)


# ---- iterative refiner and cloud-pair sampler (tests/golden/make_golden_8f.py) ---------------------
def test_pose_update_matches_reference(golden):
    z = golden("iterative")
    tp = O.calculate_T_pred(T(z["out"]).double(), T(z["T_init"]).double())
    # the reference can only run this function in float32 (R_k.float(), utility.py:124): compare at fp32 accuracy
    assert np.abs(tp.numpy() - z["T_pred_ref_fp32"]).max() <= 2e-5
    o = T(z["out"]).double().requires_grad_(True)
    (g,) = torch.autograd.grad(O.calculate_T_pred(o, T(z["T_init"]).double()), o, T(z["g_T"]))
    ref = z["d_out_ref_fp32"]
    scale = np.abs(ref).max(axis=1, keepdims=True)
    assert (np.abs(g.numpy() - ref) / scale).max() <= 2e-3       # torch's fp32 svd_backward is the noisy side here


def test_add_l1_losses_match_reference(golden):
    z = golden("iterative")
    tg, tp, v = T(z["T_gt"]).double(), T(z["T_pred_ref_fp32"]).double(), T(z["verts"]).double()
    assert np.abs(O.transform_pts(tg, v).numpy() - z["transform_pts"]).max() <= 1e-14
    a = tp.clone().requires_grad_(True)
    add = O.compute_ADD_L1_loss(tg, a, v)
    (ga,) = torch.autograd.grad(add, a)
    assert abs(add.item() - float(z["add_l1"])) <= 1e-14 and np.abs(ga.numpy() - z["add_l1_grad_pred"]).max() <= 1e-14
    assert np.abs(O.compute_ADD_L1_loss(tg, tp, v, use_batch_mean=False).numpy() - z["add_l1_items"]).max() <= 1e-14
    a = tp.clone().requires_grad_(True)
    dis = O.compute_disentangled_ADD_L1_loss(a, tg, v)
    (gd,) = torch.autograd.grad(dis, a)
    assert abs(dis.item() - float(z["disentangled"])) <= 1e-14
    assert np.abs(gd.numpy() - z["disentangled_grad_pred"]).max() <= 1e-14


def test_sampler_map_matches_reference(golden):
    z = golden("sampler")
    m = O.rotation_matrices_from_axis_angle(T(z["theta"]).double(), T(z["axis"]).double())
    assert np.abs(m.numpy() - z["matrix"]).max() <= 2e-6          # the reference evaluates this in float32
    pc = torch.rand(8, 50, 3, dtype=torch.float64)
    r = m[:8]
    want = torch.bmm(r[:, None].expand(8, 50, 3, 3).reshape(-1, 3, 3), pc.reshape(-1, 3, 1)).view(8, 50, 3)
    assert np.abs(O.rotate_cloud(r, pc).numpy() - want.numpy()).max() <= 1e-15   # point_cloud/main.py:173-181


def test_round2_fixture_matches_reference(golden):
    """tests/golden/round2.npz (make_golden_r2.py): 4-D transform_pts, so3_exp_map with other clamp bounds, and the
    training step with its metric (Comparison/main.py:59-70)."""
    z = golden("round2")
    got = O.transform_pts(T(z["tp4_T"]).double(), T(z["tp4_pts"]).double()).numpy()
    assert got.shape == z["tp4_out"].shape == (6, 3, 50, 3) and np.abs(got - z["tp4_out"]).max() <= 1e-12
    for tag, eps in (("1em2", 1e-2), ("1em6", 1e-6)):
        v = T(z["exp_v"]).double().requires_grad_(True)
        r = O.so3_exp_map(v, eps)
        (g,) = torch.autograd.grad((r * T(z["exp_w"])).sum(), v)
        assert np.abs(r.detach().numpy() - z[f"exp_r_{tag}"]).max() <= 1e-12
        assert np.abs(g.numpy() - z[f"exp_g_{tag}"]).max() <= 1e-10
    x, rt = T(z["step_x"]), T(z["step_r_true"])
    r64 = O.symmetric_orthogonalization(x.double())
    assert abs(O.loss_frobenius(r64, rt.double()).item() - float(z["step_loss_f64"])) <= 1e-12
    assert np.abs(O.angle_error(r64, rt.double()).numpy() - z["step_angle_f64"]).max() <= 1e-8
    # the fp32 pipeline of the scripts: outR in fp32, angles in float64 from it
    r32 = O.symmetric_orthogonalization(x)
    assert abs(O.loss_frobenius(r32, rt).item() - float(z["step_loss_fp32"])) <= 1e-6
    assert np.abs(O.angle_error(r32, rt).numpy() - z["step_angle_fp32_pipeline"]).max() <= 1e-9
