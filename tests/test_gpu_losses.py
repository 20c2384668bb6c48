"""GPU parity for the SO(3) distances / losses (BASELINE config 3) and Kabsch (config 4)."""
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


def test_distances_match_reference_golden(rr, golden):
    z = golden("distances")
    r1, r2, direct = T(z["r1"]).cuda(), T(z["r2"]).cuda(), T(z["direct"]).cuda()
    geo = rr.compute_geodesic_distance_from_two_matrices(r1, r2)
    assert geo.dtype == torch.float32 and geo.shape == (512,)
    # acos amplifies fp32 rounding of the cosine near 0 and pi: 1e-3 rad absolute there, 1e-5 elsewhere
    mid = (z["geo_rr"] > 0.05) & (z["geo_rr"] < np.pi - 0.05)
    assert np.abs(geo.cpu().numpy() - z["geo_rr"])[mid].max() <= 1e-5
    assert np.abs(geo.cpu().numpy() - z["geo_rr"]).max() <= 1e-3
    gd = rr.compute_geodesic_distance_from_two_matrices(direct, r2).cpu().numpy()
    assert np.abs(gd - z["geo_rr_direct"]).max() <= 1e-3
    ae = rr.angle_error(r1, r2)
    assert ae.dtype == torch.float64 and ae.shape == (512,)
    assert np.abs(ae.cpu().numpy() - z["angle_error"]).max() <= 1e-9        # float64 arithmetic on the same fp32 inputs
    assert np.abs(rr.angle_error(direct, r2).cpu().numpy() - z["angle_error_direct"]).max() <= 1e-9
    assert abs(rr.mean_angle_error(r1, r2).item() - z["angle_error"].mean()) <= 1e-9   # north_star: 1e-4 degrees
    none = rr.geodesic(r1, r2, "none").cpu().numpy()
    assert np.abs(none - z["geo_pc_none"])[mid].max() <= 1e-5
    assert abs(rr.geodesic(r1, r2, "mean").item() - float(z["geo_pc_mean"])) <= 1e-5
    assert abs(rr.geodesic(r1, r2, "sum").item() - float(z["geo_pc_sum"])) <= 512 * 1e-5
    assert rr.geodesic(r1, r2, "bogus") is None
    with pytest.raises(ValueError, match="angle out of range"):
        rr.angle_error(T(z["bad"]).cuda(), r2[:4])
    with pytest.raises(ValueError):
        rr.mean_angle_error(T(z["bad"]).cuda(), r2[:4])
    assert torch.isfinite(rr.mean_angle_error(T(z["bad"]).cuda(), r2[:4], check=False))


def test_frobenius_matches_reference_golden(rr, golden):
    z = golden("distances")
    a = T(z["direct"]).cuda().requires_grad_(True)
    b = T(z["r2"]).cuda().requires_grad_(True)
    v = rr.loss_frobenius(a, b)
    v.backward()
    assert v.dtype == torch.float32 and v.dim() == 0
    assert abs(v.item() - float(z["frob"])) <= 1e-6
    assert np.abs(a.grad.cpu().numpy() - z["frob_grad_pred"]).max() <= 1e-8
    assert np.abs(b.grad.cpu().numpy() - z["frob_grad_true"]).max() <= 1e-8
    # only one side requires grad; identical inputs give a zero (sub)gradient like torch's norm
    a2 = T(z["direct"]).cuda().requires_grad_(True)
    rr.loss_frobenius(a2, T(z["r2"]).cuda()).backward()
    assert torch.equal(a2.grad, a.grad)
    c = T(z["r1"]).cuda().requires_grad_(True)
    rr.loss_frobenius(c, T(z["r1"]).cuda()).backward()
    assert float(c.grad.abs().max()) == 0.0


def test_geodesic_gradients(rr, golden):
    z = golden("distances")
    w = z["geo_rr_w"]
    n = len(w)
    r1 = T(z["direct"][64:]).cuda().requires_grad_(True)      # the golden gradients are for (direct[64:], r2[64:])
    r2 = T(z["r2"][64:]).cuda()
    assert len(z["direct"]) - 64 == n
    th = rr.compute_geodesic_distance_from_two_matrices(r1, r2)
    (th * T(w).float().cuda()).sum().backward()
    ref = z["geo_rr_grad_first"]
    # d theta/d cos = -1/sin(theta): relative error grows like eps/sin^2; the golden subset excludes theta near 0, pi
    err = np.abs(r1.grad.cpu().numpy() - ref).reshape(n, -1).max(1) / (np.abs(ref).reshape(n, -1).max(1) + 1e-30)
    assert np.quantile(err, 0.9) <= 1e-4
    r1b = T(z["direct"][64:]).cuda().requires_grad_(True)
    rr.geodesic(r1b, r2, "mean").backward()
    refm = z["geo_pc_grad_first"]
    errm = np.abs(r1b.grad.cpu().numpy() - refm).reshape(n, -1).max(1) / (np.abs(refm).reshape(n, -1).max(1) + 1e-30)
    assert np.quantile(errm, 0.9) <= 1e-4


@pytest.mark.parametrize("n", [1, 257, 100003])
def test_distances_vs_oracle(rr, n):
    g = torch.Generator().manual_seed(n)
    r1 = O.sample_rotations_axis_angle(n, g).float()
    r2 = O.sample_rotations_axis_angle(n, g).float()
    ae = rr.angle_error(r1.cuda(), r2.cuda()).cpu()
    ref = O.angle_error(r1, r2)
    assert (ae - ref).abs().max() <= 1e-9
    assert abs(rr.mean_angle_error(r1.cuda(), r2.cuda()).item() - ref.mean().item()) <= 1e-9
    fr = rr.loss_frobenius(r1.cuda(), r2.cuda()).item()
    assert abs(fr - O.loss_frobenius(r1.double(), r2.double()).item()) <= 2e-6
    gm = rr.geodesic(r1.cuda(), r2.cuda(), "mean").item()
    assert abs(gm - O.geodesic(r1.double(), r2.double(), "mean").item()) <= 1e-5


def test_kabsch_known_answer_golden(rr, golden):
    z = golden("kabsch_kat")
    r, t = rr.kabsch(T(z["pc1"]).cuda(), T(z["pc2"]).cuda())
    assert np.abs(r.cpu().numpy() - z["gt_rmat"]).max() <= 1e-5
    assert np.abs(t.cpu().numpy()).max() <= 1e-5
    # far from the origin and translated: the shift-by-first-point accumulation must not cancel
    shift = torch.tensor([10.0, -7.0, 3.0])
    r2, t2 = rr.kabsch(T(z["pc1"]).cuda() + 5.0, T(z["pc2"]).cuda() + shift.cuda())
    assert np.abs(r2.cpu().numpy() - z["gt_rmat"]).max() <= 2e-5


@pytest.mark.parametrize("b,npts", [(1, 4), (3, 64), (40, 1024), (33, 1000), (5, 1023), (2000, 256), (7, 3)])
def test_kabsch_vs_oracle(rr, b, npts):
    g = torch.Generator().manual_seed(b * 10007 + npts)
    p = torch.rand(b, npts, 3, generator=g) - 0.5
    rot = O.sample_rotations_axis_angle(b, g).float()
    q = p @ rot.transpose(1, 2) + 0.01 * torch.randn(b, npts, 3, generator=g) + torch.randn(b, 1, 3, generator=g)
    r, t = rr.kabsch(p.cuda(), q.cuda())
    r_ref, t_ref = O.kabsch(p, q)
    if npts >= 4:
        # north_star's 1e-5 wherever the cloud determines the rotation well (N >= 64 points); a handful of noisy points
        # leaves H = sum (q - qbar)(p - pbar)^T badly conditioned and every fp32 method loses digits there
        tol = 1e-5 if npts >= 64 else 2e-5
        assert (r.cpu().double() - r_ref).abs().max() <= tol
        assert (t.cpu().double() - t_ref).abs().max() <= 2e-5     # |t| ~ 1-3: its own fp32 rounding plus |pbar| * err(R)
    rd = r.cpu().double()
    assert (rd @ rd.transpose(1, 2) - torch.eye(3, dtype=torch.float64)).abs().max() <= 1e-5
    assert (torch.linalg.det(rd) - 1).abs().max() <= 1e-5


def test_kabsch_full_size_round_trip(rr):
    """BASELINE config 4 size: 65536 pairs x 1024 points; rotate -> kabsch recovers the rotation."""
    b, npts = 65536, 1024
    torch.manual_seed(0)
    p = torch.rand(b, npts, 3, device="cuda") - 0.5
    g = torch.Generator().manual_seed(1)
    rot = O.sample_rotations_axis_angle(b, g).float().cuda()
    q = torch.bmm(p, rot.transpose(1, 2))
    r, t = rr.kabsch(p, q)
    assert float((r - rot).abs().max()) <= 1e-5
    assert float(t.abs().max()) <= 1e-5
    # float64 acos of a trace built from fp32 entries has a ~0.03 degree floor near zero angle (sqrt(2 * 3e-7) rad)
    assert float(rr.mean_angle_error(r, rot).item()) <= 0.05


def test_kabsch_degenerate_clouds_give_finite_proper_rotations(rr):
    """Clouds that do not determine a rotation - one point, all points equal, collinear points, a planar cloud against its
    mirror image (the determinant fix decides), zero-size coordinates, huge coordinates: the result must still be a
    finite proper rotation with a finite translation (the reference's SVD route returns SOME completion there), for
    the plain and the weighted form, forward and backward."""
    g = torch.Generator(device="cuda").manual_seed(17)
    npts = 96
    line = torch.linspace(-1, 1, npts, device="cuda").view(1, npts, 1) * torch.tensor([[[0.3, -0.5, 0.8]]], device="cuda")
    plane = torch.randn(1, npts, 3, device="cuda", generator=g) * torch.tensor([[[1.0, 1.0, 0.0]]], device="cuda")
    mirror = plane * torch.tensor([[[1.0, -1.0, 1.0]]], device="cuda")
    same = torch.full((1, npts, 3), 0.25, device="cuda")
    cases = {
        "one point": (torch.randn(2, 1, 3, device="cuda", generator=g), torch.randn(2, 1, 3, device="cuda", generator=g)),
        "all equal": (same, same + 1.0),
        "collinear": (line, line.flip(1) + 2.0),
        "planar vs mirror": (plane, mirror),
        "zeros": (torch.zeros(1, npts, 3, device="cuda"), torch.zeros(1, npts, 3, device="cuda")),
        # the covariance H is a sum of squares of the coordinates and its pullback divides by it: float32 carries that
        # for coordinates of 1e-12 .. 1e15 (at 1e-20 H is denormal and dH = O(1 / H) overflows - as the 1 / (s_i^2 - s_j^2)
        # of torch's own svd backward does)
        "tiny": (1e-12 * torch.randn(1, npts, 3, device="cuda", generator=g), 1e-12 * torch.randn(1, npts, 3, device="cuda", generator=g)),
        "huge": (1e15 * torch.randn(1, npts, 3, device="cuda", generator=g), 1e15 * torch.randn(1, npts, 3, device="cuda", generator=g)),
    }
    eye = torch.eye(3, device="cuda", dtype=torch.float64)
    for name, (p, q) in cases.items():
        for weights in (None, torch.rand(p.shape[0], p.shape[1], device="cuda", generator=g) + 0.1):
            pg, qg = p.clone().requires_grad_(True), q.clone().requires_grad_(True)
            r, t = rr.kabsch(pg, qg, weights=weights)
            rd = r.double()
            assert torch.isfinite(r).all() and torch.isfinite(t).all(), name
            assert (rd @ rd.transpose(1, 2) - eye).abs().max().item() <= 1e-5, name
            assert (torch.linalg.det(rd) - 1).abs().max().item() <= 1e-5, name
            (r.sum() + t.sum()).backward()
            assert torch.isfinite(pg.grad).all() and torch.isfinite(qg.grad).all(), name
    # the mirror case has a well-defined answer in the plane: a planar cloud and its reflection across the x axis are
    # matched exactly by the half turn about x (which also flips z, invisible to a planar cloud)
    r, t = rr.kabsch(plane, mirror)
    assert (torch.bmm(plane, r.transpose(1, 2)) + t.view(1, 1, 3) - mirror).abs().max().item() <= 1e-5
