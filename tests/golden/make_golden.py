"""Generate tests/golden/*.npz by running the UNMODIFIED reference functions in this container.

Run once here (``python tests/golden/make_golden.py``); the GPU box has no /root/reference, so the outputs are
committed.  Nothing in the test-suite or the product imports this script.

How the reference code is executed without copying it:
  * ``rotation_representation.py`` is imported as-is from /root/reference (it needs only torch and numpy);
  * ``Tensor.cuda`` is patched to the identity for the duration of the run, which is the only way the hard-coded
    ``.cuda()`` sites (rotation_representation.py:84,129,218,219; point_cloud/prepare.py:23,25) can run on a
    machine without a GPU.  No arithmetic changes;
  * functions that live in un-importable scripts (module-level training code, missing third-party packages) are
    pulled out with ``ast`` and exec'd from their own source text at generation time:
    ``loss_frobenius`` (Comparison/main.py:12-16), ``geodesic`` (point_cloud/main.py:61-73),
    ``normalize_vector`` + ``get_sampled_rotation_matrices_by_axisAngle`` (point_cloud/prepare.py:9-49).

Parity definition (BASELINE.json north_star): the reference evaluated in float64 on the same fp32 inputs.
"""
import ast
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


def _extract(path, names):
    src = open(path).read()
    tree = ast.parse(src)
    ns = {"torch": torch, "np": np}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in names:
            exec(compile(ast.Module([node], []), path, "exec"), ns)
    return [ns[n] for n in names]


def _grad(fn, x64, g64):
    x = x64.clone().requires_grad_(True)
    y = fn(x)
    (gx,) = torch.autograd.grad(y, x, g64)
    return y.detach(), gx


def main():
    torch.Tensor.cuda = lambda self, *a, **k: self          # see module docstring
    sys.path.insert(0, REF)
    import rotation_representation as rr                    # the reference, unmodified

    loss_frobenius, = _extract(f"{REF}/Comparison/main.py", ["loss_frobenius"])
    geodesic_pc, = _extract(f"{REF}/point_cloud/main.py", ["geodesic"])
    normalize_vector_pc, sampler = _extract(f"{REF}/point_cloud/prepare.py",
                                            ["normalize_vector", "get_sampled_rotation_matrices_by_axisAngle"])
    sampler.__globals__["normalize_vector"] = normalize_vector_pc

    torch.manual_seed(20261104)
    np.random.seed(20261104)
    B = 512

    # ---------------------------------------------------------------- a1 SVD orthogonalisation
    x = torch.randn(B, 9)
    rots = rr.symmetric_orthogonalization(torch.randn(64, 9).double()).float().reshape(64, 9)
    special = torch.cat([
        torch.eye(3).reshape(1, 9),                                            # identity
        torch.diag(torch.tensor([1.0, 1.0, -1.0])).reshape(1, 9),              # pure reflection
        torch.tensor([[0.3598, 0.8291, -0.428, 0.0176, 0.4526, 0.8915, 0.9329, -0.3283, 0.1483]]),  # 3D-Pose/loss.py:28-30
        rots,                                                                  # proper rotations -> themselves
        rots + 0.1 * torch.randn(64, 9),                                       # what a trained net emits
        1e-3 * torch.randn(32, 9), 1e3 * torch.randn(32, 9),                   # scale invariance
    ])
    # rank-2 inputs (third singular value exactly representable as ~0): outer-product construction
    a, b = torch.randn(32, 3, 2), torch.randn(32, 2, 3)
    special = torch.cat([special, (a @ b).reshape(32, 9)])
    xs = torch.cat([x, special]).float()
    g = torch.randn(xs.shape[0], 3, 3).float()
    r64, dx64 = _grad(rr.symmetric_orthogonalization, xs.double(), g.double())
    m = xs.double().reshape(-1, 3, 3)
    u, s, v = torch.svd(m)
    det = torch.det(u @ v.transpose(1, 2))
    r32 = rr.symmetric_orthogonalization(xs)
    np.savez_compressed(f"{OUT}/svd_orth.npz", x=xs.numpy(), g=g.numpy(), r=r64.numpy(), dx=dx64.numpy(),
                        det=det.numpy(), s=s.numpy(), r_fp32=r32.numpy(), n_random=B)

    # ---------------------------------------------------------------- a2 6D
    x = torch.randn(B, 6).float()
    g = torch.randn(B, 3, 3).float()
    r64, dx64 = _grad(rr.compute_rotation_matrix_from_ortho6d, x.double(), g.double())
    np.savez_compressed(f"{OUT}/ortho6d.npz", x=x.numpy(), g=g.numpy(), r=r64.numpy(), dx=dx64.numpy())

    # ---------------------------------------------------------------- a3 quaternion (last 8 rows hit the 1e-8 floor)
    x = torch.randn(B, 4).float()
    x[-8:] = x[-8:] * 1e-10
    g = torch.randn(B, 3, 3).float()
    r64, dx64 = _grad(rr.compute_rotation_matrix_from_quaternion, x.double(), g.double())
    np.savez_compressed(f"{OUT}/quaternion.npz", x=x.numpy(), g=g.numpy(), r=r64.numpy(), dx=dx64.numpy())

    # ---------------------------------------------------------------- a4 Euler
    x = ((torch.rand(B, 3) * 2 - 1) * np.pi).float()
    g = torch.randn(B, 3, 3).float()
    r64, dx64 = _grad(rr.compute_rotation_matrix_from_euler, x.double(), g.double())
    np.savez_compressed(f"{OUT}/euler.npz", x=x.numpy(), g=g.numpy(), r=r64.numpy(), dx=dx64.numpy())
    # KAT: Comparison/dataset.py:7-20 getR(x, y, z) is the same formula with angle order (x, z, y) -> (e0, e2, e1)
    getR, = _extract(f"{REF}/Comparison/dataset.py", ["getR"])
    ang = x[:16].double().numpy()
    kat = np.stack([getR(a0, a2, a1) for a0, a1, a2 in ang])
    np.savez_compressed(f"{OUT}/euler_getR.npz", x=x[:16].numpy(), r=kat)

    # ---------------------------------------------------------------- 8f: 5D and exponential map
    x = torch.randn(B, 5).float()
    g = torch.randn(B, 3, 3).float()
    r64, dx64 = _grad(rr.compute_rotation_matrix_from_ortho5d, x.double(), g.double())
    np.savez_compressed(f"{OUT}/ortho5d.npz", x=x.numpy(), g=g.numpy(), r=r64.numpy(), dx=dx64.numpy())
    x = torch.randn(B, 3).float()
    x[-16:] *= 1e-3                                                            # below the eps=1e-4 clamp
    r64, dx64 = _grad(rr.vec_3d_to_SO3, x.double(), g.double())
    np.savez_compressed(f"{OUT}/expmap.npz", x=x.numpy(), g=g.numpy(), r=r64.numpy(), dx=dx64.numpy())

    # ---------------------------------------------------------------- a5-a8 losses and distances
    r1 = rr.symmetric_orthogonalization(torch.randn(B, 9).double()).float()
    r2 = rr.symmetric_orthogonalization(torch.randn(B, 9).double()).float()
    r2[:32] = r1[:32] + 1e-4 * torch.randn(32, 3, 3)                           # tiny angles
    r2[32:40] = r1[32:40]                                                      # zero angle / zero Frobenius distance
    direct = (r1 + 0.02 * torch.randn(B, 3, 3)).float()                        # 'Direct' representation: not a rotation
    out = {}
    out["r1"], out["r2"], out["direct"] = r1.numpy(), r2.numpy(), direct.numpy()
    out["geo_rr"] = rr.compute_geodesic_distance_from_two_matrices(r1.double(), r2.double()).numpy()
    out["geo_rr_direct"] = rr.compute_geodesic_distance_from_two_matrices(direct.double(), r2.double()).numpy()
    out["angle_error"] = rr.angle_error(r1, r2).numpy()                        # reference casts to float64 itself
    out["angle_error_direct"] = rr.angle_error(direct, r2).numpy()
    for red in ("none", "mean", "sum"):
        out[f"geo_pc_{red}"] = np.asarray(geodesic_pc(r1.double(), r2.double(), red).numpy())
    a = direct.double().clone().requires_grad_(True)
    bb = r2.double().clone().requires_grad_(True)
    val = loss_frobenius(a, bb)
    ga, gb = torch.autograd.grad(val, (a, bb))
    out["frob"], out["frob_grad_pred"], out["frob_grad_true"] = val.detach().numpy(), ga.numpy(), gb.numpy()
    # geodesic (point_cloud) gradient wrt the first argument, mean reduction, away from the clamp
    a = direct[64:].double().clone().requires_grad_(True)
    val = geodesic_pc(a, r2[64:].double(), "mean")
    (ga,) = torch.autograd.grad(val, a)
    out["geo_pc_grad_first"] = ga.numpy()
    a = direct[64:].double().clone().requires_grad_(True)
    val = rr.compute_geodesic_distance_from_two_matrices(a, r2[64:].double())
    w = torch.randn(val.shape[0]).double()
    (ga,) = torch.autograd.grad(val, a, w)
    out["geo_rr_w"], out["geo_rr_grad_first"] = w.numpy(), ga.numpy()
    # angle_error must raise for improper input
    bad = (3.0 * r1[:4]).float()
    try:
        rr.angle_error(bad, r2[:4])
        raised = False
    except ValueError:
        raised = True
    out["angle_error_raises_bad"], out["bad"] = np.asarray(raised), bad.numpy()
    np.savez_compressed(f"{OUT}/distances.npz", **out)

    # ---------------------------------------------------------------- a9 Kabsch KAT from the reference's data construction
    batch, point_num = 16, 64
    pc1 = (torch.rand(batch, point_num, 3) - 0.5).float()                      # unit box like pc_normalize (prepare.py:51-56)
    gt_rmat = sampler(batch)                                                   # prepare.py:21-49, unmodified
    gt_rmats = gt_rmat.contiguous().view(batch, 1, 3, 3).expand(batch, point_num, 3, 3).contiguous().view(-1, 3, 3)
    pc2 = torch.bmm(gt_rmats, pc1.view(-1, 3, 1)).view(batch, point_num, 3)    # point_cloud/main.py:176-181
    np.savez_compressed(f"{OUT}/kabsch_kat.npz", pc1=pc1.numpy(), pc2=pc2.numpy(), gt_rmat=gt_rmat.numpy())

    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(f"{OUT}/{f}"))


if __name__ == "__main__":
    main()
