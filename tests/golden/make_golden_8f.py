"""Generate tests/golden/{iterative,sampler}.npz by running the UNMODIFIED reference functions of the SURVEY 8(f) rows.

Run once in the build container (``python tests/golden/make_golden_8f.py``); the outputs are committed because the GPU
box has no /root/reference.  Same technique as make_golden.py:
  * /root/reference/Iterative/loss.py imports only torch and is loaded as a module, as is;
  * ``calculate_T_pred`` and ``get_scene_parameters`` are pulled out of Iterative/utility.py with ``ast`` (the file
    itself imports matplotlib, torchvision and the dataset) and exec'd from their own source text, with the
    reference's ``symmetric_orthogonalization`` in scope.  The helper ``combine`` they call (utility.py:63-71) reads
    undefined globals and cannot run: it is replaced, for the generation only, by a stand-in that does what its body
    evidently means (ones, rotation block, translation column, T[:, 3, :3] = 0).  That stand-in is the one piece of
    this fixture that is not reference code;
  * the sampler of point_cloud/prepare.py:21-49 draws its random numbers inside the function; ``np.random.uniform``
    and ``torch.randn`` are wrapped to RECORD what they return, so the fixture holds (theta, axis, matrix).
"""
import ast
import importlib.util
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


def _extract(path, names, ns):
    tree = ast.parse(open(path).read())
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in names:
            exec(compile(ast.Module([node], []), path, "exec"), ns)
    return [ns[n] for n in names]


def main():
    torch.Tensor.cuda = lambda self, *a, **k: self
    sys.path.insert(0, REF)
    import rotation_representation as rr

    spec = importlib.util.spec_from_file_location("ref_iter_loss", f"{REF}/Iterative/loss.py")
    L = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(L)

    def combine(R, tx, ty, tz, device='cpu'):                 # stand-in, see the module docstring
        T = torch.ones((R.shape[0], 4, 4), dtype=R.dtype)
        T[:, :3, :3] = R
        T[:, 0, 3], T[:, 1, 3], T[:, 2, 3] = tx, ty, tz
        T[:, 3, :3] = 0
        return T

    ns = {"torch": torch, "np": np, "symmetric_orthogonalization": rr.symmetric_orthogonalization, "combine": combine}
    get_scene_parameters, calculate_T_pred = _extract(f"{REF}/Iterative/utility.py",
                                                      ["get_scene_parameters", "calculate_T_pred"], ns)

    torch.manual_seed(20261105)
    np.random.seed(20261105)
    B, N = 48, 200
    # poses like the dataset's: a rotation, a translation in front of the camera (z in [1.5, 3])
    def poses(b):
        T = torch.zeros(b, 4, 4, dtype=torch.float64)
        T[:, :3, :3] = rr.symmetric_orthogonalization(torch.randn(b, 9).double())
        T[:, :2, 3] = 0.3 * torch.randn(b, 2).double()
        T[:, 2, 3] = 1.5 + 1.5 * torch.rand(b).double()
        T[:, 3, 3] = 1
        return T
    T_init, T_gt = poses(B).float(), poses(B).float()
    out = torch.randn(B, 12).float()
    out[:, :9] = torch.eye(3).reshape(1, 9) + 0.3 * out[:, :9]          # a refiner emits near-identity updates
    out[:, 9:11] *= 20.0                                                  # pixel-scale vx, vy
    out[:, 11] = 1.0 + 0.1 * out[:, 11]                                   # depth ratio around 1
    verts = (torch.rand(B, N, 3) - 0.5).float()

    o64 = out.double().clone().requires_grad_(True)
    # the reference casts R_k to float32 (utility.py:124: R_k.float()); for the float64 evaluation R_k.float() would
    # mix dtypes inside einsum, so T_init is passed as float32-representable float64 values and .float() is a no-op
    # on the value; einsum needs equal dtypes -> evaluate the reference in float32 AND our float64 restatement
    T32 = calculate_T_pred(out, T_init, 'cpu')
    g_T = torch.randn(B, 4, 4).double()
    res = {"out": out.numpy(), "T_init": T_init.numpy(), "T_gt": T_gt.numpy(), "verts": verts.numpy(),
           "T_pred_ref_fp32": T32.numpy(), "g_T": g_T.numpy()}

    # losses: the reference functions, unmodified, in float64 on the fp32 inputs
    Tp = T32.double().clone().requires_grad_(True)
    add = L.compute_ADD_L1_loss(T_gt.double(), Tp, verts.double())
    (g_add,) = torch.autograd.grad(add, Tp)
    res["add_l1"], res["add_l1_grad_pred"] = add.detach().numpy(), g_add.numpy()
    res["add_l1_items"] = L.compute_ADD_L1_loss(T_gt.double(), Tp.detach(), verts.double(), use_batch_mean=False).numpy()
    Tp = T32.double().clone().requires_grad_(True)
    dis = L.compute_disentangled_ADD_L1_loss(Tp, T_gt.double(), verts.double())
    (g_dis,) = torch.autograd.grad(dis, Tp)
    res["disentangled"], res["disentangled_grad_pred"] = dis.detach().numpy(), g_dis.numpy()
    res["transform_pts"] = L.transform_pts(T_gt.double(), verts.double()).numpy()
    # gradient of calculate_T_pred through the reference's own autograd, float32 (its only runnable dtype)
    o32 = out.clone().requires_grad_(True)
    T32g = calculate_T_pred(o32, T_init, 'cpu')
    (g_o,) = torch.autograd.grad(T32g, o32, g_T.float())
    res["d_out_ref_fp32"] = g_o.numpy()
    np.savez_compressed(f"{OUT}/iterative.npz", **res)

    # ---------------------------------------------------------------- sampler with recorded random numbers
    ns2 = {"torch": torch, "np": np}
    normalize_vector_pc, sampler = _extract(f"{REF}/point_cloud/prepare.py",
                                            ["normalize_vector", "get_sampled_rotation_matrices_by_axisAngle"], ns2)
    rec = {}
    real_uniform, real_randn = np.random.uniform, torch.randn

    def rec_uniform(*a, **k):
        v = real_uniform(*a, **k)
        rec["u"] = np.array(v)
        return v

    def rec_randn(*a, **k):
        v = real_randn(*a, **k)
        rec["axis"] = v.clone()
        return v
    np.random.uniform, torch.randn = rec_uniform, rec_randn
    try:
        mat = sampler(256)
    finally:
        np.random.uniform, torch.randn = real_uniform, real_randn
    theta = torch.FloatTensor(rec["u"] * np.pi)                         # prepare.py:23, same expression
    np.savez_compressed(f"{OUT}/sampler.npz", theta=theta.numpy(), axis=rec["axis"].numpy(), matrix=mat.numpy())
    for f in ("iterative.npz", "sampler.npz"):
        print(f, os.path.getsize(f"{OUT}/{f}"))


if __name__ == "__main__":
    main()
