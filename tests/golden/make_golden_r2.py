"""Generate tests/golden/round2.npz by running UNMODIFIED reference functions (same technique as make_golden.py /
make_golden_8f.py; run once in the build container, the output is committed because the GPU box has no /root/reference):

  * ``transform_pts`` of /root/reference/Iterative/loss.py:51-69 with the 4-D pose form (B, K, 4, 4), :59-61;
  * ``so3_exp_map(log_rot, eps)`` of /root/reference/rotation_representation.py:245-272 with eps != 1e-4, forward and
    the gradient of a random linear functional through the reference's own autograd;
  * the training step with its metric as Comparison/main.py:59-70 computes it: ``symmetric_orthogonalization`` (fp32, the
    dtype the scripts run in), ``loss_frobenius`` (:12-16, verbatim from the loaded source) and ``angle_error`` on the
    same outR; plus the float64 evaluation of the same chain.
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
    sys.path.insert(0, REF)
    import rotation_representation as rr
    spec = importlib.util.spec_from_file_location("ref_iter_loss", f"{REF}/Iterative/loss.py")
    L = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(L)
    (loss_frobenius,) = _extract(f"{REF}/Comparison/main.py", ["loss_frobenius"], {"torch": torch, "np": np})

    torch.manual_seed(20261106)
    res = {}
    # ---- transform_pts, 4-D
    B, K, N = 6, 3, 50
    T = torch.zeros(B, K, 4, 4, dtype=torch.float64)
    T[..., :3, :3] = rr.symmetric_orthogonalization(torch.randn(B * K, 9).double()).reshape(B, K, 3, 3)
    T[..., :3, 3] = torch.randn(B, K, 3).double()
    T[..., 3, 3] = 1
    pts = torch.rand(B, N, 3).double() - 0.5
    res["tp4_T"], res["tp4_pts"] = T.float().numpy(), pts.float().numpy()
    res["tp4_out"] = L.transform_pts(T.float().double(), pts.float().double()).numpy()
    # ---- so3_exp_map with other clamp bounds
    v = torch.randn(512, 3)
    v[:64] *= 1e-3          # rows below either bound
    v[64:128] *= 5e-2       # between the bounds
    res["exp_v"] = v.numpy()
    w = torch.randn(512, 3, 3).double()
    res["exp_w"] = w.numpy()
    for tag, eps in (("1em2", 1e-2), ("1em6", 1e-6)):
        vd = v.double().clone().requires_grad_(True)
        r = rr.so3_exp_map(vd, eps)
        (g,) = torch.autograd.grad((r * w).sum(), vd)
        res[f"exp_r_{tag}"], res[f"exp_g_{tag}"] = r.detach().numpy(), g.numpy()
    # ---- training step + metric
    n = 4096
    x = torch.randn(n, 9)
    r_true = rr.symmetric_orthogonalization(torch.randn(n, 9))
    out_r = rr.symmetric_orthogonalization(x)                      # fp32, as the scripts run it
    res["step_x"], res["step_r_true"] = x.numpy(), r_true.numpy()
    res["step_loss_fp32"] = loss_frobenius(out_r, r_true).numpy()
    res["step_angle_fp32_pipeline"] = rr.angle_error(out_r, r_true).numpy()      # float64 angles of the fp32 outR
    r64 = rr.symmetric_orthogonalization(x.double())
    res["step_loss_f64"] = loss_frobenius(r64, r_true.double()).numpy()
    res["step_angle_f64"] = rr.angle_error(r64, r_true.double()).numpy()
    np.savez_compressed(f"{OUT}/round2.npz", **res)
    print("round2.npz", os.path.getsize(f"{OUT}/round2.npz"))


if __name__ == "__main__":
    main()
