"""Drop-in for the rotation-mapping part of the iterative (CosyPose-style) refiner, SURVEY.md 8(f) row 2.

Mirrors the names of /root/reference/Iterative/utility.py (``calculate_T_pred``, ``get_scene_parameters``,
``angle_error``) and /root/reference/Iterative/loss.py (``loss_frobenius``, ``transform_pts``,
``compute_ADD_L1_loss``, ``compute_disentangled_ADD_L1_loss``): same positional signatures, shapes and dtypes.
All tensors must be float32 CUDA tensors; there is no CPU path.

What changes underneath:
  * ``calculate_T_pred`` is one kernel (SVD orthogonalisation, dR R_k, translation update and the 4x4 assembly), with
    an analytic backward to the network output; the reference's ``combine`` (utility.py:63-71) reads undefined
    globals, its evident intent (rotation block, translation column, last row 0 0 0 1) is what is built;
  * the ADD-L1 losses read the (B, N, 3) vertices ONCE (12 B/vertex) instead of transforming them once per pose
    (2 transforms for the plain loss, 6 for the disentangled one), and the same pass yields the gradient.
"""
from __future__ import annotations

import torch

from . import ops as _ops
from .rotation_representation import angle_error, loss_frobenius  # noqa: F401  (same names)

__all__ = ["calculate_T_pred", "get_scene_parameters", "angle_error", "loss_frobenius", "transform_pts",
           "compute_ADD_L1_loss", "compute_disentangled_ADD_L1_loss"]


def get_scene_parameters():
    """Iterative/utility.py:73-88: (fx, fy) of the 320 px / 36 mm sensor / 50 mm lens camera (two values here; the
    function of the same name in rotation_representation.py returns four)."""
    sw, img_res, flen = 36, 320, 50
    ppm = sw / img_res
    fx = fy = flen / ppm
    return fx, fy


def calculate_T_pred(model_output, T_init, device=None, rot_repr='SVD'):
    """Iterative/utility.py:90-128.  model_output: (B, >=12) network output (9 rotation entries, vx, vy, vz);
    T_init: (B, 4, 4).  Returns T_pred (B, 4, 4); differentiable w.r.t. model_output (T_init is an input pose)."""
    # `rot_repr` is accepted and ignored, as in the reference (Iterative/utility.py:90-107 always orthogonalises)
    fx, fy = get_scene_parameters()
    out = model_output[:, :12]                       # strided view, read in place
    return _ops.pose_update(out, T_init.float(), float(fx), float(fy))


def transform_pts(T, pts):
    """Iterative/loss.py:51-69: (B, 4, 4) poses applied to (B, N, 3) points -> (B, N, 3)."""
    bsz = T.shape[0]
    n_pts = pts.shape[1]
    assert pts.shape == (bsz, n_pts, 3)                                  # :58
    if T.dim() == 4:                                                     # :59-61: K poses per item, (B, K, 4, 4)
        assert T.shape[-2:] == (4, 4) and T.shape[0] == bsz
        k = T.shape[1]
        tf = T.reshape(bsz * k, 4, 4)
        pf = pts.unsqueeze(1).expand(bsz, k, n_pts, 3).reshape(bsz * k, n_pts, 3)
        return _ops.affine_points(tf[:, :3, :3], tf[:, :3, 3], pf, False).reshape(bsz, k, n_pts, 3)
    elif T.dim() == 3:
        assert T.shape == (bsz, 4, 4)                                    # :63
    else:
        raise ValueError('Unsupported shape for T', T.shape)             # :65
    return _ops.affine_points(T[:, :3, :3], T[:, :3, 3], pts, False)


def compute_ADD_L1_loss(TCO_gt, TCO_pred, points, use_batch_mean=True):
    """Iterative/loss.py:10-28: mean_{vertex, xyz} |T_gt p - T_pred p|, then the batch mean (or per item)."""
    bsz = len(TCO_gt)
    assert TCO_pred.shape == (bsz, 4, 4) and TCO_gt.shape == (bsz, 4, 4)  # :18
    assert points.dim() == 3 and points.shape[-1] == 3                   # :19
    item, s, _ = _ops.add_l1(TCO_gt, TCO_pred, points, False)
    if use_batch_mean:
        return (s / bsz).to(torch.float32).reshape(())
    return item


def compute_disentangled_ADD_L1_loss(T_CO_pred, T_CO_gt, points):
    """Iterative/loss.py:31-48: rotation + xy-translation + depth ADD-L1 terms, each a batch mean.  NOTE the argument
    order (prediction first), as in the reference."""
    bsz = len(T_CO_gt)
    assert T_CO_pred.shape == (bsz, 4, 4) and T_CO_gt.shape == (bsz, 4, 4)
    assert points.dim() == 3 and points.shape[-1] == 3
    _, s, _ = _ops.add_l1(T_CO_gt, T_CO_pred, points, True)
    return (s / bsz).to(torch.float32).reshape(())
