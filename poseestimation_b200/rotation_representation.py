"""Drop-in for the reference's ``rotation_representation`` module (/root/reference/rotation_representation.py).

Same function names, positional signatures, output shapes and dtypes, exceptions - but each mapping / distance is ONE
hand-written sm_100a kernel per direction (``poseestimation_b200.ops`` -> ``libpe_rotation.so``) instead of 8-30 small
PyTorch launches.  How a reference script picks this up: put this package's directory ahead of the
reference's on ``sys.path`` so that ``from rotation_representation import *`` (Comparison/models.py:11,
UPNA/main.py:23, 3D-Pose/main.py:23, Iterative/utility.py:4) resolves here, and call :func:`install` on modules that
carry private copies of the functions (3D-Pose/main2.py:34-62, point_cloud/main.py:43-155, ...).

Inputs must be float32 CUDA tensors: there is no CPU path (a CPU tensor raises ``NotImplementedError``), by design.
Additions that are not in the reference: ``loss_frobenius`` and ``geodesic`` (the reference keeps copies of them in
its experiment scripts), ``kabsch``, ``svd_frobenius_loss`` (fused map + loss) and ``mean_angle_error``.
"""
from __future__ import annotations

import math  # noqa: F401

import numpy as np  # noqa: F401  (the reference module's star-import hands `np` and `torch` to its importers)
import torch

try:  # importable both as poseestimation_b200.rotation_representation and as a top-level rotation_representation
    from . import ops as _ops
except ImportError:  # pragma: no cover - top-level import through sys.path
    import os as _os
    import sys as _sys
    _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
    from poseestimation_b200 import ops as _ops

# No __all__ on purpose: the reference module has none, so `from rotation_representation import *` (Comparison/models.py:11,
# UPNA/main.py:23, 3D-Pose/main.py:23) also hands `torch`, `np`, `nn`, `os`, `sys` to the importing script, and some of
# them rely on that.  Public names of this module = everything without a leading underscore.
import os  # noqa: E402,F401
import sys  # noqa: E402,F401
import torch.functional as F  # noqa: E402,F401
import torch.nn as nn  # noqa: E402,F401
from os.path import join as pjoin  # noqa: E402,F401


# ---------------------------------------------------------------------------------------------------------------------
# maps R^d -> SO(3)
def symmetric_orthogonalization(x):
    """Nearest rotation to each 3x3 block of ``x`` (rotation_representation.py:192-206).

    ``x`` is anything ``.view(-1, 3, 3)`` accepts there: (B, 9), (B, 3, 3), or the ``out[:, :9]`` slice of a wider
    network output, which is read in place through its row stride.  Returns (B, 3, 3).
    """
    if x.dim() == 2 and x.shape[1] == 9:
        rec = x                                   # possibly strided rows: no copy
    else:
        rec = x.reshape(-1, 9)                    # reference: x.view(-1, 3, 3)
    if (_ops._C is not None and rec.requires_grad and type(rec) in (torch.Tensor, torch.nn.Parameter) and rec.is_cuda
            and rec.dtype == torch.float32 and torch.is_grad_enabled() and not torch.compiler.is_compiling()):
        return _ops._C.svd_orth(rec)              # C++ autograd node (csrc/pe_torch.cpp): same kernels, no Python frames
    r = _ops.svd_orth(rec)
    if r.requires_grad and not torch.compiler.is_compiling():
        _ops.tag_svd_output(r, rec)               # lets loss_frobenius(r, R_true) take the single-pass step kernel
    return r


def _plain_cuda_f32(t) -> bool:
    return type(t) in (torch.Tensor, torch.nn.Parameter) and t.is_cuda and t.dtype == torch.float32


def _cpp_nodes() -> bool:
    """the C++ autograd nodes of csrc/pe_torch.cpp serve eager calls that will be differentiated; everything else (no
    gradient, traced graphs, tensor subclasses, argument errors) goes through the registered ops of ops.py"""
    return _ops._C is not None and torch.is_grad_enabled() and not torch.compiler.is_compiling()


def _map(x2d, map_id: int, op, eps: float = 0.0):
    if x2d.requires_grad and _plain_cuda_f32(x2d) and _cpp_nodes():
        return _ops._C.map_fwd(x2d, map_id, eps)    # same kernels, no Python frames in forward / backward
    return op(x2d) if eps == 0.0 else op(x2d, eps)


def _pair_for_cpp(a, b) -> bool:
    return (_plain_cuda_f32(a) and _plain_cuda_f32(b) and (a.requires_grad or b.requires_grad) and a.device == b.device
            and a.shape == b.shape and (a.shape[1:] == (3, 3) or (a.dim() == 2 and a.shape[1] == 9)) and _cpp_nodes())


def compute_rotation_matrix_from_ortho6d(poses):
    """Gram-Schmidt 6D map, columns (x, y, z) (rotation_representation.py:21-36 = 174-189).  [..., 6] -> [..., 3, 3]."""
    assert poses.shape[-1] == 6                   # :28
    lead = poses.shape[:-1]
    return _map(poses.reshape(-1, 6), 1, _ops.ortho6d).reshape(*lead, 3, 3)


def compute_rotation_matrix_from_quaternion(quaternion):
    """(B, 4) quaternions (w, x, y, z), norm floored at 1e-8 -> (B, 3, 3) (rotation_representation.py:139-172)."""
    return _map(quaternion.reshape(-1, 4), 2, _ops.quat)


def compute_rotation_matrix_from_euler(euler):
    """(B, 3) -> (B, 3, 3) with the reference's angle order (rotation_representation.py:92-115)."""
    return _map(euler.reshape(-1, 3), 3, _ops.euler)


def compute_rotation_matrix_from_ortho5d(a):
    """(B, 5) -> (B, 3, 3) (rotation_representation.py:120-136)."""
    return _map(a.reshape(-1, 5), 4, _ops.ortho5d)


def so3_exp_map(log_rot, eps: float = 0.0001):
    """Rodrigues formula with |v|^2 clamped at ``eps`` (rotation_representation.py:245-272)."""
    if log_rot.dim() != 2 or log_rot.shape[1] != 3:
        raise ValueError("Input tensor shape has to be Nx3.")           # :255-256
    if eps == 0.0001:
        return _map(log_rot, 5, _ops.expmap)
    return _map(log_rot, 5, _ops.expmap_eps, float(eps))


def vec_3d_to_SO3(x):
    """rotation_representation.py:306-315."""
    assert x.shape == (x.shape[0], 3)             # :311
    return so3_exp_map(x)


# small helpers that sit on the module surface but not on the hot path (used only inside the reference's own maps,
# which are fused kernels here); kept for callers that import them
def normalize_vector(v):
    """v / max(|v|, 1e-8) row-wise (rotation_representation.py:38-50)."""
    mag = torch.sqrt(v.pow(2).sum(1)).clamp_min(1e-8)
    return v / mag.view(-1, 1)


def cross_product(u, v):
    """rotation_representation.py:53-69."""
    return torch.linalg.cross(u, v, dim=1)


def stereographic_unproject(a, axis=None):
    """rotation_representation.py:72-90."""
    batch, dim = a.shape
    axis = dim if axis is None else axis
    s2 = a.pow(2).sum(1, keepdim=True)
    un = 2 * a / (s2 + 1)
    return torch.cat((un[:, :axis], (s2 - 1) / (s2 + 1), un[:, axis:]), dim=1)


def hat(v):
    """(N, 3) -> (N, 3, 3) skew matrices (rotation_representation.py:275-303)."""
    n, dim = v.shape
    if dim != 3:
        raise ValueError("Input vectors have to be 3-dimensional.")
    x, y, z = v.unbind(1)
    o = torch.zeros_like(x)
    return torch.stack((o, -z, y, z, o, -x, -y, x, o), dim=1).view(n, 3, 3)


# ---------------------------------------------------------------------------------------------------------------------
# distances and losses on SO(3)
def compute_geodesic_distance_from_two_matrices(m1, m2):
    """acos(clamp((tr(m1 m2^T) - 1)/2, -1, 1)), radians, (B,) float32 (rotation_representation.py:209-227)."""
    if _pair_for_cpp(m1, m2):
        return _ops._C.geodesic(m1, m2, -1.0, 1.0, 0)
    return _ops.geodesic_theta(m1, m2, -1.0, 1.0)


def geodesic(R1, R2, reduction="mean"):
    """point_cloud/main.py:61-73: clamp at +-(1 - 1e-7); ``none`` | ``mean`` | ``sum``."""
    eps = 1e-7
    if reduction in ("none", "mean", "sum") and _pair_for_cpp(R1, R2):
        return _ops._C.geodesic(R1, R2, -1 + eps, 1 - eps, ("none", "mean", "sum").index(reduction))
    if reduction == "none":
        return _ops.geodesic_theta(R1, R2, -1 + eps, 1 - eps)
    if reduction == "mean":
        return _ops.geodesic_reduce(R1, R2, -1 + eps, 1 - eps, True)
    if reduction == "sum":
        return _ops.geodesic_reduce(R1, R2, -1 + eps, 1 - eps, False)
    return None                                   # the reference falls off the end of the function


def angle_error(t_R1, t_R2):
    """Angle between rotations in DEGREES, computed in float64, (B,) float64 (rotation_representation.py:230-242).

    Raises ``ValueError`` if any cosine falls outside [-1.1, 1.1] (:237-239) - this reads one int back from the
    device, the same synchronisation the reference's ``torch.any`` incurs.
    """
    deg, flag = _ops.angle_error_deg(t_R1, t_R2)
    if int(flag.item()) & 1:
        raise ValueError("angle out of range, input probably not proper rotation matrices")
    return deg


def mean_angle_error(R1, R2, check: bool = True):
    """``angle_error(R1, R2).mean()`` without materialising the per-matrix angles: 0-d float64 on the device.

    With ``check=False`` nothing synchronises (the caller's ``.item()`` is the only sync)."""
    s, flag = _ops.angle_error_sum(R1, R2)
    if check and int(flag.item()) & 1:
        raise ValueError("angle out of range, input probably not proper rotation matrices")
    n = R1.reshape(-1, 9).shape[0]
    return (s / n).reshape(())


_MAP_KEYS = {"SVD": "svd", "6D": "ortho6d", "5D": "ortho5d", "3D": "expmap", "Quat": "quat", "quat": "quat",
             "Euler": "euler", "euler": "euler"}


def map_metric(rot_rep, x, R_true, metric="angle_error", reduction="mean", check: bool = True):
    """``metric(transform(x), R_true)`` reduced over the batch in ONE pass, the rotations never written.

    ``rot_rep``: a key of the reference's dispatch dicts (``'SVD'``, ``'6D'``, ``'5D'``, ``'3D'``, ``'Quat'``,
    ``'Euler'``: rotation_representation.py:323-324, Comparison/models.py:18-19, 3D-Pose/main.py:46).
    ``metric``: ``'angle_error'`` (degrees, float64, as :230-242 and its ``.mean()`` at the call sites; raises
    ``ValueError`` like the reference unless ``check=False``), ``'geodesic'`` (radians, clamp of
    point_cloud/main.py:61-73), ``'geodesic_a6'`` (clamp +-1 of :209-227) or ``'frobenius'`` (loss_frobenius).
    Returns a 0-d tensor on the device (float64 for angle_error, float32 otherwise).  Not differentiable: it is the
    evaluation path (the reference runs it under ``torch.no_grad()``)."""
    if rot_rep not in _MAP_KEYS:
        raise KeyError(rot_rep)
    map_id, width = _ops.MAP_IDS[_MAP_KEYS[rot_rep]]
    lo, hi = -1.0, 1.0
    if metric == "geodesic":
        lo, hi = -1.0 + 1e-7, 1.0 - 1e-7
    mkey = "geodesic" if metric in ("geodesic", "geodesic_a6") else metric
    if mkey not in _ops.METRIC_IDS:
        raise ValueError(f"unknown metric {metric!r}")
    if reduction not in ("mean", "sum"):
        raise ValueError("reduction must be 'mean' or 'sum'")
    rec = x.detach().reshape(-1, width)
    s, flag = _ops.map_metric_sum(rec, R_true.detach(), map_id, width, _ops.METRIC_IDS[mkey], lo, hi)
    if mkey == "angle_error" and check and int(flag.item()) & 1:
        raise ValueError("angle out of range, input probably not proper rotation matrices")
    out = s / rec.shape[0] if reduction == "mean" else s
    return out.reshape(()) if mkey == "angle_error" else out.reshape(()).to(torch.float32)


def loss_frobenius(R_pred, R_true):
    """mean_b |R_true - R_pred|_F (un-squared), 0-d float32 (Comparison/main.py:12-16 and its 7 copies).

    When one argument is the untouched output of :func:`symmetric_orthogonalization` and a gradient will be asked for
    (the reference's training step, Comparison/main.py:59-62, UPNA/main.py:54-57, 3D-Pose/main.py:60,85-91), the loss
    and d loss/d x come from ONE kernel on that call's own input (``ops.frobenius_of_svd_output``): the unchanged two-call
    surface then costs 72 + 108 B/matrix instead of 360.  Same numbers as :func:`svd_frobenius_loss`."""
    if not torch.compiler.is_compiling():          # (a traced graph goes through the registered ops)
        if _ops._C is not None and type(R_pred) is torch.Tensor and type(R_true) is torch.Tensor:
            fused = _ops._C.frobenius_of_svd_output(R_pred, R_true)     # outputs of the C++ svd_orth node
            if fused is not None:
                return fused
        fused = _ops.frobenius_of_svd_output(R_pred, R_true)            # outputs of the Python svd_orth node
        if fused is not None:
            return fused
        if _pair_for_cpp(R_pred, R_true):
            return _ops._C.frobenius_mean(R_pred, R_true)               # generic pair, C++ node
    return _ops.frobenius_mean(R_pred, R_true)


def svd_frobenius_loss(x, R_true, single_pass=True, with_angle=False, check=True):
    """``loss_frobenius(symmetric_orthogonalization(x), R_true)`` fused (Comparison/main.py:59-62 and the 7 copies).

    When ``x`` needs a gradient the default is the single-pass step: ONE decomposition per matrix yields the loss and
    d loss/d x (108 B/matrix), and ``backward`` only scales that gradient by the upstream scalar (72 B/matrix, nothing
    when it is exactly 1).  ``single_pass=False`` keeps nothing between the passes (72 B forward, 108 B backward that
    recomputes the SVD).

    ``with_angle=True`` returns ``(loss, mean_angle)``: ``angle_error(symmetric_orthogonalization(x), R_true).mean()`` in
    degrees (0-d float64) from the SAME kernel - the metric every training loop of the reference computes next to the
    loss (Comparison/main.py:59-70, 3D-Pose/main.py:85-91).  ``check`` reads the range flag back and raises the
    reference's ``ValueError`` (rotation_representation.py:237-239); ``check=False`` synchronises nothing."""
    rec = x if (x.dim() == 2 and x.shape[1] == 9) else x.reshape(-1, 9)
    if with_angle:
        loss, angle, flag = _ops.fused_step_loss(rec, R_true, True)
        if check and int(flag.item()) & 1:
            raise ValueError("angle out of range, input probably not proper rotation matrices")
        return loss, angle
    if single_pass and torch.is_grad_enabled() and rec.requires_grad:
        return _ops.fused_step_loss(rec, R_true, False)
    return _ops.svd_orth_frobenius(rec, R_true)


def svd_frobenius_loss_and_grad(x, R_true):
    """(loss, d loss / d x) of the fused step in one kernel, without autograd: for training loops that call
    ``x.backward(grad)`` themselves.  ``grad`` has the shape of ``x``."""
    rec = x if (x.dim() == 2 and x.shape[1] == 9) else x.reshape(-1, 9)
    loss, dx = _ops.svd_orth_frobenius_step(rec.detach(), R_true)
    return loss, dx.reshape(x.shape)


def kabsch(P, Q, weights=None):
    """Rotation and translation aligning cloud P onto Q: (B, N, 3) x2 -> ((B, 3, 3), (B, 3)), R p + t ~ q.

    Not in the reference; the cloud pair is the (pc1, pc_out) of point_cloud/main.py:171-181.
    Differentiable w.r.t. both clouds (poseestimation_b200.point_cloud.kabsch is the same function); with
    ``weights`` (B, N) >= 0 it solves the weighted problem and is differentiable w.r.t. the weights as well."""
    if weights is not None:
        r, t, _ = _ops.kabsch_weighted_ctx(P, Q, weights)
        return r, t
    if torch.is_grad_enabled() and (P.requires_grad or Q.requires_grad):
        r, t, _ = _ops.kabsch_ctx(P, Q)
        return r, t
    return _ops.kabsch(P, Q)


# ---------------------------------------------------------------------------------------------------------------------
transform_output = {'SVD': (9, symmetric_orthogonalization), '6D': (6, compute_rotation_matrix_from_ortho6d),
                    '3D': (3, vec_3d_to_SO3)}     # rotation_representation.py:323-324


def get_scene_parameters():
    """(fx, fy, vx, vy) of the 320 px / 36 mm / 50 mm camera (rotation_representation.py:327-343)."""
    sw, img_res, flen = 36, 320, 50
    ppm = sw / img_res
    fx = fy = flen / ppm
    vx = vy = img_res / 2
    return fx, fy, vx, vy


def parametrization(output, T_init, fx, fy, vx, vy, rot_rep):
    """CosyPose pose update (rotation_representation.py:346-373; the reference body uses undefined names and cannot
    run - this follows its formulas and the working twin Iterative/utility.py:90-128).  ``'SVD'`` is ONE native kernel
    (orthogonalisation, dR R_k, translation update, 4x4 assembly; differentiable w.r.t. ``output``); for the other keys
    the rotation goes through its native map and only the 12-number assembly is torch indexing."""
    vector_length, representation_func = transform_output[rot_rep]
    if rot_rep == 'SVD' and output.shape[1] >= 12:
        return _ops.pose_update(output[:, :12], T_init.float(), float(fx), float(fy))
    R = representation_func(output[:, 0:vector_length])
    v = output[:, vector_length:vector_length + 3]
    xk, yk, zk = T_init[:, 0, 3], T_init[:, 1, 3], T_init[:, 2, 3]
    zk1 = v[:, 2] * zk
    yk1 = (v[:, 1] / fy + yk / zk) * zk1
    xk1 = (v[:, 0] / fx + xk / zk) * zk1
    T_out = torch.zeros((output.shape[0], 4, 4), dtype=output.dtype, device=output.device)
    T_out[:, :3, :3] = R @ T_init[:, :3, :3]
    T_out[:, :3, 3] = torch.stack((xk1, yk1, zk1), dim=1)
    T_out[:, 3, 3] = 1
    return T_out


_DROP_IN = ("compute_rotation_matrix_from_ortho6d", "compute_rotation_matrix_from_euler",
            "compute_rotation_matrix_from_ortho5d", "compute_rotation_matrix_from_quaternion",
            "symmetric_orthogonalization", "compute_geodesic_distance_from_two_matrices", "angle_error",
            "vec_3d_to_SO3", "so3_exp_map", "loss_frobenius", "geodesic")


def install(namespace) -> list:
    """Rebind the reference's private copies of the hot-path functions in ``namespace`` (a module or its globals
    dict) to the native ones; returns the names that were rebound.  For scripts that shadow the import with their
    own definitions: 3D-Pose/main2.py:34-70, 3D-Pose/test.py:32-73, point_cloud/main.py:43-155, UPNA/main.py:27-31,
    Iterative/loss.py:4-7, Comparison/main.py:12-31."""
    g = namespace if isinstance(namespace, dict) else vars(namespace)
    done = []
    for name in _DROP_IN:
        if name in g:
            g[name] = globals()[name]
            done.append(name)
    # the refiner's and the point-cloud experiment's helpers (Iterative/main.py: `from utility import *`,
    # `from loss import *`; point_cloud/main.py: `from prepare import *`) live in two sibling modules
    import importlib
    pkg = __package__ or "poseestimation_b200"
    for mod, names in ((".iterative", ("calculate_T_pred", "transform_pts", "compute_ADD_L1_loss",
                                       "compute_disentangled_ADD_L1_loss")),
                       (".point_cloud", ("get_sampled_rotation_matrices_by_axisAngle",))):
        m = importlib.import_module(mod, pkg)
        for name in names:
            if name in g:
                g[name] = getattr(m, name)
                done.append(name)
    return done
