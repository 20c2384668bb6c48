"""CPU restatement of the reference's rotation-mapping path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may
import this module.  Nothing under ``poseestimation_b200/`` imports it: the product path is the CUDA extension and
fails loudly without it.

What is restated
----------------
Every function on the hot path named by SURVEY.md section 8(a).  The reference is pure Python on PyTorch, so the
restatement is PyTorch too (CPU, any float dtype; float64 for parity, float32 for the timed CPU baseline) and calls
the same library primitives the reference calls (``torch.svd``, ``torch.det``, ``torch.acos``, autograd).  The
arithmetic therefore lives in the third-party dependency **torch** (reference pin ``requirements.txt:7``
``torch==1.6.0``; installed here and on the GPU box: 2.11.0) exactly as it does for the reference.

The only deliberate differences from the reference source are the ones needed to run without a GPU:
the hard-coded ``.cuda()`` calls (rotation_representation.py:84,129,218,219) become ``device=<input>.device``.

Parity pin
----------
The reference has no tests and no golden vectors.  This port is pinned instead against
outputs of the UNMODIFIED reference functions, executed in this container by ``tests/golden/make_golden.py`` and
committed as ``tests/golden/*.npz``; ``tests/test_oracle.py`` checks every function here against those files.
Kabsch has no reference implementation at all: its restatement ends in the reference's own
symmetric orthogonalisation and its known answer is the reference's data construction ``pc2 = gt_rmat @ pc1``
(point_cloud/main.py:173-181) - for that op parity is pinned by construction, not by reference output.

All ``file:line`` citations are relative to ``/root/reference``.
"""
from __future__ import annotations

import math

import numpy as np
import torch

# --------------------------------------------------------------------------------------------------------------
# a1  symmetric orthogonalisation                                   rotation_representation.py:192-206


def symmetric_orthogonalization(x: torch.Tensor) -> torch.Tensor:
    """R = U diag(1, 1, det(U V^T)) V^T for every 3x3 block of ``x`` (rotation_representation.py:192-206)."""
    m = x.reshape(-1, 3, 3)                                   # :199  (reference: .view)
    u, _s, v = torch.svd(m)                                   # :200
    vt = v.transpose(1, 2)                                    # :201
    d = torch.det(u @ vt).reshape(-1, 1, 1)                   # :202-203
    vt = torch.cat((vt[:, :2, :], vt[:, 2:, :] * d), dim=1)   # :204  last ROW of V^T times det
    return u @ vt                                             # :205


def svd_orth_det_sign(x: torch.Tensor) -> torch.Tensor:
    """The factor ``det`` of rotation_representation.py:202 (= +-1), one per matrix."""
    m = x.reshape(-1, 3, 3)
    u, _s, v = torch.svd(m)
    return torch.det(u @ v.transpose(1, 2))


def svd_orth_pullback(x: torch.Tensor, grad_r: torch.Tensor) -> torch.Tensor:
    """Closed-form dL/dM of :func:`symmetric_orthogonalization`, float64.

    B = D U^T G V ;  W_ij = (B_ij - B_ji) / (s'_i + s'_j) ;  dM = U D W V^T  with  s' = (s1, s2, d*s3).
    Independent of autograd: tests compare it with autograd through the reference AND with the CUDA kernel.
    """
    m = x.reshape(-1, 3, 3).double()
    g = grad_r.reshape(-1, 3, 3).double()
    u, s, v = torch.svd(m)
    d = torch.det(u @ v.transpose(1, 2))
    dvec = torch.stack((torch.ones_like(d), torch.ones_like(d), d), dim=1)            # (B, 3)
    sp = s * dvec
    b = dvec[:, :, None] * (u.transpose(1, 2) @ g @ v)
    denom = sp[:, :, None] + sp[:, None, :]
    w = (b - b.transpose(1, 2)) / denom
    w = w * (1.0 - torch.eye(3, dtype=w.dtype))                                        # W_ii = 0
    return u @ (dvec[:, :, None] * w) @ v.transpose(1, 2)


def svd_orth_conditioning(x: torch.Tensor) -> torch.Tensor:
    """min_{i<j}(s'_i + s'_j) / s1 : the quantity that bounds forward and backward error."""
    m = x.reshape(-1, 3, 3).double()
    u, s, v = torch.svd(m)
    d = torch.det(u @ v.transpose(1, 2))
    s3 = d * s[:, 2]
    gap = torch.minimum(torch.minimum(s[:, 0] + s[:, 1], s[:, 0] + s3), s[:, 1] + s3)
    return gap / s[:, 0].clamp_min(1e-300)


# --------------------------------------------------------------------------------------------------------------
# a2  6D Gram-Schmidt                                               rotation_representation.py:21-36 (=174-189)


def compute_rotation_matrix_from_ortho6d(poses: torch.Tensor) -> torch.Tensor:
    assert poses.shape[-1] == 6                                              # :28
    a, b = poses[..., 0:3], poses[..., 3:6]                                  # :29-30
    x = a / torch.norm(a, p=2, dim=-1, keepdim=True)                         # :31
    z = torch.cross(x, b, dim=-1)                                            # :32
    z = z / torch.norm(z, p=2, dim=-1, keepdim=True)                         # :33
    y = torch.cross(z, x, dim=-1)                                            # :34
    return torch.stack((x, y, z), dim=-1)                                    # :35  columns x, y, z


def normalize_vector(v: torch.Tensor) -> torch.Tensor:
    """rotation_representation.py:38-50: divide by max(|v|, 1e-8)."""
    mag = v.pow(2).sum(1).sqrt()                                             # :46
    mag = torch.max(mag, torch.tensor([1e-8], dtype=torch.float32, device=v.device))   # :47 FloatTensor([1e-8])
    return v / mag.reshape(-1, 1)                                            # :48-49


def cross_product(u: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
    """rotation_representation.py:53-69."""
    i = u[:, 1] * v[:, 2] - u[:, 2] * v[:, 1]
    j = u[:, 2] * v[:, 0] - u[:, 0] * v[:, 2]
    k = u[:, 0] * v[:, 1] - u[:, 1] * v[:, 0]
    return torch.stack((i, j, k), dim=1)


# --------------------------------------------------------------------------------------------------------------
# a3  quaternion (w, x, y, z)                                       rotation_representation.py:139-172


def compute_rotation_matrix_from_quaternion(quaternion: torch.Tensor) -> torch.Tensor:
    q = normalize_vector(quaternion)                                         # :148
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]                          # :150-153
    xx, yy, zz = x * x, y * y, z * z                                         # :156-158
    xy, xz, yz = x * y, x * z, y * z                                         # :159-161
    xw, yw, zw = x * w, y * w, z * w                                         # :162-164
    r0 = torch.stack((1 - 2 * yy - 2 * zz, 2 * xy - 2 * zw, 2 * xz + 2 * yw), dim=1)   # :166
    r1 = torch.stack((2 * xy + 2 * zw, 1 - 2 * xx - 2 * zz, 2 * yz - 2 * xw), dim=1)   # :167
    r2 = torch.stack((2 * xz - 2 * yw, 2 * yz + 2 * xw, 1 - 2 * xx - 2 * yy), dim=1)   # :168
    return torch.stack((r0, r1, r2), dim=1)                                  # :170


# --------------------------------------------------------------------------------------------------------------
# a4  Euler                                                         rotation_representation.py:92-115


def compute_rotation_matrix_from_euler(euler: torch.Tensor) -> torch.Tensor:
    c1, s1 = torch.cos(euler[:, 0]), torch.sin(euler[:, 0])                  # :101-102
    c2, s2 = torch.cos(euler[:, 2]), torch.sin(euler[:, 2])                  # :103-104  (index 2!)
    c3, s3 = torch.cos(euler[:, 1]), torch.sin(euler[:, 1])                  # :105-106  (index 1!)
    r0 = torch.stack((c2 * c3, -s2, c2 * s3), dim=1)                                          # :108
    r1 = torch.stack((c1 * s2 * c3 + s1 * s3, c1 * c2, c1 * s2 * s3 - s1 * c3), dim=1)        # :109
    r2 = torch.stack((s1 * s2 * c3 - c1 * s3, s1 * c2, s1 * s2 * s3 + c1 * c3), dim=1)        # :110
    return torch.stack((r0, r1, r2), dim=1)                                  # :112


# --------------------------------------------------------------------------------------------------------------
# 5D (section 8f "next" row)                                        rotation_representation.py:72-90, 120-136


def stereographic_unproject(a: torch.Tensor, axis=None) -> torch.Tensor:
    batch, dim = a.shape
    if axis is None:
        axis = dim                                                           # :81-82
    s2 = a.pow(2).sum(1)                                                     # :83
    unproj = 2 * a / (s2 + 1).reshape(batch, 1)                              # :85
    new = ((s2 - 1) / (s2 + 1)).reshape(batch, 1)                            # :88
    return torch.cat((unproj[:, :axis], new, unproj[:, axis:]), dim=1)       # :86-89 (zeros().cuda() + slices)


def compute_rotation_matrix_from_ortho5d(a: torch.Tensor) -> torch.Tensor:
    batch = a.shape[0]
    scale = torch.tensor([math.sqrt(2) + 1, math.sqrt(2) + 1, math.sqrt(2)],
                         dtype=torch.float32, device=a.device)               # :128-129 FloatTensor(...).cuda()
    u = stereographic_unproject(a[:, 2:5] * scale.reshape(1, 3), axis=0)     # :131
    norm = u[:, 1:].pow(2).sum(1).sqrt()                                     # :132
    u = u / norm.reshape(batch, 1)                                           # :133
    b = torch.cat((a[:, 0:2], u), dim=1)                                     # :134
    return compute_rotation_matrix_from_ortho6d(b)                           # :135


# --------------------------------------------------------------------------------------------------------------
# exponential map (section 8f "next" row)                           rotation_representation.py:245-320


def hat(v: torch.Tensor) -> torch.Tensor:
    n, dim = v.shape
    if dim != 3:
        raise ValueError("Input vectors have to be 3-dimensional.")           # :289-290
    x, y, z = v.unbind(1)
    zero = torch.zeros_like(x)
    return torch.stack((torch.stack((zero, -z, y), 1), torch.stack((z, zero, -x), 1),
                        torch.stack((-y, x, zero), 1)), 1)                   # :292-302


def so3_exp_map(log_rot: torch.Tensor, eps: float = 0.0001) -> torch.Tensor:
    _, dim = log_rot.shape
    if dim != 3:
        raise ValueError("Input tensor shape has to be Nx3.")                 # :255-256
    nrms = (log_rot * log_rot).sum(1)                                        # :258
    ang = torch.clamp(nrms, eps).sqrt()                                      # :260
    inv = 1.0 / ang                                                          # :261
    fac1 = inv * ang.sin()                                                   # :262
    fac2 = inv * inv * (1.0 - ang.cos())                                     # :263
    k = hat(log_rot)                                                         # :264
    kk = torch.bmm(k, k)                                                     # :265
    eye = torch.eye(3, dtype=log_rot.dtype, device=log_rot.device)[None]
    return fac1[:, None, None] * k + fac2[:, None, None] * kk + eye          # :267-272


def vec_3d_to_SO3(x: torch.Tensor) -> torch.Tensor:
    assert x.shape == (x.shape[0], 3)                                        # :311
    return so3_exp_map(x)                                                    # :313


# --------------------------------------------------------------------------------------------------------------
# a5-a8  losses and distances


def loss_frobenius(r_pred: torch.Tensor, r_true: torch.Tensor) -> torch.Tensor:
    """mean_b |R_true - R_pred|_F, un-squared (Comparison/main.py:12-16 and 7 copies)."""
    return torch.linalg.matrix_norm(r_true - r_pred, ord="fro").mean()


def compute_geodesic_distance_from_two_matrices(m1: torch.Tensor, m2: torch.Tensor) -> torch.Tensor:
    """radians, input dtype (rotation_representation.py:209-227)."""
    m = torch.bmm(m1, m2.transpose(1, 2))                                    # :215
    cos = (m[:, 0, 0] + m[:, 1, 1] + m[:, 2, 2] - 1) / 2                     # :217
    one = torch.ones(m1.shape[0], dtype=torch.float32, device=m1.device)     # :218-219 torch.ones(batch).cuda()
    cos = torch.min(cos, one)                                                # :218
    cos = torch.max(cos, -one)                                               # :219
    return torch.acos(cos)                                                   # :222


def angle_error(r1: torch.Tensor, r2: torch.Tensor) -> torch.Tensor:
    """degrees, float64, ValueError outside [-1.1, 1.1] (rotation_representation.py:230-242)."""
    off = torch.matmul(r1.transpose(1, 2).double(), r2.double())             # :232-233
    tr = off.reshape(-1, 9)[:, ::4].sum(dim=1)                               # :234-235
    cos = (tr - 1) / 2                                                       # :236
    if torch.any(cos < -1.1) or torch.any(cos > 1.1):                        # :237
        raise ValueError("angle out of range, input probably not proper rotation matrices")   # :238-239
    return torch.acos(torch.clamp(cos, -1, 1)) * (180 / np.pi)               # :240-242


def geodesic(r1: torch.Tensor, r2: torch.Tensor, reduction: str = "mean"):
    """point_cloud/main.py:61-73 (eps-clamped, optional reduction)."""
    eps = 1e-7
    tr = (r1 @ r2.permute(0, 2, 1)).diagonal(dim1=-2, dim2=-1).sum(-1)       # :63-65
    d = torch.acos(torch.clamp((tr - 1) / 2, -1 + eps, 1 - eps))             # :66-67
    if reduction == "none":
        return d
    if reduction == "mean":
        return d.mean()
    if reduction == "sum":
        return d.sum()
    return None                                                              # reference falls off the end


# --------------------------------------------------------------------------------------------------------------
# a9  Kabsch - NO reference implementation.  Restatement: centroids, cross-covariance, then the
# reference's own det-fixed SVD projection.  Site and known answer: point_cloud/main.py:171-181.


def kabsch(p: torch.Tensor, q: torch.Tensor):
    """Rotation R (B,3,3) and translation t (B,3) minimising sum_n |R p_n + t - q_n|^2, computed in float64."""
    p64, q64 = p.double(), q.double()
    pc, qc = p64.mean(dim=1, keepdim=True), q64.mean(dim=1, keepdim=True)
    c = (q64 - qc).transpose(1, 2) @ (p64 - pc)               # C = sum (q - qc)(p - pc)^T  = H^T
    r = symmetric_orthogonalization(c.reshape(-1, 9))
    t = qc.squeeze(1) - (r @ pc.transpose(1, 2)).squeeze(2)
    return r, t


def kabsch_weighted(p: torch.Tensor, q: torch.Tensor, w: torch.Tensor):
    """Weighted variant (no reference function): weighted centroids and covariance, then the
    reference's own projection.  Pairs whose weights sum to zero get the projection of the zero matrix."""
    p, q, w = p.double(), q.double(), w.double()
    wsum = w.sum(dim=1, keepdim=True)
    inv = torch.where(wsum > 0, 1.0 / wsum.clamp_min(torch.finfo(w.dtype).tiny), torch.zeros_like(wsum))
    pbar = (w.unsqueeze(2) * p).sum(dim=1, keepdim=True) * inv.unsqueeze(2)
    qbar = (w.unsqueeze(2) * q).sum(dim=1, keepdim=True) * inv.unsqueeze(2)
    c = torch.einsum("bk,bki,bkj->bij", w, q - qbar, p - pbar)
    r = symmetric_orthogonalization(c.reshape(-1, 9))
    t = qbar.squeeze(1) - torch.einsum("bij,bj->bi", r, pbar.squeeze(1))
    return r, t


def sample_rotations_axis_angle(batch: int, generator: torch.Generator | None = None) -> torch.Tensor:
    """The cloud-pair rotation sampler of point_cloud/prepare.py:21-49, seeded and on the CPU.

    Quirk kept: the reference uses (cos t, axis*sin t) as the quaternion, i.e. the rotation angle is 2t.
    """
    theta = (torch.rand(batch, generator=generator, dtype=torch.float64) * 2 - 1) * math.pi   # :23
    axis = torch.randn(batch, 3, generator=generator, dtype=torch.float64)                    # :25
    axis = axis / axis.norm(dim=1, keepdim=True).clamp_min(1e-8)                              # :26
    q = torch.cat((torch.cos(theta)[:, None], axis * torch.sin(theta)[:, None]), dim=1)       # :27-30
    return compute_rotation_matrix_from_quaternion(q)                                         # :33-47 same rows


def rotation_matrices_from_axis_angle(theta: torch.Tensor, axis: torch.Tensor) -> torch.Tensor:
    """Deterministic part of point_cloud/prepare.py:21-49 for given theta (B,) and un-normalised axis (B, 3)."""
    axis = axis / axis.norm(dim=1, keepdim=True).clamp_min(1e-8)                              # :26 (normalize_vector)
    q = torch.cat((torch.cos(theta)[:, None], axis * torch.sin(theta)[:, None]), dim=1)       # :27-30
    # :33-47 are the rows of the unit-quaternion matrix WITHOUT re-normalising q (|q| = 1 by construction)
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    xx, yy, zz, xy, xz, yz, xw, yw, zw = x * x, y * y, z * z, x * y, x * z, y * z, x * w, y * w, z * w
    row0 = torch.stack((1 - 2 * yy - 2 * zz, 2 * xy - 2 * zw, 2 * xz + 2 * yw), 1)
    row1 = torch.stack((2 * xy + 2 * zw, 1 - 2 * xx - 2 * zz, 2 * yz - 2 * xw), 1)
    row2 = torch.stack((2 * xz - 2 * yw, 2 * yz + 2 * xw, 1 - 2 * xx - 2 * yy), 1)
    return torch.stack((row0, row1, row2), 1)


def rotate_cloud(rmat: torch.Tensor, pc: torch.Tensor) -> torch.Tensor:
    """pc2 = gt_rmat @ pc1 per point (point_cloud/main.py:173-181, without the (B*N, 3, 3) expansion)."""
    return (rmat @ pc.transpose(1, 2)).transpose(1, 2)


# --------------------------------------------------------------------------------------------------------------
# 8f.2  iterative refiner: pose update and ADD-L1 losses           Iterative/utility.py:73-128, Iterative/loss.py:10-69


def get_scene_parameters_iterative():
    """Iterative/utility.py:73-88: (fx, fy)."""
    sw, img_res, flen = 36, 320, 50
    ppm = sw / img_res
    fx = fy = flen / ppm
    return fx, fy


def calculate_T_pred(model_output: torch.Tensor, T_init: torch.Tensor) -> torch.Tensor:
    """Iterative/utility.py:90-128.  ``combine`` (:63-71) reads undefined globals (``model_output``, ``R_new``) and
    cannot run; its evident intent - ones((B,4,4)), rotation block, translation column, T[:, 3, :3] = 0 - is
    restated inline."""
    dR = symmetric_orthogonalization(model_output[:, :9])                       # :107
    transl = model_output[:, 9:12]
    vx, vy, vz = transl[:, 0], transl[:, 1], transl[:, 2]                       # :110
    R_k = T_init[:, :3, :3].to(model_output.dtype)                              # :112, :124
    fx, fy = get_scene_parameters_iterative()
    z_k = T_init[:, 2, 3]
    z_new = vz * z_k                                                            # :118
    x_k, y_k = T_init[:, 0, 3], T_init[:, 1, 3]
    x_new = (vx / fx + x_k / z_k) * z_new                                       # :122
    y_new = (vy / fy + y_k / z_k) * z_new                                       # :123
    R_new = torch.einsum('bij,bjk->bik', dR, R_k)                               # :125
    top = torch.cat((R_new, torch.stack((x_new, y_new, z_new), 1)[:, :, None]), dim=2)
    last = torch.zeros(model_output.shape[0], 1, 4, dtype=model_output.dtype)
    last[:, 0, 3] = 1
    return torch.cat((top, last), dim=1)


def transform_pts(T: torch.Tensor, pts: torch.Tensor) -> torch.Tensor:
    """Iterative/loss.py:51-69: T of shape (B, 4, 4) -> (B, N, 3), or (B, K, 4, 4) -> (B, K, N, 3) (:59-61)."""
    bsz, n_pts = T.shape[0], pts.shape[1]
    assert pts.shape == (bsz, n_pts, 3)                                      # :58
    if T.dim() == 4:
        pts = pts.unsqueeze(1)                                               # :60
        assert T.shape[-2:] == (4, 4)                                        # :61
    elif T.dim() == 3:
        assert T.shape == (bsz, 4, 4)                                        # :63
    else:
        raise ValueError('Unsupported shape for T', T.shape)                 # :65
    pts = pts.unsqueeze(-1)                                                  # :66
    T = T.unsqueeze(-3)                                                      # :67
    return (T[..., :3, :3] @ pts + T[..., :3, [-1]]).squeeze(-1)             # :68-69


def compute_ADD_L1_loss(TCO_gt, TCO_pred, points, use_batch_mean=True):
    """Iterative/loss.py:10-28."""
    dists = (transform_pts(TCO_gt, points) - transform_pts(TCO_pred, points)).abs().mean(dim=(-1, -2))
    return dists.mean() if use_batch_mean else dists


def compute_disentangled_ADD_L1_loss(T_CO_pred, T_CO_gt, points):
    """Iterative/loss.py:31-48."""
    rot, dep, tra = T_CO_gt.clone(), T_CO_gt.clone(), T_CO_gt.clone()
    rot[:, :3, :3] = T_CO_pred[:, :3, :3]
    tra[:, :2, 3] = T_CO_pred[:, :2, 3]
    dep[:, 2, 3] = T_CO_pred[:, 2, 3]
    return (compute_ADD_L1_loss(T_CO_gt, rot, points) + compute_ADD_L1_loss(T_CO_gt, tra, points)
            + compute_ADD_L1_loss(T_CO_gt, dep, points))


TRANSFORM_OUTPUT = {"SVD": (9, symmetric_orthogonalization), "6D": (6, compute_rotation_matrix_from_ortho6d),
                    "3D": (3, vec_3d_to_SO3)}                                # rotation_representation.py:323-324
