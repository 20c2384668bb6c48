"""Drop-in for the rotation-mapping part of the point-cloud experiment, SURVEY.md 8(f) rows 3 and 4.

Mirrors /root/reference/point_cloud/prepare.py (``get_sampled_rotation_matrices_by_axisAngle``, ``normalize_vector``)
and the call sites of /root/reference/point_cloud/main.py:171-197 (cloud-pair construction, ``geodesic``,
``angle_error``, ``loss_frobenius``); adds ``rotate_cloud`` and a differentiable ``kabsch``.
"""
from __future__ import annotations

import math

import torch

from . import ops as _ops
from .rotation_representation import (angle_error, compute_geodesic_distance_from_two_matrices, geodesic,  # noqa: F401
                                      loss_frobenius, normalize_vector, symmetric_orthogonalization)

__all__ = ["get_sampled_rotation_matrices_by_axisAngle", "rotation_matrices_from_axis_angle", "rotate_cloud",
           "make_cloud_pair", "kabsch", "normalize_vector", "geodesic", "angle_error", "loss_frobenius",
           "compute_geodesic_distance_from_two_matrices", "symmetric_orthogonalization"]


def rotation_matrices_from_axis_angle(theta, axis):
    """Deterministic part of prepare.py:21-49: theta (B,), un-normalised axis (B, 3) -> (B, 3, 3)."""
    return _ops.axis_angle_matrix(theta, axis)


def get_sampled_rotation_matrices_by_axisAngle(batch, device=None, generator=None):
    """point_cloud/prepare.py:21-49: theta ~ U(-pi, pi), axis ~ N(0, I) normalised, matrix of the quaternion
    (cos theta, axis sin theta).  Random numbers are drawn ON the device (the reference draws theta with numpy on the
    host and ships it over); the map to matrices is one kernel."""
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    theta = (torch.rand(batch, device=device, generator=generator) * 2.0 - 1.0) * math.pi
    axis = torch.randn(batch, 3, device=device, generator=generator)
    return _ops.axis_angle_matrix(theta, axis)


def rotate_cloud(rmat, pc):
    """pc2 = rmat @ pc1 for (B, 3, 3) rotations and (B, N, 3) clouds (point_cloud/main.py:173-181), without expanding
    the rotations to (B*N, 3, 3).  24 B per point.  Differentiable in both arguments."""
    return _ops.affine_points(rmat, None, pc, False)


def make_cloud_pair(pc1, generator=None):
    """(gt_rmat, pc_out) of point_cloud/main.py:171-181 for a batch of clouds pc1 (B, N, 3)."""
    gt_rmat = get_sampled_rotation_matrices_by_axisAngle(pc1.shape[0], pc1.device, generator)
    return gt_rmat, rotate_cloud(gt_rmat, pc1)


def kabsch(P, Q, weights=None):
    """Closed-form alignment of cloud pairs: R (B, 3, 3), t (B, 3) with R p + t ~ q.  Differentiable w.r.t. P and Q
    (analytic pullback through the covariance and the orthogonalisation; 48 B per point pair in the backward).

    ``weights`` (B, N), >= 0: the weighted problem  min sum_k w_k |R p_k + t - q_k|^2  (weighted centroids and
    covariance), differentiable w.r.t. P, Q and the weights - the layer a correspondence-weighting network feeds."""
    if weights is None:
        r, t, _ = _ops.kabsch_ctx(P, Q)
    else:
        r, t, _ = _ops.kabsch_weighted_ctx(P, Q, weights)
    return r, t
