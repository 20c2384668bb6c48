"""B200-native rotation-mapping path of henrikgruner/PoseEstimation.

``poseestimation_b200.rotation_representation`` mirrors the reference module of that name; ``poseestimation_b200.ops``
holds the ``torch.library`` custom ops; ``libpe_rotation.so`` (C ABI in ``include/pe_rotation.h``) holds the sm_100a
kernels.  Importing the package loads the shared library and fails if it has not been built - there is no fallback.
"""
from . import _lib, ops, rotation_representation, iterative, point_cloud  # noqa: F401
from ._lib import launch_count, version  # noqa: F401

__all__ = ["ops", "rotation_representation", "iterative", "point_cloud", "launch_count", "version"]
