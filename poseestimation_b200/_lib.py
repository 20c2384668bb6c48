"""ctypes binding of ``libpe_rotation.so`` (C ABI: ``include/pe_rotation.h``).

There is no fallback: if the shared library is missing, or an entry point is missing from it, importing this module
raises.  Build it with ``python poseestimation_b200/build.py`` (or ``__graft_entry__.build()``).
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PE_ROTATION_LIB", os.path.join(_HERE, "libpe_rotation.so"))

_f = ctypes.c_void_p      # device pointers travel as integers
_i64 = ctypes.c_int64
_flt = ctypes.c_float
_int = ctypes.c_int
_dbl = ctypes.c_double

# name -> argtypes; every compute entry point returns int (0 = ok).  Mirrors include/pe_rotation.h one to one;
# tests/test_abi.py checks the two against each other.
SIGNATURES = {
    "pe_svd_orth_fwd": [_f, _i64, _f, _f, _i64, _f],
    "pe_svd_orth_bwd": [_f, _i64, _f, _f, _f, _i64, _f],
    "pe_ortho6d_fwd": [_f, _i64, _f, _i64, _f],
    "pe_ortho6d_bwd": [_f, _i64, _f, _f, _i64, _f],
    "pe_quat_fwd": [_f, _i64, _f, _i64, _f],
    "pe_quat_bwd": [_f, _i64, _f, _f, _i64, _f],
    "pe_euler_fwd": [_f, _i64, _f, _i64, _f],
    "pe_euler_bwd": [_f, _i64, _f, _f, _i64, _f],
    "pe_ortho5d_fwd": [_f, _i64, _f, _i64, _f],
    "pe_ortho5d_bwd": [_f, _i64, _f, _f, _i64, _f],
    "pe_expmap_fwd": [_f, _i64, _f, _i64, _f],
    "pe_expmap_bwd": [_f, _i64, _f, _f, _i64, _f],
    "pe_expmap_eps_fwd": [_f, _i64, _flt, _f, _i64, _f],
    "pe_expmap_eps_bwd": [_f, _i64, _f, _flt, _f, _i64, _f],
    "pe_frobenius_fwd": [_f, _f, _f, _i64, _f],
    "pe_frobenius_mean_fwd": [_f, _f, _f, _f, _i64, _f],
    "pe_frobenius_fwd_static": [_f, _f, _f, _i64, _f],
    "pe_frobenius_bwd": [_f, _f, _f, _flt, _f, _f, _i64, _f],
    "pe_geodesic_fwd": [_f, _f, _flt, _flt, _f, _f, _i64, _f],
    "pe_geodesic_reduce_fwd": [_f, _f, _flt, _flt, _f, _f, _dbl, _i64, _f],
    "pe_geodesic_bwd": [_f, _f, _flt, _flt, _f, _int, _flt, _f, _f, _i64, _f],
    "pe_angle_error_fwd": [_f, _f, _f, _f, _f, _i64, _f],
    "pe_svd_orth_frobenius_fwd": [_f, _i64, _f, _f, _f, _i64, _f],
    "pe_svd_orth_frobenius_bwd": [_f, _i64, _f, _f, _flt, _f, _i64, _f],
    "pe_svd_orth_frobenius_step": [_f, _i64, _f, _flt, _f, _f, _i64, _f],
    "pe_svd_orth_frobenius_step_mean": [_f, _i64, _f, _flt, _f, _f, _f, _i64, _f],
    "pe_svd_orth_frobenius_step_metrics": [_f, _i64, _f, _flt, _f, _f, _f, _f, _f, _f, _i64, _f],
    "pe_scale_by_scalar": [_f, _f, _f, _i64, _f],
    "pe_scale_by_scalar_inplace": [_f, _f, _i64, _f],
    "pe_map_metric_fwd": [_int, _int, _f, _i64, _f, _flt, _flt, _f, _f, _i64, _f],
    "pe_kabsch_fwd": [_f, _f, _i64, _i64, _f, _f, _f],
    "pe_kabsch_fwd_ctx": [_f, _f, _i64, _i64, _f, _f, _f, _f],
    "pe_kabsch_bwd_coeffs": [_f, _f, _f, _i64, _f, _f, _i64, _f],
    "pe_kabsch_weighted_fwd": [_f, _f, _f, _i64, _i64, _f, _f, _f, _f],
    "pe_kabsch_weighted_bwd_coeffs": [_f, _f, _f, _f, _f, _i64, _f],
    "pe_kabsch_weighted_bwd_points": [_f, _f, _f, _f, _f, _f, _i64, _i64, _f, _f, _f, _f],
    "pe_cloud_outer_sum": [_f, _f, _i64, _i64, _f, _f, _f],
    "pe_affine_points": [_f, _i64, _int, _int, _f, _i64, _int, _f, _i64, _i64, _f, _f],
    "pe_axis_angle_fwd": [_f, _f, _f, _i64, _f],
    "pe_pose_update_fwd": [_f, _i64, _f, _flt, _flt, _f, _i64, _f],
    "pe_pose_update_bwd": [_f, _i64, _f, _f, _flt, _flt, _f, _i64, _f],
    "pe_add_l1": [_f, _f, _f, _i64, _i64, _int, _f, _f, _f, _f],
}


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: the CUDA extension has not been built (python poseestimation_b200/build.py). "
            "There is no CPU or PyTorch fallback for the rotation-mapping path.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is missing: fail loudly
        fn.argtypes = argtypes
        fn.restype = ctypes.c_int
    lib.pe_version.argtypes = []
    lib.pe_version.restype = ctypes.c_int
    lib.pe_last_error.argtypes = []
    lib.pe_last_error.restype = ctypes.c_char_p
    lib.pe_launch_count.argtypes = []
    lib.pe_launch_count.restype = ctypes.c_longlong
    lib.pe_init.argtypes = [ctypes.c_int]
    lib.pe_init.restype = ctypes.c_int
    return lib


lib = _load()


def check(rc: int, what: str) -> None:
    """Map the C return convention to an exception (0 ok, >0 cudaError_t, <0 PE_ERR_*)."""
    if rc == 0:
        return
    msg = lib.pe_last_error().decode("utf-8", "replace")
    kind = "CUDA error" if rc > 0 else "bad call"
    raise RuntimeError(f"{what}: {kind} {rc}: {msg}")


_initialised: set = set()


def init_device(index: int) -> None:
    """``pe_init(index)`` once per device: the library's 128 KB of launch slots are allocated here, explicitly, so
    that no kernel launch ever allocates (and a launch may be captured into a CUDA graph)."""
    if index not in _initialised:
        check(lib.pe_init(index), "pe_init")
        _initialised.add(index)


def launch_count() -> int:
    return int(lib.pe_launch_count())


def version() -> int:
    return int(lib.pe_version())
