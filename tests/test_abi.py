"""The C-ABI library loads and exports every symbol include/pe_rotation.h declares; the ctypes table matches the
header.  No compute calls (no GPU needed) except the argument-validation paths, which return before any CUDA call."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pe_rotation.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    out = {}
    for m in re.finditer(r"\b(int|long long|const char\s*\*)\s+(pe_\w+)\s*\(([^)]*)\)\s*;", src):
        args = [a.strip() for a in m.group(3).split(",") if a.strip() and a.strip() != "void"]
        out[m.group(2)] = args
    return out


@pytest.fixture(scope="module")
def built():
    from poseestimation_b200 import build
    return build.build()


def test_header_declares_the_expected_surface():
    d = _declared()
    for name in ("pe_version", "pe_last_error", "pe_launch_count", "pe_svd_orth_fwd", "pe_svd_orth_bwd",
                 "pe_ortho6d_fwd", "pe_quat_fwd", "pe_euler_fwd", "pe_frobenius_fwd", "pe_geodesic_fwd",
                 "pe_angle_error_fwd", "pe_svd_orth_frobenius_fwd", "pe_svd_orth_frobenius_bwd", "pe_svd_orth_frobenius_step",
                 "pe_scale_by_scalar", "pe_kabsch_fwd"):
        assert name in d
    assert len(d) >= 34


def test_library_exports_every_declared_symbol(built):
    lib = ctypes.CDLL(built)
    for name in _declared():
        assert hasattr(lib, name), f"{name} declared in pe_rotation.h but not exported"


def test_ctypes_table_matches_header(built):
    from poseestimation_b200 import _lib
    d = _declared()
    compute = {k: v for k, v in d.items() if k not in ("pe_version", "pe_last_error", "pe_launch_count", "pe_init")}
    assert set(compute) == set(_lib.SIGNATURES)
    for name, args in compute.items():
        sig = _lib.SIGNATURES[name]
        assert len(sig) == len(args), name
        for c_decl, ct in zip(args, sig):
            if "*" in c_decl:
                assert ct is ctypes.c_void_p, (name, c_decl)
            elif c_decl.startswith("int64_t"):
                assert ct is ctypes.c_int64, (name, c_decl)
            elif c_decl.startswith("float"):
                assert ct is ctypes.c_float, (name, c_decl)
            elif c_decl.startswith("double"):
                assert ct is ctypes.c_double, (name, c_decl)
            elif c_decl.startswith("int"):
                assert ct is ctypes.c_int, (name, c_decl)
            else:
                raise AssertionError(f"unhandled C type {c_decl!r}")


def test_version_and_error_convention(built):
    from poseestimation_b200 import _lib
    assert _lib.version() == 200
    # negative count is rejected before any CUDA call
    rc = _lib.lib.pe_svd_orth_fwd(None, 9, None, None, -1, None)
    assert rc == -1
    assert b"negative" in _lib.lib.pe_last_error()
    with pytest.raises(RuntimeError, match="bad call -1"):
        _lib.check(rc, "pe_svd_orth_fwd")
    assert _lib.lib.pe_kabsch_fwd(None, None, 4, 0, None, None, None) == -1
    assert _lib.lib.pe_map_metric_fwd(9, 0, 1024, 6, 2048, -1.0, 1.0, 4096, None, 1, None) == -1
    assert b"unknown map" in _lib.lib.pe_last_error()
    assert _lib.lib.pe_map_metric_fwd(1, 7, 1024, 6, 2048, -1.0, 1.0, 4096, None, 1, None) == -1
    # misaligned arrays are an error code, not a device fault (checked before any CUDA call)
    assert _lib.lib.pe_frobenius_fwd(1024, 2048, 4096 + 4, 1, None) == -1
    assert b"8 bytes" in _lib.lib.pe_last_error()
    assert _lib.lib.pe_euler_fwd(1024 + 2, 3, 2048, 1, None) == -1
    assert _lib.lib.pe_angle_error_fwd(1024, 2048, 4096 + 4, None, 8192, 1, None) == -1
    assert _lib.lib.pe_add_l1(1024, 2048, 4096, 1, 8, 0, None, 8192 + 4, None, None) == -1


def test_no_cpu_fallback():
    """CPU tensors must not be computed by some other path: the dispatcher has no CPU kernel for these ops."""
    import torch
    import poseestimation_b200.rotation_representation as rr
    with pytest.raises(NotImplementedError):
        rr.symmetric_orthogonalization(torch.randn(4, 9))
    with pytest.raises(NotImplementedError):
        rr.loss_frobenius(torch.randn(4, 3, 3), torch.randn(4, 3, 3))
    with pytest.raises(NotImplementedError):
        rr.kabsch(torch.randn(2, 8, 3), torch.randn(2, 8, 3))


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "poseestimation_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "reference_port" not in text, f


def test_fake_kernels_give_shapes():
    """register_fake: shape/dtype inference without a device (what torch.compile traces)."""
    import torch
    from torch._subclasses.fake_tensor import FakeTensorMode
    from poseestimation_b200 import ops
    with FakeTensorMode():
        x = torch.empty(5, 9, device="cuda")
        assert ops.svd_orth(x).shape == (5, 3, 3)
        assert ops.ortho6d(torch.empty(5, 6, device="cuda")).shape == (5, 3, 3)
        assert ops.frobenius_mean(torch.empty(5, 3, 3, device="cuda"), torch.empty(5, 3, 3, device="cuda")).shape == ()
        deg, flag = ops.angle_error_deg(torch.empty(5, 3, 3, device="cuda"), torch.empty(5, 3, 3, device="cuda"))
        assert deg.dtype == torch.float64 and deg.shape == (5,) and flag.dtype == torch.int32
        r, t = ops.kabsch(torch.empty(2, 16, 3, device="cuda"), torch.empty(2, 16, 3, device="cuda"))
        assert r.shape == (2, 3, 3) and t.shape == (2, 3)


def test_no_cpu_fallback_for_the_8f_modules():
    import torch
    import poseestimation_b200.iterative as it
    import poseestimation_b200.point_cloud as pc
    T = torch.eye(4).repeat(2, 1, 1)
    with pytest.raises(NotImplementedError):
        it.calculate_T_pred(torch.randn(2, 12), T)
    with pytest.raises(NotImplementedError):
        it.compute_disentangled_ADD_L1_loss(T, T, torch.randn(2, 8, 3))
    with pytest.raises(NotImplementedError):
        it.transform_pts(T, torch.randn(2, 8, 3))
    with pytest.raises(NotImplementedError):
        pc.rotate_cloud(torch.eye(3).repeat(2, 1, 1), torch.randn(2, 8, 3))
    with pytest.raises(NotImplementedError):
        pc.rotation_matrices_from_axis_angle(torch.randn(2), torch.randn(2, 3))
    with pytest.raises(NotImplementedError):
        pc.kabsch(torch.randn(2, 8, 3).requires_grad_(True), torch.randn(2, 8, 3))


def test_fake_kernels_of_the_8f_ops():
    import torch
    from torch._subclasses.fake_tensor import FakeTensorMode
    from poseestimation_b200 import ops
    with FakeTensorMode():
        c = lambda *s: torch.empty(*s, device="cuda")
        assert ops.affine_points(c(2, 3, 3), c(2, 3), c(2, 16, 3), False).shape == (2, 16, 3)
        assert ops.affine_points(c(2, 3, 3), None, c(2, 16, 3), True).shape == (2, 16, 3)
        m, s = ops.cloud_outer_sum(c(2, 16, 3), c(2, 16, 3))
        assert m.shape == (2, 3, 3) and s.shape == (2, 3)
        r, t, ctx = ops.kabsch_ctx(c(2, 16, 3), c(2, 16, 3))
        assert r.shape == (2, 3, 3) and t.shape == (2, 3) and ctx.shape == (2, 15)
        r, t, ctx = ops.kabsch_weighted_ctx(c(2, 16, 3), c(2, 16, 3), c(2, 16))
        assert r.shape == (2, 3, 3) and t.shape == (2, 3) and ctx.shape == (2, 16)
        dp, dq, dw = ops.kabsch_weighted_bwd(c(2, 16, 3), c(2, 16, 3), c(2, 16), r, ctx, c(2, 3, 3), c(2, 3), False,
                                             True, True)
        assert dp.shape == (0,) and dq.shape == (2, 16, 3) and dw.shape == (2, 16)
        assert ops.axis_angle_matrix(c(7), c(7, 3)).shape == (7, 3, 3)
        assert ops.pose_update(c(4, 12), c(4, 4, 4), 444.0, 444.0).shape == (4, 4, 4)
        assert ops.pose_update_bwd(c(4, 12), c(4, 4, 4), c(4, 4, 4), 444.0, 444.0).shape == (4, 12)
        item, tot, g = ops.add_l1(c(4, 4, 4), c(4, 4, 4), c(4, 32, 3), True)
        assert item.shape == (4,) and tot.dtype == torch.float64 and g.shape == (4, 3, 4)
        sm, fl = ops.map_metric_sum(c(5, 6), c(5, 3, 3), 1, 6, 2, -1.0, 1.0)
        assert sm.shape == (1,) and sm.dtype == torch.float64 and fl.dtype == torch.int32
        loss, dx = ops.svd_orth_frobenius_step(c(5, 9), c(5, 3, 3))
        assert loss.shape == () and dx.shape == (5, 9)


def test_iterative_module_mirrors_the_reference_signatures():
    """Same positional parameters as Iterative/utility.py:90 and Iterative/loss.py:10,31,51."""
    import inspect
    import poseestimation_b200.iterative as it
    assert list(inspect.signature(it.calculate_T_pred).parameters) == ["model_output", "T_init", "device", "rot_repr"]
    assert list(inspect.signature(it.compute_ADD_L1_loss).parameters) == ["TCO_gt", "TCO_pred", "points", "use_batch_mean"]
    assert list(inspect.signature(it.compute_disentangled_ADD_L1_loss).parameters) == ["T_CO_pred", "T_CO_gt", "points"]
    assert list(inspect.signature(it.transform_pts).parameters) == ["T", "pts"]
    assert it.get_scene_parameters() == (50 / (36 / 320), 50 / (36 / 320))


def test_install_rebinds_private_copies_including_the_8f_helpers():
    import poseestimation_b200.rotation_representation as rr
    import poseestimation_b200.iterative as it
    import poseestimation_b200.point_cloud as pc
    ns = {"symmetric_orthogonalization": None, "calculate_T_pred": None, "compute_disentangled_ADD_L1_loss": None,
          "get_sampled_rotation_matrices_by_axisAngle": None, "unrelated": 7}
    done = rr.install(ns)
    assert set(done) == set(ns) - {"unrelated"} and ns["unrelated"] == 7
    assert ns["symmetric_orthogonalization"] is rr.symmetric_orthogonalization
    assert ns["calculate_T_pred"] is it.calculate_T_pred
    assert ns["compute_disentangled_ADD_L1_loss"] is it.compute_disentangled_ADD_L1_loss
    assert ns["get_sampled_rotation_matrices_by_axisAngle"] is pc.get_sampled_rotation_matrices_by_axisAngle
