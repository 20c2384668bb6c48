"""pe_frobenius_fwd_static: the allocation-free variant of pe_frobenius_fwd (scratch in a file-scope static __device__
array, csrc/pe_static.cu), called through the C ABI and checked against the existing entry point on two calls in a row.

Collected LAST on purpose (the file name sorts behind every other test file): a
failure here must not hide another test.  Passes alone in a fresh process (it needs neither pe_init nor any earlier launch)
and after the other tests (it shares nothing with the launch-slot pool they use)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_RECORDS = 8388608            # PE_STATIC_MAX_RECORDS of include/pe_rotation.h


@pytest.fixture(scope="module")
def abi():
    from poseestimation_b200 import _lib
    return _lib


def _rotations(n, gen):
    """random proper rotations from unit quaternions (plain torch: this test only exercises the two loss entry points)"""
    q = torch.randn(n, 4, device="cuda", generator=gen)
    q = q / q.norm(dim=1, keepdim=True)
    w, x, y, z = q.unbind(1)
    r = torch.stack((1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
                     2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
                     2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)), dim=1)
    return r.contiguous()


@pytest.mark.parametrize("n", [1, 255, 257, 70001, 1 << 20, MAX_RECORDS])
def test_static_scratch_entry_point_matches_pe_frobenius_fwd_twice_in_a_row(abi, n):
    lib = abi.lib
    gen = torch.Generator(device="cuda").manual_seed(n)
    a, b = _rotations(n, gen), _rotations(n, gen)
    st = torch.cuda.current_stream().cuda_stream
    want = torch.full((1,), -1.0, dtype=torch.float64, device="cuda")
    abi.check(lib.pe_frobenius_fwd(a.data_ptr(), b.data_ptr(), want.data_ptr(), n, st), "pe_frobenius_fwd")
    got = [torch.full((1,), -7.0, dtype=torch.float64, device="cuda") for _ in range(2)]
    for g in got:                                        # two calls in a row share the one static array
        rc = lib.pe_frobenius_fwd_static(a.data_ptr(), b.data_ptr(), g.data_ptr(), n, st)
        assert rc == 0, lib.pe_last_error()
    torch.cuda.synchronize()
    ref = want.item()
    assert ref > 0
    # the tolerance the sum comparisons of tests/test_gpu_guard.py use: float64 sums of the same float32 terms
    for g in got:
        assert abs(g.item() - ref) <= 1e-9 * n
    assert got[0].item() == got[1].item()                # fixed summation order: bitwise reproducible
    # against plain torch in float64 as well (the mean of loss_frobenius, Comparison/main.py:12-16, to 1e-6)
    direct = (a.double() - b.double()).view(n, 9).norm(dim=1).sum().item()
    assert abs(got[0].item() - direct) / n <= 1e-6
    # arrays that are only 4-byte aligned take the scalar-load path of the kernel: same sum, bit for bit
    if n <= 1 << 20:
        ao = torch.empty(n * 9 + 1, device="cuda")[1:].view(n, 9).copy_(a)
        bo = torch.empty(n * 9 + 1, device="cuda")[1:].view(n, 9).copy_(b)
        odd = torch.zeros(1, dtype=torch.float64, device="cuda")
        assert lib.pe_frobenius_fwd_static(ao.data_ptr(), bo.data_ptr(), odd.data_ptr(), n, st) == 0
        torch.cuda.synchronize()
        assert odd.item() == got[0].item()
    # a different batch right after: nothing of the previous call's partials leaks into it
    m = max(1, n // 3)
    small = torch.full((1,), -7.0, dtype=torch.float64, device="cuda")
    assert lib.pe_frobenius_fwd_static(a.data_ptr(), b.data_ptr(), small.data_ptr(), m, st) == 0
    abi.check(lib.pe_frobenius_fwd(a.data_ptr(), b.data_ptr(), want.data_ptr(), m, st), "pe_frobenius_fwd")
    torch.cuda.synchronize()
    assert abs(small.item() - want.item()) <= 1e-9 * m


def test_static_scratch_entry_point_limits_and_no_allocation(abi):
    lib = abi.lib
    st = torch.cuda.current_stream().cuda_stream
    n = 4096
    gen = torch.Generator(device="cuda").manual_seed(5)
    a, b = _rotations(n, gen), _rotations(n, gen)
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    # beyond the maximum: a clear error, before any CUDA call
    rc = lib.pe_frobenius_fwd_static(a.data_ptr(), b.data_ptr(), out.data_ptr(), MAX_RECORDS + 1, st)
    assert rc == -1 and b"PE_STATIC_MAX_RECORDS" in lib.pe_last_error()
    assert lib.pe_frobenius_fwd_static(a.data_ptr(), b.data_ptr(), out.data_ptr(), -1, st) == -1
    assert lib.pe_frobenius_fwd_static(a.data_ptr(), b.data_ptr(), None, n, st) == -1
    assert lib.pe_frobenius_fwd_static(a.data_ptr(), b.data_ptr(), out.data_ptr() + 4, n, st) == -1
    # an empty batch defines its sum
    out.fill_(3.0)
    assert lib.pe_frobenius_fwd_static(None, None, out.data_ptr(), 0, st) == 0
    torch.cuda.synchronize()
    assert out.item() == 0.0
    # no device memory is taken by a call: the free-memory reading of the driver does not move over 50 calls
    assert lib.pe_frobenius_fwd_static(a.data_ptr(), b.data_ptr(), out.data_ptr(), n, st) == 0   # symbol resolved here
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(50):
        assert lib.pe_frobenius_fwd_static(a.data_ptr(), b.data_ptr(), out.data_ptr(), n, st) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
    # and it can be captured into a CUDA graph and replayed (after the one call above, outside the capture)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    cap = torch.zeros(1, dtype=torch.float64, device="cuda")
    with torch.cuda.graph(graph, stream=s):
        rc = lib.pe_frobenius_fwd_static(a.data_ptr(), b.data_ptr(), cap.data_ptr(), n, s.cuda_stream)
    assert rc == 0
    for _ in range(3):
        cap.fill_(-1.0)
        graph.replay()
        torch.cuda.synchronize()
        assert cap.item() == out.item()
