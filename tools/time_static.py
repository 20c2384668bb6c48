import sys, torch
sys.path.insert(0, '.')
from poseestimation_b200 import _lib
lib = _lib.lib
n = 1 << 23
a = torch.randn(n, 9, device='cuda'); b = torch.randn(n, 9, device='cuda')
s = torch.zeros(1, dtype=torch.float64, device='cuda'); st = torch.cuda.current_stream().cuda_stream
def t(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
ms0 = t(lambda: lib.pe_frobenius_fwd(a.data_ptr(), b.data_ptr(), s.data_ptr(), n, st))
ms1 = t(lambda: lib.pe_frobenius_fwd_static(a.data_ptr(), b.data_ptr(), s.data_ptr(), n, st))
print('n=2^23: pe_frobenius_fwd %.4f ms (%.0f GB/s), pe_frobenius_fwd_static %.4f ms (%.0f GB/s)' % (ms0, 72*n/ms0/1e6, ms1, 72*n/ms1/1e6))
