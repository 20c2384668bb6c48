"""time the SVD-family kernels only (A/B of library variants selected with PE_ROTATION_LIB); prints one JSON line"""
import json, os, sys, torch
sys.path.insert(0, '.')
from poseestimation_b200 import _lib
lib, check = _lib.lib, _lib.check
n = 1 << 24
dev = torch.device('cuda', 0); st = torch.cuda.current_stream().cuda_stream
g = torch.Generator(device=dev).manual_seed(0)
x = torch.randn(n, 9, device=dev, generator=g); up = torch.randn(n, 9, device=dev, generator=g)
rt = torch.randn(n, 9, device=dev, generator=g)
check(lib.pe_svd_orth_fwd(rt.data_ptr(), 9, rt.data_ptr(), None, n, st), '')
out = torch.empty(n, 9, device=dev); sums = torch.zeros(3, dtype=torch.float64, device=dev)
P = lambda t: t.data_ptr()
def t(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return round(a.elapsed_time(b) / reps, 4)
wide = torch.randn(n, 12, device=dev, generator=g)          # the out[:, :9] view of a (B, 12) network output
res = {"lib": os.path.basename(_lib.LIB_PATH)}
for rep in range(2):
    res[f"fwd_{rep}"] = t(lambda: lib.pe_svd_orth_fwd(P(x), 9, P(out), None, n, st))
    res[f"bwd_{rep}"] = t(lambda: lib.pe_svd_orth_bwd(P(x), 9, None, P(up), P(out), n, st))
    res[f"step_{rep}"] = t(lambda: lib.pe_svd_orth_frobenius_step(P(x), 9, P(rt), 1.0 / n, P(out), P(sums), n, st))
    res[f"stepm_{rep}"] = t(lambda: lib.pe_svd_orth_frobenius_step_metrics(P(x), 9, P(rt), 1.0 / n, P(out), P(sums), None, None, None, None, n, st))
    res[f"wfwd_{rep}"] = t(lambda: lib.pe_svd_orth_fwd(P(wide), 12, P(out), None, n, st))
    res[f"wbwd_{rep}"] = t(lambda: lib.pe_svd_orth_bwd(P(wide), 12, None, P(up), P(out), n, st))
    res[f"wstep_{rep}"] = t(lambda: lib.pe_svd_orth_frobenius_step(P(wide), 12, P(rt), 1.0 / n, P(out), P(sums), n, st))
    res[f"wstepm_{rep}"] = t(lambda: lib.pe_svd_orth_frobenius_step_metrics(P(wide), 12, P(rt), 1.0 / n, P(out), P(sums), None, None, None, None, n, st))
print(json.dumps(res))
