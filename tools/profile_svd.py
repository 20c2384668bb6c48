#!/usr/bin/env python
"""Launch each SVD kernel of the path twice on 2^24 resident matrices (for `ncu --set full`, see profiles/README).
Usage: ncu --set full --clock-control none --import-source on -o out python tools/profile_svd.py [--which a,b,c]"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from poseestimation_b200 import _lib  # noqa: E402

lib, check = _lib.lib, _lib.check


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1 << 24)
    ap.add_argument("--which", default="fwd,ffwd,fbwd,step,stepm")
    args = ap.parse_args()
    n, dev = args.n, torch.device("cuda", 0)
    st = torch.cuda.current_stream().cuda_stream
    g = torch.Generator(device=dev).manual_seed(0)
    x = torch.randn(n, 9, device=dev, generator=g)
    rt = torch.empty(n, 9, device=dev)
    check(lib.pe_svd_orth_fwd(torch.randn(n, 9, device=dev, generator=g).data_ptr(), 9, rt.data_ptr(), None, n, st), "")
    out = torch.empty(n, 9, device=dev)
    sums = torch.zeros(2, dtype=torch.float64, device=dev)
    one = torch.ones(1, device=dev)
    P = lambda t: t.data_ptr()
    calls = {
        "fwd": lambda: lib.pe_svd_orth_fwd(P(x), 9, P(out), None, n, st),
        "ffwd": lambda: lib.pe_svd_orth_frobenius_fwd(P(x), 9, P(rt), None, P(sums), n, st),
        "fbwd": lambda: lib.pe_svd_orth_frobenius_bwd(P(x), 9, P(rt), P(one), 1.0 / n, P(out), n, st),
    }
    sums3 = torch.zeros(3, dtype=torch.float64, device=dev)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    calls["step"] = lambda: lib.pe_svd_orth_frobenius_step(P(x), 9, P(rt), 1.0 / n, P(out), P(sums), n, st)
    calls["stepm"] = lambda: lib.pe_svd_orth_frobenius_step_metrics(P(x), 9, P(rt), 1.0 / n, P(out), P(sums3), None, None,
                                                                    None, None, n, st)
    calls["svdangle"] = lambda: lib.pe_map_metric_fwd(0, 2, P(x), 9, P(rt), -1.0, 1.0, P(sums), P(flag), n, st)
    if "wide" in args.which:
        wide = torch.randn(n, 12, device=dev, generator=g)
        calls["widefwd"] = lambda: lib.pe_svd_orth_fwd(P(wide), 12, P(out), None, n, st)
        calls["widebwd"] = lambda: lib.pe_svd_orth_bwd(P(wide), 12, None, P(rt), P(out), n, st)
        calls["widestep0"] = lambda: lib.pe_svd_orth_frobenius_step(P(wide), 12, P(rt), 1.0 / n, P(out), P(sums), n, st)
        calls["widestep"] = lambda: lib.pe_svd_orth_frobenius_step_metrics(P(wide), 12, P(rt), 1.0 / n, P(out), P(sums3),
                                                                           None, None, None, None, n, st)
    if "quatangle" in args.which or "eulerangle" in args.which:
        x4, x3 = torch.randn(n, 4, device=dev, generator=g), torch.randn(n, 3, device=dev, generator=g)
        calls["quatangle"] = lambda: lib.pe_map_metric_fwd(2, 2, P(x4), 4, P(rt), -1.0, 1.0, P(sums), P(flag), n, st)
        calls["eulerangle"] = lambda: lib.pe_map_metric_fwd(3, 2, P(x3), 3, P(rt), -1.0, 1.0, P(sums), P(flag), n, st)
    for name in args.which.split(","):
        if name in calls:
            for _ in range(2):
                check(calls[name](), name)
    torch.cuda.synchronize()


if __name__ == "__main__":
    main()
