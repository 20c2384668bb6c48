#!/usr/bin/env python
"""Launch the fused map + metric kernels (BASELINE config 3) and the plain maps twice on 2^24 resident records (for
`ncu --set full`, see profiles/README).
Usage: ncu --set full --clock-control none --import-source on -k regex:"Metric|Ortho6dFwd|QuatFwd|EulerFwd" -o out \
       python tools/profile_maps.py"""
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
    args = ap.parse_args()
    n, dev = args.n, torch.device("cuda", 0)
    st = torch.cuda.current_stream().cuda_stream
    g = torch.Generator(device=dev).manual_seed(0)
    P = lambda t: t.data_ptr()
    rt = torch.empty(n, 9, device=dev)
    check(lib.pe_svd_orth_fwd(P(torch.randn(n, 9, device=dev, generator=g)), 9, P(rt), None, n, st), "")
    out = torch.empty(n, 9, device=dev)
    sums = torch.zeros(2, dtype=torch.float64, device=dev)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    # (map id, record width) as in ops.MAP_IDS; metric ids 0 frobenius, 1 geodesic, 2 angle_error (ops.METRIC_IDS)
    for name, mid, w in (("svd", 0, 9), ("ortho6d", 1, 6), ("quat", 2, 4), ("euler", 3, 3)):
        x = torch.randn(n, w, device=dev, generator=g)
        for metric in (2, 1, 0):
            for _ in range(2):
                check(lib.pe_map_metric_fwd(mid, metric, P(x), w, P(rt), -1.0, 1.0, P(sums), P(flag), n, st), name)
        if name != "svd":
            for _ in range(2):
                check(getattr(lib, f"pe_{name}_fwd")(P(x), w, P(out), n, st), name)
        del x
    torch.cuda.synchronize()


if __name__ == "__main__":
    main()
