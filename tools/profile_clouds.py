#!/usr/bin/env python
"""Launch each cloud / refiner kernel twice at 65 536 x 1024 resident points (for `ncu --set full`, see profiles/README).
Usage: ncu --set full --clock-control none --import-source on -o out python tools/profile_clouds.py [--which a,b]"""
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
    ap.add_argument("--b", type=int, default=65536)
    ap.add_argument("--npts", type=int, default=1024)
    ap.add_argument("--which", default="kabsch,kabsch_w,wbwd,affine,outer,addl1")
    args = ap.parse_args()
    b, npts, dev = args.b, args.npts, torch.device("cuda", 0)
    st = torch.cuda.current_stream().cuda_stream
    g = torch.Generator(device=dev).manual_seed(0)
    p = torch.rand(b, npts, 3, device=dev, generator=g) - 0.5
    q = torch.rand(b, npts, 3, device=dev, generator=g) - 0.5
    w = torch.rand(b, npts, device=dev, generator=g)
    o1, o2, ow = torch.empty_like(p), torch.empty_like(p), torch.empty_like(w)
    r, t = torch.empty(b, 9, device=dev), torch.empty(b, 3, device=dev)
    ctx = torch.empty(b, 16, device=dev)
    cq, cp = torch.randn(b, 12, device=dev, generator=g), torch.randn(b, 12, device=dev, generator=g)
    tg = torch.eye(4, device=dev).repeat(b, 1, 1) + 0.1 * torch.randn(b, 4, 4, device=dev, generator=g)
    tp = torch.eye(4, device=dev).repeat(b, 1, 1) + 0.1 * torch.randn(b, 4, 4, device=dev, generator=g)
    item, gsd = torch.empty(b, device=dev), torch.empty(b, 12, device=dev)
    sums = torch.zeros(2, dtype=torch.float64, device=dev)
    P = lambda x: x.data_ptr()
    calls = {
        "kabsch": lambda: lib.pe_kabsch_fwd(P(p), P(q), b, npts, P(r), P(t), st),
        "kabsch_w": lambda: lib.pe_kabsch_weighted_fwd(P(p), P(q), P(w), b, npts, P(r), P(t), P(ctx), st),
        "wbwd": lambda: lib.pe_kabsch_weighted_bwd_points(P(p), P(q), P(w), P(cq), P(cp), P(ctx), b, npts, P(o1), P(o2),
                                                          P(ow), st),
        "affine": lambda: lib.pe_affine_points(P(cq), 12, 3, 0, P(cq) + 36, 12, 1, P(p), b, npts, P(o1), st),
        "outer": lambda: lib.pe_cloud_outer_sum(P(q), P(p), b, npts, P(r), P(t), st),
        "addl1": lambda: lib.pe_add_l1(P(tg), P(tp), P(p), b, npts, 1, P(item), P(sums), P(gsd), st),
    }
    for name in args.which.split(","):
        if name in calls:
            for _ in range(2):
                check(calls[name](), name)
    torch.cuda.synchronize()


if __name__ == "__main__":
    main()
