#!/usr/bin/env python
"""Time every kernel of libpe_rotation.so on resident synthetic data and print achieved algorithmic GB/s against the
measured HBM peak (MEASURED_PEAKS.json).  BASELINE configs 2 (SVD), 3 (maps + metrics, 2^24) and 4 (Kabsch).
Usage: python tools/kernel_table.py [--n 16777216] [--json out.json]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from poseestimation_b200 import _lib  # noqa: E402
import poseestimation_b200.rotation_representation as rr  # noqa: E402

lib, check = _lib.lib, _lib.check


def timeit(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1 << 24)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    n = args.n
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    dev = torch.device("cuda", 0)
    st = torch.cuda.current_stream().cuda_stream
    g = torch.Generator(device=dev).manual_seed(0)
    rnd = lambda *s: torch.randn(*s, device=dev, generator=g)
    x9, g9 = rnd(n, 9), rnd(n, 9)
    r1 = rr.symmetric_orthogonalization(rnd(n, 9)).reshape(n, 9)
    r2 = rr.symmetric_orthogonalization(rnd(n, 9)).reshape(n, 9)
    o9, o9b = torch.empty(n, 9, device=dev), torch.empty(n, 9, device=dev)
    sums = torch.zeros(2, dtype=torch.float64, device=dev)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    th = torch.empty(n, device=dev)
    deg = torch.empty(n, dtype=torch.float64, device=dev)
    one = torch.ones(1, device=dev)
    P = lambda t: t.data_ptr()
    rows = []

    def add(name, bytes_per, fn, units=n, unit_name="rec"):
        ms = timeit(fn)
        gbs = bytes_per * units / (ms * 1e-3) / 1e9
        rows.append({"kernel": name, "bytes_per_unit": bytes_per, "ms": round(ms, 4), "GBps": round(gbs, 1),
                     "frac_of_measured_peak": round(gbs / peak, 3), "G_units_per_s": round(units / ms / 1e6, 3),
                     "unit": unit_name})

    add("svd_orth_fwd", 72, lambda: check(lib.pe_svd_orth_fwd(P(x9), 9, P(o9), None, n, st), ""))
    add("svd_orth_bwd (saved R; moves 144 B)", 108, lambda: check(lib.pe_svd_orth_bwd(P(x9), 9, P(o9), P(g9), P(o9b), n, st), ""))
    add("svd_orth_bwd (recompute)", 108, lambda: check(lib.pe_svd_orth_bwd(P(x9), 9, None, P(g9), P(o9b), n, st), ""))
    add("svd_orth_frobenius_fwd", 72, lambda: check(lib.pe_svd_orth_frobenius_fwd(P(x9), 9, P(r1), None, P(sums), n, st), ""))
    add("svd_orth_frobenius_bwd", 108, lambda: check(lib.pe_svd_orth_frobenius_bwd(P(x9), 9, P(r1), P(one), 1.0 / n, P(o9b), n, st), ""))
    add("svd_orth_frobenius_step (loss + grad, one SVD)", 108, lambda: check(lib.pe_svd_orth_frobenius_step(P(x9), 9, P(r1), 1.0 / n, P(o9b), P(sums), n, st), ""))
    sums3 = torch.zeros(3, dtype=torch.float64, device=dev)
    add("svd_orth_frobenius_step_metrics (loss + mean angle + grad)", 108, lambda: check(lib.pe_svd_orth_frobenius_step_metrics(P(x9), 9, P(r1), 1.0 / n, P(o9b), P(sums3), None, None, None, None, n, st), ""))
    w12 = rnd(n, 12)
    add("svd_orth_fwd on out[:, :9] of (n, 12) (wide TMA stage)", 72, lambda: check(lib.pe_svd_orth_fwd(P(w12), 12, P(o9), None, n, st), ""))
    add("svd_orth_bwd (recompute) on the (n, 12) view", 108, lambda: check(lib.pe_svd_orth_bwd(P(w12), 12, None, P(g9), P(o9b), n, st), ""))
    add("svd_orth_frobenius_step on the (n, 12) view", 108, lambda: check(lib.pe_svd_orth_frobenius_step(P(w12), 12, P(r1), 1.0 / n, P(o9b), P(sums), n, st), ""))
    add("step_metrics on the (n, 12) view", 108, lambda: check(lib.pe_svd_orth_frobenius_step_metrics(P(w12), 12, P(r1), 1.0 / n, P(o9b), P(sums3), None, None, None, None, n, st), ""))
    del w12
    add("scale_by_scalar (backward of the step)", 72, lambda: check(lib.pe_scale_by_scalar(P(o9b), P(one), P(o9), 9 * n, st), ""))
    for name, w in (("ortho6d", 6), ("quat", 4), ("euler", 3), ("ortho5d", 5), ("expmap", 3)):
        xi, dxi = rnd(n, w), torch.empty(n, w, device=dev)
        f, b = getattr(lib, f"pe_{name}_fwd"), getattr(lib, f"pe_{name}_bwd")
        add(f"{name}_fwd", 4 * w + 36, lambda f=f, xi=xi: check(f(P(xi), xi.shape[1], P(o9), n, st), ""))
        add(f"{name}_bwd", 8 * w + 36, lambda b=b, xi=xi, dxi=dxi: check(b(P(xi), xi.shape[1], P(g9), P(dxi), n, st), ""))
    add("frobenius_fwd", 72, lambda: check(lib.pe_frobenius_fwd(P(r1), P(r2), P(sums), n, st), ""))
    add("frobenius_bwd (one grad)", 108, lambda: check(lib.pe_frobenius_bwd(P(r1), P(r2), P(one), 1.0 / n, P(o9), None, n, st), ""))
    add("geodesic_fwd (theta)", 76, lambda: check(lib.pe_geodesic_fwd(P(r1), P(r2), -1.0, 1.0, P(th), None, n, st), ""))
    add("geodesic_fwd (fused mean)", 72, lambda: check(lib.pe_geodesic_fwd(P(r1), P(r2), -1.0, 1.0, None, P(sums), n, st), ""))
    add("geodesic_bwd (one grad)", 112, lambda: check(lib.pe_geodesic_bwd(P(r1), P(r2), -1.0, 1.0, P(th), 0, 1.0, P(o9), None, n, st), ""))
    add("angle_error_fwd (deg, f64)", 80, lambda: check(lib.pe_angle_error_fwd(P(r1), P(r2), P(deg), None, P(flag), n, st), ""))
    add("angle_error_fwd (fused mean)", 72, lambda: check(lib.pe_angle_error_fwd(P(r1), P(r2), None, P(sums), P(flag), n, st), ""))
    # fused map + metric (config 3): nothing written
    for name, mid, w in (("svd", 0, 9), ("ortho6d", 1, 6), ("quat", 2, 4), ("euler", 3, 3)):
        xi = x9 if w == 9 else rnd(n, w)
        add(f"{name} + angle_error mean (fused, f64)", 4 * w + 36,
            lambda mid=mid, xi=xi, w=w: check(lib.pe_map_metric_fwd(mid, 2, P(xi), w, P(r1), -1.0, 1.0, P(sums), P(flag), n, st), ""))
    x6 = rnd(n, 6)
    add("ortho6d + geodesic mean (fused)", 60, lambda: check(lib.pe_map_metric_fwd(1, 1, P(x6), 6, P(r1), -1.0, 1.0, P(sums), None, n, st), ""))
    add("ortho6d + frobenius mean (fused)", 60, lambda: check(lib.pe_map_metric_fwd(1, 0, P(x6), 6, P(r1), -1.0, 1.0, P(sums), None, n, st), ""))
    del x6
    del x9, g9, r1, r2, o9, o9b
    b, npts = 65536, 1024
    pc = torch.rand(b, npts, 3, device=dev, generator=g) - 0.5
    qc = torch.rand(b, npts, 3, device=dev, generator=g) - 0.5
    rk, tk = torch.empty(b, 9, device=dev), torch.empty(b, 3, device=dev)
    add("kabsch_fwd (65536 x 1024)", 24 * npts + 48, lambda: check(lib.pe_kabsch_fwd(P(pc), P(qc), b, npts, P(rk), P(tk), st), ""),
        units=b, unit_name="pair")
    # SURVEY 8(f) rows at the same cloud shape
    ctx, cq, cp = torch.empty(b, 15, device=dev), torch.empty(b, 12, device=dev), torch.empty(b, 12, device=dev)
    gk = torch.randn(b, 12, device=dev, generator=g)
    oc = torch.empty(b, npts, 3, device=dev)
    add("kabsch_fwd + context", 24 * npts + 108, lambda: check(lib.pe_kabsch_fwd_ctx(P(pc), P(qc), b, npts, P(rk), P(tk), P(ctx), st), ""),
        units=b, unit_name="pair")
    add("kabsch_bwd_coeffs (per pair)", 4 * (36 + 24), lambda: check(lib.pe_kabsch_bwd_coeffs(P(ctx), P(rk), P(gk), npts, P(cq), P(cp), b, st), ""),
        units=b, unit_name="pair")
    add("affine_points (rotate cloud / dQ / dP)", 24, lambda: check(lib.pe_affine_points(P(cq), 12, 3, 0, P(cq) + 36, 12, 1, P(pc), b, npts, P(oc), st), ""),
        units=b * npts, unit_name="point")
    add("cloud_outer_sum (dA, dt of affine)", 24 * npts + 48, lambda: check(lib.pe_cloud_outer_sum(P(qc), P(pc), b, npts, P(rk), P(tk), st), ""),
        units=b, unit_name="pair")
    wk = torch.rand(b, npts, device=dev, generator=g)
    ctxw, dwk, oc2 = torch.empty(b, 16, device=dev), torch.empty(b, npts, device=dev), torch.empty(b, npts, 3, device=dev)
    add("kabsch_weighted_fwd + context", 28 * npts + 112, lambda: check(lib.pe_kabsch_weighted_fwd(P(pc), P(qc), P(wk), b, npts, P(rk), P(tk), P(ctxw), st), ""),
        units=b, unit_name="pair")
    check(lib.pe_kabsch_weighted_bwd_coeffs(P(ctxw), P(rk), P(gk), P(cq), P(cp), b, st), "")
    add("kabsch_weighted_bwd_points (dp, dq, dw)", 56, lambda: check(lib.pe_kabsch_weighted_bwd_points(P(pc), P(qc), P(wk), P(cq), P(cp), P(ctxw), b, npts, P(oc), P(oc2), P(dwk), st), ""),
        units=b * npts, unit_name="point")
    add("kabsch_weighted_bwd_points (dw only)", 32, lambda: check(lib.pe_kabsch_weighted_bwd_points(P(pc), P(qc), P(wk), P(cq), P(cp), P(ctxw), b, npts, None, None, P(dwk), st), ""),
        units=b * npts, unit_name="point")
    del wk, dwk, oc2
    tg = torch.eye(4, device=dev).repeat(b, 1, 1) + 0.1 * torch.randn(b, 4, 4, device=dev, generator=g)
    tp = torch.eye(4, device=dev).repeat(b, 1, 1) + 0.1 * torch.randn(b, 4, 4, device=dev, generator=g)
    item, gsd = torch.empty(b, device=dev), torch.empty(b, 12, device=dev)
    for dis in (0, 1):
        add(f"add_l1 loss + grad ({'disentangled' if dis else 'plain'})", 12 * npts + 128 + 52,
            lambda dis=dis: check(lib.pe_add_l1(P(tg), P(tp), P(pc), b, npts, dis, P(item), P(sums), P(gsd), st), ""),
            units=b, unit_name="item")
    mo, tpo, dmo = torch.randn(n // 8, 12, device=dev, generator=g), torch.empty(n // 8, 16, device=dev), torch.empty(n // 8, 12, device=dev)
    tin = (torch.eye(4, device=dev).repeat(n // 8, 1, 1) + 0.1 * torch.randn(n // 8, 4, 4, device=dev, generator=g)).contiguous()
    gto = torch.randn(n // 8, 16, device=dev, generator=g)
    add("pose_update_fwd", 4 * (12 + 16 + 16), lambda: check(lib.pe_pose_update_fwd(P(mo), 12, P(tin), 444.4, 444.4, P(tpo), n // 8, st), ""),
        units=n // 8)
    add("pose_update_bwd", 4 * (12 + 16 + 16 + 12), lambda: check(lib.pe_pose_update_bwd(P(mo), 12, P(tin), P(gto), 444.4, 444.4, P(dmo), n // 8, st), ""),
        units=n // 8)
    th1, ax3 = torch.randn(n, device=dev, generator=g), torch.randn(n, 3, device=dev, generator=g)
    o9c = torch.empty(n, 9, device=dev)
    add("axis_angle_fwd (sampler map)", 16 + 36, lambda: check(lib.pe_axis_angle_fwd(P(th1), P(ax3), P(o9c), n, st), ""))
    print(f"{'kernel':42s} {'B/unit':>7s} {'ms':>8s} {'GB/s':>8s} {'of peak':>8s}  (peak {peak} GB/s measured, n = {n})")
    for r in rows:
        print(f"{r['kernel']:42s} {r['bytes_per_unit']:7d} {r['ms']:8.4f} {r['GBps']:8.1f} {r['frac_of_measured_peak']:8.3f}")
    if args.json:
        json.dump({"n": n, "peak_gbs": peak, "rows": rows}, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()
