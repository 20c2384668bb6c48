#!/usr/bin/env python
"""Per-call latency of the public API at the batch sizes the reference trains with (32-128 matrices per step:
Comparison/main.py:100, UPNA/main.py:149, 3D-Pose configs, Iterative/main.py:216), where a call is launch- and
host-bound, not HBM-bound.  Prints one JSON object: microseconds per call (wall clock over `reps` back-to-back calls,
synchronised at both ends) for

  * svd_frobenius_loss(x, R_true).backward()         the fused training step (2 launches), on the eager fast path
    (ops._Eager) and through the registered torch.library op (the route torch.compile traces)
  * symmetric_orthogonalization(x) -> loss_frobenius -> backward   the unfused drop-in route (4 launches)
  * the same with the whole step captured in a CUDA graph and replayed
  * stock torch on the same GPU (torch.svd + det + matmul + matrix_norm, the calls of rotation_representation.py:192-206
    and Comparison/main.py:12-16) for scale

Usage: python tools/latency.py [--reps 300] [--batches 32,64,128,4096]"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import poseestimation_b200.rotation_representation as rr  # noqa: E402
from poseestimation_b200 import ops  # noqa: E402


def stock_torch_step(x, rt):
    m = x.view(-1, 3, 3)
    u, s, v = torch.svd(m)
    vt = v.transpose(1, 2)
    det = torch.det(torch.matmul(u, vt)).view(-1, 1, 1)
    vt = torch.cat((vt[:, :2, :], vt[:, -1:, :] * det), 1)
    r = torch.matmul(u, vt)
    loss = torch.linalg.matrix_norm(rt - r, ord="fro").mean()
    loss.backward()
    return loss


def timed(fn, reps):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e6


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=300)
    ap.add_argument("--batches", default="32,64,128,4096")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    res = {}
    for b in [int(v) for v in args.batches.split(",")]:
        x = torch.randn(b, 9, device=dev, requires_grad=True)
        rt = rr.symmetric_orthogonalization(torch.randn(b, 9, device=dev)).detach()

        def fused():
            x.grad = None
            rr.svd_frobenius_loss(x, rt).backward()

        def unfused():
            x.grad = None
            rr.loss_frobenius(rr.symmetric_orthogonalization(x), rt).backward()

        def no_autograd():
            rr.svd_frobenius_loss_and_grad(x, rt)

        def dispatcher():
            x.grad = None
            ops.svd_orth_frobenius_step.op(x, rt)[0].backward()

        def stock():
            x.grad = None
            stock_torch_step(x, rt)

        row = {"fused_step_us": timed(fused, args.reps), "fused_step_via_dispatcher_us": timed(dispatcher, args.reps),
               "unfused_dropin_us": timed(unfused, args.reps),
               "loss_and_grad_no_autograd_us": timed(no_autograd, args.reps),
               "stock_torch_gpu_us": timed(stock, max(20, args.reps // 10))}
        # the whole fused step as one CUDA graph (static input buffers, as a training loop would hold them)
        xs = x.detach().clone()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                rr.svd_frobenius_loss_and_grad(xs, rt)
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            loss, grad = rr.svd_frobenius_loss_and_grad(xs, rt)
        ref_loss, ref_grad = rr.svd_frobenius_loss_and_grad(xs, rt)
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(grad, ref_grad) and abs(float(loss) - float(ref_loss)) <= 1e-6
        row["loss_and_grad_cuda_graph_us"] = timed(graph.replay, args.reps)

        def step_with_metric():
            x.grad = None
            loss, ang = rr.svd_frobenius_loss(x, rt, with_angle=True, check=False)
            loss.backward()

        row["fused_step_with_mean_angle_us"] = timed(step_with_metric, args.reps)

        def unfused_python_nodes():         # the unchanged surface through the Python nodes of ops.py
            keep, ops._C = ops._C, None
            try:
                x.grad = None
                rr.loss_frobenius(rr.symmetric_orthogonalization(x), rt).backward()
            finally:
                ops._C = keep

        row["unfused_dropin_python_nodes_us"] = timed(unfused_python_nodes, args.reps)

        def python_node():                  # the Python autograd.Function the C++ node (csrc/pe_torch.cpp) replaces
            x.grad = None
            ops._FusedStep.apply(x, rt, None).backward()

        row["fused_step_python_node_us"] = timed(python_node, args.reps)
        row["cpp_node"] = float(ops._C is not None)

        def floor():                        # PyTorch's own floor: one CUDA op with autograd
            x.grad = None
            (x * 2).sum().backward()

        row["torch_one_op_autograd_floor_us"] = timed(floor, args.reps)
        res[str(b)] = {k: round(v, 2) for k, v in row.items()}
    # Kabsch at the reference's cloud batch (point_cloud/main.py:257: 128 clouds x 1024 points) and smaller: one warp per
    # pair, so 128 pairs occupy 16 of 148 SMs - is a call launch-bound (then nothing is to gain) or stream-bound?
    kab = {}
    for b, npts in ((32, 1024), (128, 1024), (128, 4096), (1024, 1024)):
        p = torch.rand(b, npts, 3, device=dev) - 0.5
        q = torch.rand(b, npts, 3, device=dev) - 0.5
        w = torch.rand(b, npts, device=dev)
        pg, qg = p.clone().requires_grad_(True), q.clone().requires_grad_(True)

        def fwd_bwd():
            pg.grad = qg.grad = None
            r, t = rr.kabsch(pg, qg)
            (r.sum() + t.sum()).backward()

        graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            rr.kabsch(p, q)
        torch.cuda.current_stream().wait_stream(side)
        with torch.cuda.graph(graph):
            rr.kabsch(p, q)
        kab[f"{b}x{npts}"] = {"kabsch_fwd_us": round(timed(lambda: rr.kabsch(p, q), args.reps), 2),
                              "kabsch_fwd_cuda_graph_us": round(timed(graph.replay, args.reps), 2),
                              "kabsch_weighted_fwd_us": round(timed(lambda: rr.kabsch(p, q, weights=w), args.reps), 2),
                              "kabsch_fwd_bwd_us": round(timed(fwd_bwd, args.reps), 2)}
    # the other representations of Comparison/main.py at its batch size: map -> loss_frobenius -> backward
    maps = {}
    b = 64
    rt = rr.symmetric_orthogonalization(torch.randn(b, 9, device=dev)).detach()
    for key, width, fn in (("6D", 6, rr.compute_rotation_matrix_from_ortho6d), ("Quat", 4, rr.compute_rotation_matrix_from_quaternion),
                           ("Euler", 3, rr.compute_rotation_matrix_from_euler), ("5D", 5, rr.compute_rotation_matrix_from_ortho5d),
                           ("3D", 3, rr.so3_exp_map)):
        xm = torch.randn(b, width, device=dev, requires_grad=True)

        def step(fn=fn, xm=xm):
            xm.grad = None
            rr.loss_frobenius(fn(xm), rt).backward()

        def step_geo(fn=fn, xm=xm):
            xm.grad = None
            rr.geodesic(fn(xm), rt).backward()

        def fwd_only(fn=fn, xm=xm):
            with torch.no_grad():
                fn(xm)

        maps[key] = {"map_frobenius_step_us": round(timed(step, args.reps), 2),
                     "map_geodesic_step_us": round(timed(step_geo, args.reps), 2),
                     "map_forward_no_grad_us": round(timed(fwd_only, args.reps), 2)}
    print(json.dumps({"what": "microseconds per call, wall clock, 1 B200", "reps": args.reps, "batch": res,
                      "kabsch": kab, "maps_batch64": maps}))


if __name__ == "__main__":
    main()
