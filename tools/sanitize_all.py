"""Drive every kernel of the library once at small, ragged sizes - the workload for compute-sanitizer where it is
available:

    compute-sanitizer --tool memcheck  --error-exitcode 3 python tools/sanitize_all.py
    compute-sanitizer --tool racecheck --error-exitcode 3 python tools/sanitize_all.py

(tools/run_sanitizers.sh runs all four tools and keeps the logs; profiles/r2_sanitizer_*.log are its output on the GPU
pool.  tests/test_gpu_guard.py - guard bands around every output of every entry point - runs with the GPU tests.)

Sizes cover: one record, a ragged tail (n % 256 != 0, n % 512 != 0), several tiles per CTA (so the shared-memory rings
wrap and the tile scheduler hands out more tiles than there are CTAs), unaligned / strided inputs (the direct kernels),
clouds whose size is not a multiple of the staging chunk.  Prints the number of launches it made.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import poseestimation_b200.iterative as it                      # noqa: E402
import poseestimation_b200.point_cloud as pc                    # noqa: E402
import poseestimation_b200.rotation_representation as rr        # noqa: E402
from poseestimation_b200 import ops                             # noqa: E402
from poseestimation_b200._lib import lib                        # noqa: E402


def main():
    torch.manual_seed(0)
    dev = torch.device("cuda", 0)
    big = int(os.environ.get("PE_SANITIZE_N", 200003))          # > 148 CTAs x 512 records: tiles are re-drawn
    start = lib.pe_launch_count()
    for n in (1, 257, 1000, big):
        x = torch.randn(n, 9, device=dev, requires_grad=True)
        rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device=dev))
        r = rr.symmetric_orthogonalization(x)
        (r * torch.randn_like(r)).sum().backward()
        for single in (True, False):
            x.grad = None
            rr.svd_frobenius_loss(x, rt, single_pass=single).backward()
        loss, g = rr.svd_frobenius_loss_and_grad(x.detach(), rt)
        # the step with its metric, the drop-in route that parks its gradient on the svd_orth node, a second traversal
        x.grad = None
        loss, ang = rr.svd_frobenius_loss(x, rt, with_angle=True)
        loss.backward()
        x.grad = None
        rloss = rr.loss_frobenius(rr.symmetric_orthogonalization(x), rt)
        rloss.backward(retain_graph=True)
        rloss.backward()
        # 12-word rows whose storage ends with the last matrix (wide TMA stage: last tile by plain loads)
        tight = torch.randn(n * 12 - 3, device=dev).as_strided((n, 9), (12, 1))
        rr.symmetric_orthogonalization(tight)
        rr.svd_frobenius_loss_and_grad(tight, rt)
        rr.map_metric("SVD", tight, rt)
        # strided rows (read in place) and an unaligned base (direct kernels)
        wide = torch.randn(n, 12, device=dev)
        rr.symmetric_orthogonalization(wide[:, :9])
        odd = torch.randn(n * 9 + 1, device=dev)[1:].reshape(n, 9)
        rr.symmetric_orthogonalization(odd)
        for fname, w in (("compute_rotation_matrix_from_ortho6d", 6), ("compute_rotation_matrix_from_quaternion", 4),
                         ("compute_rotation_matrix_from_euler", 3), ("compute_rotation_matrix_from_ortho5d", 5),
                         ("vec_3d_to_SO3", 3)):
            xi = torch.randn(n, w, device=dev, requires_grad=True)
            ri = getattr(rr, fname)(xi)
            (ri * torch.randn_like(ri)).sum().backward()
            oddi = torch.randn(n * w + 1, device=dev)[1:].reshape(n, w)
            getattr(rr, fname)(oddi)
        r1 = r.detach().clone().requires_grad_(True)
        r2 = rt.clone()
        rr.loss_frobenius(r1, r2).backward()
        rr.geodesic(r1, r2, reduction="mean").backward()
        rr.geodesic(r1, r2, reduction="none").sum().backward()
        rr.compute_geodesic_distance_from_two_matrices(r1, r2)
        rr.angle_error(r1.detach(), r2)
        rr.mean_angle_error(r1.detach(), r2)
        # refiner
        out = torch.randn(n, 14, device=dev, requires_grad=True)
        t_init = torch.eye(4, device=dev).repeat(n, 1, 1) + 0.1 * torch.randn(n, 4, 4, device=dev)
        t_init[:, 3] = torch.tensor([0.0, 0, 0, 1], device=dev)
        t_init[:, 2, 3] += 2.0
        t_pred = it.calculate_T_pred(out, t_init)
        (t_pred * torch.randn_like(t_pred)).sum().backward()
        th, ax = torch.randn(n, device=dev), torch.randn(n, 3, device=dev)
        pc.rotation_matrices_from_axis_angle(th, ax)
    for b, npts in ((1, 1), (3, 5), (7, 1024), (33, 1021), (150, 2048 + 4), (300, 256)):
        p = (torch.rand(b, npts, 3, device=dev) - 0.5).requires_grad_(True)
        rot, q = pc.make_cloud_pair(p.detach())
        q = (q + 0.01 * torch.randn_like(q)).requires_grad_(True)
        r, t = pc.kabsch(p, q)
        ((r * torch.randn_like(r)).sum() + t.sum()).backward()
        rr.kabsch(p.detach(), q.detach())
        w = torch.rand(b, npts, device=dev, requires_grad=True)
        r, t = pc.kabsch(p, q, weights=w)
        ((r * torch.randn_like(r)).sum() + t.sum()).backward()
        r, t = pc.kabsch(p.detach(), q.detach(), weights=w)
        (r.sum() + t.sum()).backward()
        a = torch.randn(b, 3, 3, device=dev, requires_grad=True)
        tt = torch.randn(b, 3, device=dev, requires_grad=True)
        ops.affine_points(a, tt, p, False).sum().backward()
        ops.affine_points(a, None, p, True).sum().backward()
        poses = torch.eye(4, device=dev).repeat(b, 1, 1) + 0.1 * torch.randn(b, 4, 4, device=dev)
        poses2 = (poses + 0.05 * torch.randn_like(poses)).requires_grad_(True)
        it.transform_pts(poses, p.detach())
        it.compute_ADD_L1_loss(poses, poses2, p.detach()).backward()
        it.compute_ADD_L1_loss(poses, poses2, p.detach(), use_batch_mean=False)
        it.compute_disentangled_ADD_L1_loss(poses2, poses, p.detach()).backward()
    # a captured step replayed next to eager launches on another stream (launch slots: per stream / per captured launch)
    n = 40000
    xa, xb = torch.randn(n, 9, device=dev), torch.randn(n, 9, device=dev)
    rt = rr.symmetric_orthogonalization(torch.randn(n, 9, device=dev))
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    with torch.cuda.stream(s1):
        rr.svd_frobenius_loss_and_grad(xa, rt)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=s1):
        rr.svd_frobenius_loss_and_grad(xa, rt)
    for _ in range(int(os.environ.get("PE_SANITIZE_REPLAYS", 20))):
        with torch.cuda.stream(s1):
            graph.replay()
        with torch.cuda.stream(s2):
            rr.svd_frobenius_loss_and_grad(xb, rt)
    torch.cuda.synchronize()
    print(f"sanitize_all: {lib.pe_launch_count() - start} launches, ok")


if __name__ == "__main__":
    main()
