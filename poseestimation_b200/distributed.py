"""Batch-sharded use of the path across GPUs: the only exchange is one all-reduce of a 3-element float64 vector.

Mirrors ``reduce_loss`` of the reference's DDP smoke script (/root/reference/3D-Pose/main_DDP.py:56-60: ``dist.reduce``
of the scalar loss, then ``/ world_size``), but all-reduces SUMS and COUNTS so that ragged shards give the exact global
mean.  Works with any ``torch.distributed`` backend: NCCL on the GPUs, gloo in the CPU test-suite (the reduction logic,
not the kernels, is what runs there).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_bounds(n: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous split of ``n`` records: rank r owns [lo, hi).  The first ``n % world`` ranks get one extra."""
    if world < 1 or not (0 <= rank < world) or n < 0:
        raise ValueError("bad shard request")
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def pack_partials(loss_sum: torch.Tensor, angle_sum: torch.Tensor | None, count: int) -> torch.Tensor:
    """[sum_b |.|_F, sum_b theta_b, count] as one float64 vector on the partials' device (what gets all-reduced)."""
    v = torch.zeros(3, dtype=torch.float64, device=loss_sum.device)
    v[0] = loss_sum.reshape(-1)[0].to(torch.float64)
    if angle_sum is not None:
        v[1] = angle_sum.reshape(-1)[0].to(torch.float64)
    v[2] = float(count)
    return v


def global_means(partials: torch.Tensor, group=None) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """All-reduce the packed partials (in place, on the current stream) -> (mean loss, mean angle, global count).
    Without an initialised process group this is the single-rank identity."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(partials, op=dist.ReduceOp.SUM, group=group)
    n = partials[2].clamp_min(1.0)
    return partials[0] / n, partials[1] / n, partials[2]


def backward_scale(global_count: int) -> float:
    """The factor 1/B_global the backward kernels take: known before the step, so the backward needs no collective."""
    return 1.0 / max(int(global_count), 1)
