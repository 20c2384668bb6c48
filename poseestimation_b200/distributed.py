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


class PendingMeans:
    """Handle of an all-reduce that is still in flight (see :func:`global_means_async`)."""

    def __init__(self, partials: torch.Tensor, work):
        self.partials, self.work = partials, work

    def wait(self) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        """Make the current stream wait for the reduction (no host block on CUDA) -> (mean loss, mean angle, count)."""
        if self.work is not None:
            self.work.wait()
            self.work = None
        n = self.partials[2].clamp_min(1.0)
        return self.partials[0] / n, self.partials[1] / n, self.partials[2]


def global_means_async(partials: torch.Tensor, group=None) -> PendingMeans:
    """Start the all-reduce of the packed partials and return at once.  Nothing in a training step depends on the
    reduced loss (the backward scale 1/B_global is known beforehand), so the caller launches the next kernels first
    and calls ``.wait()`` only where the value is consumed (logging): the 24-byte exchange then overlaps the next
    step's kernel instead of sitting between two of them.  ``partials`` must stay untouched until ``.wait()``."""
    work = None
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        work = dist.all_reduce(partials, op=dist.ReduceOp.SUM, group=group, async_op=True)
    return PendingMeans(partials, work)


def backward_scale(global_count: int) -> float:
    """The factor 1/B_global the backward kernels take: known before the step, so the backward needs no collective."""
    return 1.0 / max(int(global_count), 1)
