"""N > 1 host logic on CPU: world_size-2 gloo.  The kernels need a GPU; the sharding + reduction logic does not."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from poseestimation_b200 import distributed as D


def test_shard_bounds_cover_everything():
    for n in (0, 1, 7, 16, 1000003):
        for world in (1, 2, 3, 8):
            spans = [D.shard_bounds(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        D.shard_bounds(4, 2, 2)


def test_single_rank_identity():
    p = D.pack_partials(torch.tensor([6.0], dtype=torch.float64), torch.tensor([30.0], dtype=torch.float64), 3)
    loss, ang, n = D.global_means(p)
    assert loss.item() == 2.0 and ang.item() == 10.0 and n.item() == 3.0
    assert D.backward_scale(4) == 0.25
    p2 = D.pack_partials(torch.tensor([6.0], dtype=torch.float64), None, 3)
    assert D.global_means_async(p2).wait()[0].item() == 2.0


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(0)
        per_record_loss = torch.rand(n, generator=g, dtype=torch.float64)      # same on every rank
        per_record_angle = torch.rand(n, generator=g, dtype=torch.float64) * 180
        lo, hi = D.shard_bounds(n, rank, world)
        part = D.pack_partials(per_record_loss[lo:hi].sum().reshape(1), per_record_angle[lo:hi].sum().reshape(1), hi - lo)
        pending = D.global_means_async(part.clone())          # the overlapped form gives the same numbers
        loss, ang, cnt = D.global_means(part)
        loss2, ang2, cnt2 = pending.wait()
        ok = (loss2.item() == loss.item() and ang2.item() == ang.item() and cnt2.item() == cnt.item() and
              abs(loss.item() - per_record_loss.mean().item()) < 1e-12 and
              abs(ang.item() - per_record_angle.mean().item()) < 1e-9 and int(cnt.item()) == n)
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [10, 1001])
def test_two_rank_gloo_all_reduce_gives_global_means(n):
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), n, out), nprocs=world, join=True)
    assert dict(out) == {0: True, 1: True}
