#!/usr/bin/env python
"""Copy-only ceiling of bench.py's end-to-end leg: the same 2 x 604 MB pinned cudaMemcpyAsync per rank and step with NO
kernels, at N = 1 / 2 / 4 / 8 ranks (one `cudaMemcpyAsync` per array, as the bench issues them - not batched).
Prints one JSON line (rank 0): GB/s per GPU (max-over-ranks time) for
  * default pinned host memory (torch `pin_memory`, what bench.py uses),
  * cudaHostAllocPortable | cudaHostAllocWriteCombined buffers,
  * both with the allocating thread bound to the GPU's NVML-local CPUs (when NVML reports a narrower set).
If the copy-only figure equals the e2e leg's H2D rate, the host / PCIe path is the limiter of e2e, not the pipeline.

    python tools/h2d_ceiling.py                      # N = 1
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 tools/h2d_ceiling.py
"""
import ctypes
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import bind_near_gpu  # noqa: E402

N = 1 << 24
BYTES = N * 36


def host_alloc(nbytes, flags):
    """cudaHostAlloc with explicit flags -> (uint8 tensor over the buffer, pointer for cudaFreeHost)."""
    rt = ctypes.CDLL("libcudart.so.12")
    p = ctypes.c_void_p()
    rc = rt.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(nbytes), ctypes.c_uint(flags))
    if rc != 0:
        raise RuntimeError(f"cudaHostAlloc -> {rc}")
    buf = (ctypes.c_uint8 * nbytes).from_address(p.value)
    return torch.frombuffer(buf, dtype=torch.uint8), (rt, p)


def measure(dev, hosts, reps=6):
    dsts = [torch.empty(BYTES, dtype=torch.uint8, device=dev) for _ in hosts]
    for d, h in zip(dsts, hosts):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    if dist.is_initialized():
        dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        for d, h in zip(dsts, hosts):
            d.copy_(h, non_blocking=True)
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / reps], dtype=torch.float64, device=dev)
    if dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    return {"ms_per_step": ms, "gbs_per_gpu": len(hosts) * BYTES / (ms * 1e-3) / 1e9}


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    out = {"what": "H2D copy-only ceiling of the e2e leg: 2 x %d bytes per rank per step, max over ranks" % BYTES,
           "n_gpus": world}
    h = [torch.empty(BYTES, dtype=torch.uint8).pin_memory() for _ in range(2)]
    for t in h:
        t.fill_(1)
    out["pinned_default"] = measure(dev, h)
    del h
    prev, note = bind_near_gpu(dev)
    out["numa_binding"] = note
    if prev is not None:
        h = [torch.empty(BYTES, dtype=torch.uint8).pin_memory() for _ in range(2)]
        for t in h:
            t.fill_(1)
        out["pinned_default_numa_local"] = measure(dev, h)
        del h
    try:
        keep = [host_alloc(BYTES, 0x01 | 0x04) for _ in range(2)]          # Portable | WriteCombined
        for t, _ in keep:
            t.fill_(1)
        out["portable_write_combined"] = measure(dev, [t for t, _ in keep])
        for t, (rt, p) in keep:
            del t
            rt.cudaFreeHost(p)
    except Exception as e:  # noqa: BLE001
        out["portable_write_combined"] = f"unavailable ({e})"
    if prev is not None:
        os.sched_setaffinity(0, prev)
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
