#!/usr/bin/env python
"""Benchmark of the rotation-mapping hot path.  Contract: one JSON line on stdout (rank 0).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--n MATRICES_PER_GPU] [--impl native|reference]

Metric (BASELINE.json): 3x3 SVD orthogonalisations per second, forward + backward.
A *step* is one pass of the reference's training step over one resident batch (Comparison/main.py:59-70,
UPNA/main.py:54-58, 3D-Pose/main.py:60,85-91):
    outR = symmetric_orthogonalization(out); loss = loss_frobenius(outR, R); loss.backward(); angle_error(outR, R)
as the launches the public API makes for it through the C ABI: the single-pass step kernel with the metric fused in
(reads x, R_true; ONE decomposition per matrix; writes dL/dx for a unit upstream gradient and the 24-byte vector
[sum |.|_F, sum angle_error, range flag]: 108 B/matrix) and the backward's in-place scaling by the upstream scalar,
which tests the scalar ON THE DEVICE and does nothing when it is exactly 1 - the case of `loss.backward()`, i.e. of
every call site of the reference; for N > 1, one NCCL all-reduce of that vector per step between them.

  value          whole-job matrices/s with x, R_true resident in HBM (CUDA events, max over ranks), BASELINE config 2
                 per GPU (weak scaling over N)
  e2e            same step through the public API (rotation_representation.svd_frobenius_loss + autograd) with HOST
                 pinned inputs: H2D of x and R_true and D2H of loss + mean angle inside the timed region (the copies of
                 step i+1 are issued on a copy stream while step i computes, as a data loader would)
  roofline       dominant kernel (the step kernel): algorithmic bytes / event-timed duration vs MEASURED_PEAKS.json;
                 roofline.pair_180B = the literal 180 B definition of SURVEY 8(d) (svd_orth fwd + its bwd from M, G)
  dropin         the UNCHANGED call surface timed as the reference writes it:
                 loss_frobenius(symmetric_orthogonalization(x), R).backward()
  configs        BASELINE configs 3 (maps + metrics, 2^24) and 4 (Kabsch, 65536 x 1024): event-timed kernels
  config5        BASELINE config 5: 2^30 matrices in total sharded over the N GPUs (strong scaling), same step +
                 all-reduce; skipped with --no-config5 or when it does not fit
  cpu_baseline   the oracle port (= the reference's torch code) in fp32 on this box's host cores, per BASELINE.md 3
  --impl reference   times that CPU implementation as the headline instead (rank 0 only)
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "svd_orthogonalizations_per_sec_fwd_bwd"
UNIT = "matrices/s"
BYTES_FWD, BYTES_BWD = 72, 108            # algorithmic bytes per matrix per launch
BYTES_STEP, BYTES_SCALE = 108, 72         # single-pass step kernel; scaling pass of its backward
N_DEFAULT = 1 << 24                       # BASELINE config 2: 16M matrices on one B200
TOTAL_CONFIG5 = 1 << 30                   # BASELINE config 5
# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the step kernel at n = 2^24 (ncu --set full).  Not
# measured by this run - it is the committed capture named in TRAFFIC_SOURCE, scaled by n.
TRAFFIC_STEP, TRAFFIC_SOURCE = 1.7552e9, "profiles/r1m_ncu_full_summary.csv (ncu --set full, 1.208 GB read + 0.548 GB written)"
_tp = os.path.join(ROOT, "profiles", "step_kernel_traffic.json")
if os.path.exists(_tp):                   # refreshed by tools/ncu_summary.py from the latest capture
    try:
        _t = json.load(open(_tp))
        TRAFFIC_STEP, TRAFFIC_SOURCE = float(_t["dram_bytes_per_launch_at_2p24"]), str(_t["source"])
    except (KeyError, ValueError):
        pass


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--n", "--per-gpu", dest="n", type=int, default=N_DEFAULT,
                    help="matrices per GPU (under torchrun spell it --per-gpu: its own parser rejects the prefix --n)")
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=1 << 18, help="matrices per step of --impl reference")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the config 3 / 4 kernel timings and the drop-in leg")
    ap.add_argument("--no-config5", action="store_true", help="skip the 2^30-matrix strong-scaling leg")
    ap.add_argument("--config5-total", type=int, default=TOTAL_CONFIG5)
    ap.add_argument("--total", type=int, default=0,
                    help="headline on a fixed total instead: per-GPU count = total / N (strong scaling); overrides --n")
    ap.add_argument("--lean", action="store_true",
                    help="only the timed step: no explanation kernels, no e2e, no CPU baseline, no extra configs")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks + throttle reasons while the timed region runs (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [t.strip() for t in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except (KeyError, ValueError):
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---------------------------------------------------------------------------------------------------------------------
def cpu_reference_table(threads: int):
    """the reference's CPU implementation (oracle port = the reference's torch calls), fp32,
    n = 4096 (config 1) and n = 2^20, forward+loss and forward+loss+backward, 2 warm-up + best of 5."""
    import torch
    from oracle import reference_port as O
    torch.set_num_threads(threads)
    out = {}
    for label, n in (("config1_n4096", 4096), ("n2p20", 1 << 20)):
        torch.manual_seed(0)
        x = torch.randn(n, 9)
        r_true = O.symmetric_orthogonalization(torch.randn(n, 9))

        def fwd_loss():
            with torch.no_grad():
                return O.loss_frobenius(O.symmetric_orthogonalization(x), r_true)

        def fwd_loss_bwd():
            xr = x.clone().requires_grad_(True)
            O.loss_frobenius(O.symmetric_orthogonalization(xr), r_true).backward()

        row = {"matrices": n}
        for key, fn in (("fwd_loss", fwd_loss), ("fwd_loss_bwd", fwd_loss_bwd)):
            for _ in range(2):
                fn()
            best = float("inf")
            for _ in range(5):
                t0 = time.perf_counter()
                fn()
                best = min(best, time.perf_counter() - t0)
            row[key + "_ms"] = best * 1e3
            row[key + "_matrices_per_s"] = n / best
        out[label] = row
    return out


def run_reference(args):
    """--impl reference: the CPU implementation is the thing measured.  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    from oracle import reference_port as O
    torch.set_num_threads(cores)
    n = args.cpu_sample
    g = torch.Generator().manual_seed(0)
    x = torch.randn(n, 9, generator=g)
    r_true = O.symmetric_orthogonalization(torch.randn(n, 9, generator=g))

    def step():
        xr = x.clone().requires_grad_(True)
        O.loss_frobenius(O.symmetric_orthogonalization(xr), r_true).backward()

    for _ in range(max(args.warmup, 1)):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    val = n / dt
    sample = f"{n} matrices per step (bounded sample of the {args.n}-matrix workload; CPU throughput is flat in n)"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "svd_orth+frobenius fwd+bwd (reference torch code on host cores)",
                   "matrices_per_step": n},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                         "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ---------------------------------------------------------------------------------------------------------------------
def bind_near_gpu(dev):
    """Pin the calling thread to the CPUs NVML reports as local to `dev`, so that the pinned host buffers of the e2e
    leg are allocated (first touch) on the GPU's own NUMA node.  Returns (previous affinity or None, description)."""
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(dev).uuid)
        handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        prev = os.sched_getaffinity(0)
        words = (max(prev) // 64) + 1
        mask = pynvml.nvmlDeviceGetCpuAffinity(handle, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1} & prev
        if not cpus or cpus == prev:
            return None, "nvml: no narrower CPU set than the process already has"
        os.sched_setaffinity(0, cpus)
        return prev, f"nvml: {len(cpus)} of {len(prev)} CPUs local to the GPU"
    except Exception as e:  # no NVML, no permission: leave the placement to the OS
        return None, f"unavailable ({type(e).__name__})"


class Ctx:
    """What every leg needs: torch, the library, rank / world, device, stream."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.args = torch, dist, args
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device - the native arm has no CPU path")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.stdout_fd = None
        if self.world > 1:
            # NCCL (C code) writes its version / debug lines straight to file descriptor 1.  stdout must carry exactly
            # one JSON line, so fd 1 points at stderr until that line is printed.
            sys.stdout.flush()
            self.stdout_fd = os.dup(1)
            os.dup2(2, 1)
            import datetime
            # a mismatch in the collective sequence must fail in five minutes, not in NCCL's default ten (ranks of a fresh box can start a minute apart)
            dist.init_process_group("nccl", device_id=self.dev, timeout=datetime.timedelta(seconds=300))
        from poseestimation_b200 import _lib, launch_count
        import poseestimation_b200.rotation_representation as rr
        self.lib, self.check, self.launch_count, self.rr = _lib.lib, _lib.check, launch_count, rr
        _lib.init_device(self.local)
        self.stream = torch.cuda.current_stream().cuda_stream
        self.peak, self.peak_src = measured_peak_gbs()

    def sync_all(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, values):
        t = self.torch.tensor(values, dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(v) for v in t.tolist()]

    def time_kernel(self, fn, reps=10, warm=2):
        torch = self.torch
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    def free(self):
        import gc
        gc.collect()
        self.torch.cuda.empty_cache()


class StepLoop:
    """K steps of: metrics step kernel -> (N > 1: asynchronous all-reduce of its 24-byte vector) -> backward's scaling
    (upstream gradient 1: a device-side no-op), on resident x, R_true.  Used by the headline leg and by config 5."""

    def __init__(self, c: Ctx, n: int, seed: int):
        torch = c.torch
        self.c, self.n = c, n
        gen = torch.Generator(device=c.dev).manual_seed(seed)
        self.x = torch.randn(n, 9, device=c.dev, generator=gen)
        self.r_true = torch.randn(n, 9, device=c.dev, generator=gen)
        c.check(c.lib.pe_svd_orth_fwd(self.r_true.data_ptr(), 9, self.r_true.data_ptr(), None, n, c.stream),
                "targets")                              # in place: rotations as targets
        self.dx_unit = torch.empty(n, 9, device=c.dev)
        # [sum |.|_F, sum angle_error, range flag]: the all-reduced vector; two of them, because the all-reduce of step
        # i is asynchronous and overlaps step i + 1 (nothing in a step depends on the reduced values: 1/B_global is known)
        self.vec = [torch.zeros(3, dtype=torch.float64, device=c.dev) for _ in range(2)]
        self.gout = torch.ones(1, dtype=torch.float32, device=c.dev)
        self.inv_n_global = 1.0 / (n * c.world)
        self.pend = [None, None]

    def k_step(self, buf):
        c = self.c
        c.check(c.lib.pe_svd_orth_frobenius_step_metrics(self.x.data_ptr(), 9, self.r_true.data_ptr(),
                                                         self.inv_n_global, self.dx_unit.data_ptr(), buf.data_ptr(),
                                                         None, None, None, None, self.n, c.stream), "step")

    def k_bwd(self):
        c = self.c
        c.check(c.lib.pe_scale_by_scalar_inplace(self.dx_unit.data_ptr(), self.gout.data_ptr(), 9 * self.n, c.stream),
                "scale")

    def step(self, i, ev=None):
        b = i % 2
        if self.pend[b] is not None:
            self.pend[b].wait()              # the all-reduce of step i - 2: frees this buffer (long finished)
        if ev:
            ev[0].record()
        self.k_step(self.vec[b])
        if ev:
            ev[1].record()
        if self.c.world > 1:                 # 24 bytes, asynchronous: overlaps the rest of this step and the next one
            self.pend[b] = self.c.dist.all_reduce(self.vec[b], async_op=True)
        self.k_bwd()
        if ev:
            ev[2].record()

    def drain(self):
        for b in range(2):
            if self.pend[b] is not None:
                self.pend[b].wait()
                self.pend[b] = None

    def timed(self, steps: int, warmup: int, sampler: ClockSampler | None = None):
        """-> dict(total_ms, step_ms (kernel), bwd_ms, launches, means) ; times are max over ranks."""
        c, torch = self.c, self.c.torch
        for i in range(warmup):
            self.step(i)
        self.drain()
        c.sync_all()
        if sampler is not None:
            sampler.start()
            time.sleep(0.25)
        # the sampler's start-up left rank 0's GPU idle for a quarter of a second: three more untimed steps ON EVERY RANK
        # (each step carries a collective), so that the timed region starts on devices that are awake, then the barrier +
        # synchronize the contract asks for
        for i in range(3):
            self.step(i)
        self.drain()
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)]
        launches0 = c.launch_count()
        c.sync_all()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for i in range(steps):
            self.step(i, ev[i])
        self.drain()                         # every all-reduce of the K steps completes inside the timed region
        t1.record()
        c.sync_all()
        launches = c.launch_count() - launches0
        total = t0.elapsed_time(t1)
        per_k = [e[0].elapsed_time(e[1]) for e in ev]
        k_ms = statistics.mean(per_k)
        b_ms = statistics.mean(e[1].elapsed_time(e[2]) for e in ev)
        # SURVEY 8d config 2 asks for median and best next to the mean (this rank's; the mean is the max over ranks)
        spread = {"step_kernel_ms_median": statistics.median(per_k), "step_kernel_ms_best": min(per_k),
                  "step_kernel_ms_worst": max(per_k)}
        clocks = sampler.stop() if sampler is not None else None
        vec = self.vec[(steps - 1) % 2].tolist()
        total, k_ms, b_ms = c.max_over_ranks([total, k_ms, b_ms])
        nglob = self.n * c.world
        return {"total_ms": total, "ms_per_step": total / steps, "kernel_ms": k_ms, "bwd_ms": b_ms,
                "launches": int(launches), "clocks": clocks, "spread": spread,
                "loss": vec[0] / nglob, "mean_angle_error_deg": vec[1] / nglob, "range_flag": vec[2]}


# ---------------------------------------------------------------------------------------------------------------------
def leg_explain(c: Ctx, L: StepLoop):
    """roofline.pair_180B and the other single-GPU explanation numbers around the headline (device-resident)."""
    torch, lib, check, n, st = c.torch, c.lib, c.check, L.n, c.stream
    x, rt = L.x, L.r_true
    r_out, dx = torch.empty(n, 9, device=c.dev), torch.empty(n, 9, device=c.dev)
    g_up = torch.randn(n, 9, device=c.dev)
    sums = torch.zeros(3, dtype=torch.float64, device=c.dev)
    P = lambda t: t.data_ptr()
    o = {}
    o["svd_orth_fwd_ms"] = c.time_kernel(lambda: check(lib.pe_svd_orth_fwd(P(x), 9, P(r_out), None, n, st), "f"))
    o["svd_orth_bwd_recompute_ms"] = c.time_kernel(
        lambda: check(lib.pe_svd_orth_bwd(P(x), 9, None, P(g_up), P(dx), n, st), "b"))
    o["svd_orth_bwd_savedR_ms"] = c.time_kernel(
        lambda: check(lib.pe_svd_orth_bwd(P(x), 9, P(r_out), P(g_up), P(dx), n, st), "b"))
    o["step_without_metric_ms"] = c.time_kernel(
        lambda: check(lib.pe_svd_orth_frobenius_step(P(x), 9, P(rt), 1.0 / n, P(dx), P(sums), n, st), "s"))
    o["backward_upstream_not_one_ms"] = c.time_kernel(
        lambda: check(lib.pe_scale_by_scalar(P(L.dx_unit), P(L.gout), P(dx), 9 * n, st), "scale"))
    o["recompute_fwd_ms"] = c.time_kernel(
        lambda: check(lib.pe_svd_orth_frobenius_fwd(P(x), 9, P(rt), None, P(sums), n, st), "f"))
    o["recompute_bwd_ms"] = c.time_kernel(
        lambda: check(lib.pe_svd_orth_frobenius_bwd(P(x), 9, P(rt), P(L.gout), 1.0 / n, P(dx), n, st), "b"))
    # SURVEY config 2, second distribution: near-rotations R + 0.1 randn (what a trained network emits)
    near = rt + 0.1 * torch.randn(n, 9, device=c.dev)
    o["svd_orth_fwd_near_rotation_ms"] = c.time_kernel(
        lambda: check(lib.pe_svd_orth_fwd(P(near), 9, P(r_out), None, n, st), "f"))
    # the reference's real input layout: out[:, :9] of a (B, 12) network output (3D-Pose/main2.py:94)
    wide = torch.randn(n, 12, device=c.dev)
    o["step_on_strided_view_ms"] = c.time_kernel(
        lambda: check(lib.pe_svd_orth_frobenius_step_metrics(P(wide), 12, P(rt), 1.0 / n, P(dx), P(sums), None, None,
                                                             None, None, n, st), "s"))
    o["svd_orth_fwd_on_strided_view_ms"] = c.time_kernel(
        lambda: check(lib.pe_svd_orth_fwd(P(wide), 12, P(r_out), None, n, st), "f"))
    del near, wide, r_out, dx, g_up
    pair_ms = o["svd_orth_fwd_ms"] + o["svd_orth_bwd_recompute_ms"]
    pair = {"definition": "svd_orth forward (72 B) + its backward from M and G (108 B): SURVEY 8(d)'s 180 B/matrix",
            "fwd_ms": o["svd_orth_fwd_ms"], "bwd_ms": o["svd_orth_bwd_recompute_ms"], "ms": pair_ms,
            "matrices_per_s": n / (pair_ms * 1e-3), "frac": 180 * n / (pair_ms * 1e-3) / 1e9 / c.peak,
            "fwd_frac": 72 * n / (o["svd_orth_fwd_ms"] * 1e-3) / 1e9 / c.peak,
            "bwd_frac": 108 * n / (o["svd_orth_bwd_recompute_ms"] * 1e-3) / 1e9 / c.peak}
    return pair, o


def leg_dropin(c: Ctx, L: StepLoop):
    """The unchanged call surface, exactly as the reference writes its step (Comparison/main.py:59-62)."""
    torch, rr, n = c.torch, c.rr, L.n
    xd = L.x.detach().requires_grad_(True)
    rt = L.r_true.view(n, 3, 3)

    def step():
        xd.grad = None
        out_r = rr.symmetric_orthogonalization(xd)
        loss = rr.loss_frobenius(out_r, rt)
        loss.backward()

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    reps = 5
    l0 = c.launch_count()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        step()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / reps
    launches = (c.launch_count() - l0) / reps
    del xd
    return {"api": "loss_frobenius(symmetric_orthogonalization(x), R).backward()", "ms_per_step": ms,
            "matrices_per_s": n / (ms * 1e-3), "launches": launches,
            "bytes_per_matrix": "72 (svd_orth fwd, R materialised) + 108 (single-pass step on the node's own input)"}


def leg_e2e(c: Ctx, L: StepLoop):
    torch, rr, n, dev, world = c.torch, c.rr, L.n, c.dev, c.world
    # host buffers on the GPU's own NUMA node: only with several ranks (one rank keeps every core for its threads)
    prev_affinity, numa_note = bind_near_gpu(dev) if world > 1 else (None, "not set (single rank)")
    hx = torch.randn(n, 9).pin_memory()
    ht = L.r_true.cpu().pin_memory()
    hloss = torch.empty((), dtype=torch.float32).pin_memory()
    hang = torch.empty((), dtype=torch.float64).pin_memory()
    if prev_affinity is not None:
        os.sched_setaffinity(0, prev_affinity)         # the CPU baseline uses every host core again

    # Input pipeline as a data loader would run it: the copies of step i + 1 go out on a copy stream while the
    # kernels of step i run (two device buffer pairs).  Every step still moves its own 2 x 36 B per matrix from
    # pinned host memory inside the timed region, and the region ends only when both streams have drained.
    cur = torch.cuda.current_stream()
    copy_stream = torch.cuda.Stream(device=dev)
    dbuf = [(torch.empty(n, 9, device=dev), torch.empty(n, 9, device=dev)) for _ in range(2)]
    ready = [torch.cuda.Event() for _ in range(2)]
    free = [torch.cuda.Event() for _ in range(2)]
    for e in free:
        e.record(cur)

    def prefetch(i):
        b = i & 1
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(free[b])          # the step that last used this pair has finished
            dbuf[b][0].copy_(hx, non_blocking=True)
            dbuf[b][1].copy_(ht, non_blocking=True)
            ready[b].record(copy_stream)

    def e2e_step(i):
        b = i & 1
        prefetch(i + 1)
        cur.wait_event(ready[b])
        xd = dbuf[b][0].detach().requires_grad_(True)
        loss, ang = rr.svd_frobenius_loss(xd, dbuf[b][1], with_angle=True, check=False)
        loss.backward()
        hloss.copy_(loss.detach(), non_blocking=True)
        hang.copy_(ang, non_blocking=True)
        free[b].record(cur)
        return xd.grad

    prefetch(0)
    for i in range(2):
        e2e_step(i)
    cur.wait_stream(copy_stream)
    c.sync_all()
    k = max(3, min(c.args.steps, 5))
    k += k & 1                                       # even: the buffer parity is the same at both ends
    prefetch(0)                                      # step 0's inputs; the region then holds k sets of copies
    cur.wait_stream(copy_stream)
    c.sync_all()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(k):
        e2e_step(i)
    cur.wait_stream(copy_stream)                     # k steps and k sets of copies (those of steps 1 .. k)
    b.record()
    c.sync_all()
    (ms,) = c.max_over_ranks([a.elapsed_time(b) / k])
    return {"value": n * world / (ms * 1e-3), "unit": UNIT,
            "h2d_bytes_per_step": 2 * n * 36, "d2h_bytes_per_step": 12, "ms_per_step": ms,
            "h2d_gbs_per_gpu": 2 * n * 36 / (ms * 1e-3) / 1e9,
            "api": "rotation_representation.svd_frobenius_loss(x, R_true, with_angle=True).backward(); loss and mean "
                   "angle error read back",
            "host_buffers": "pinned; CPU affinity while allocating: " + numa_note,
            "pipeline": "H2D copies of step i+1 overlap the kernels of step i (copy stream, two device buffer pairs)",
            "limiter": "the host-to-device copies (PCIe); copy-only ceiling per N: profiles/r2_h2d_ceiling.json"}


def leg_config3(c: Ctx, n: int):
    """BASELINE config 3: Gram-Schmidt 6D + quaternion / Euler maps with geodesic mean-angle-error, 2^24 batch."""
    torch, lib, check, st, dev = c.torch, c.lib, c.check, c.stream, c.dev
    g = torch.Generator(device=dev).manual_seed(0)
    P = lambda t: t.data_ptr()
    rt = torch.randn(n, 9, device=dev, generator=g)
    check(lib.pe_svd_orth_fwd(P(rt), 9, P(rt), None, n, st), "targets")
    o9, g9 = torch.empty(n, 9, device=dev), torch.randn(n, 9, device=dev, generator=g)
    sums = torch.zeros(3, dtype=torch.float64, device=dev)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    rows = {}

    def add(name, nbytes, fn):
        ms = c.time_kernel(fn)
        rows[name] = {"bytes": nbytes, "ms": ms, "frac": nbytes * n / (ms * 1e-3) / 1e9 / c.peak,
                      "records_per_s": n / (ms * 1e-3)}

    for name, mid, w in (("ortho6d", 1, 6), ("quat", 2, 4), ("euler", 3, 3)):
        xi = torch.randn(n, w, device=dev, generator=g) if name != "euler" else \
            (torch.rand(n, w, device=dev, generator=g) * 2 - 1) * 3.14159265
        dxi = torch.empty(n, w, device=dev)
        f, b = getattr(lib, f"pe_{name}_fwd"), getattr(lib, f"pe_{name}_bwd")
        add(f"{name}_fwd", 4 * w + 36, lambda: check(f(P(xi), w, P(o9), n, st), ""))
        add(f"{name}_bwd", 8 * w + 36, lambda: check(b(P(xi), w, P(g9), P(dxi), n, st), ""))
        add(f"{name}+angle_error_mean (fused, f64)", 4 * w + 36,
            lambda: check(lib.pe_map_metric_fwd(mid, 2, P(xi), w, P(rt), -1.0, 1.0, P(sums), P(flag), n, st), ""))
        add(f"{name}+geodesic_mean (fused)", 4 * w + 36,
            lambda: check(lib.pe_map_metric_fwd(mid, 1, P(xi), w, P(rt), -1.0 + 1e-7, 1.0 - 1e-7, P(sums), None, n, st), ""))
        del xi, dxi
    x9 = torch.randn(n, 9, device=dev, generator=g)
    add("svd+angle_error_mean (fused, f64)", 72,
        lambda: check(lib.pe_map_metric_fwd(0, 2, P(x9), 9, P(rt), -1.0, 1.0, P(sums), P(flag), n, st), ""))
    check(lib.pe_svd_orth_fwd(P(x9), 9, P(x9), None, n, st), "r")
    add("angle_error_mean (f64)", 72, lambda: check(lib.pe_angle_error_fwd(P(x9), P(rt), None, P(sums), P(flag), n, st), ""))
    add("geodesic_mean", 72,
        lambda: check(lib.pe_geodesic_fwd(P(x9), P(rt), -1.0 + 1e-7, 1.0 - 1e-7, None, P(sums), n, st), ""))
    return {"workload": f"{n} records per kernel, resident, CUDA events, 10 reps (BASELINE config 3)", "kernels": rows}


def leg_config4(c: Ctx):
    """BASELINE config 4: point_cloud Kabsch, 65536 synthetic cloud pairs x 1024 points."""
    torch, lib, check, st, dev = c.torch, c.lib, c.check, c.stream, c.dev
    import poseestimation_b200.point_cloud as pc
    b, npts = 65536, 1024
    g = torch.Generator(device=dev).manual_seed(0)
    p = torch.rand(b, npts, 3, device=dev, generator=g) - 0.5
    rot, q = pc.make_cloud_pair(p)
    r, t = torch.empty(b, 9, device=dev), torch.empty(b, 3, device=dev)
    ms = c.time_kernel(lambda: check(lib.pe_kabsch_fwd(p.data_ptr(), q.data_ptr(), b, npts, r.data_ptr(), t.data_ptr(),
                                                       st), "kabsch"))
    mae = float(c.rr.mean_angle_error(r.view(b, 3, 3), rot).item())
    nbytes = 24 * npts + 48
    return {"workload": "65536 cloud pairs x 1024 points, Q = R_gt P (BASELINE config 4)", "bytes_per_pair": nbytes,
            "ms": ms, "pairs_per_s": b / (ms * 1e-3), "point_pairs_per_s": b * npts / (ms * 1e-3),
            "frac": nbytes * b / (ms * 1e-3) / 1e9 / c.peak, "mean_angle_error_vs_R_gt_deg": mae}


def leg_config5(c: Ctx, total: int, weak_per_gpu_rate: float):
    """BASELINE config 5: `total` matrices sharded over the N GPUs, the same step + all-reduce (strong scaling)."""
    torch, world = c.torch, c.world
    if total % world:
        return {"skipped": f"{total} not divisible by {world} ranks"}
    n = total // world
    need = 3 * n * 36 + (1 << 30)
    free_b, _ = torch.cuda.mem_get_info()
    fits = c.max_over_ranks([0.0 if free_b >= need else 1.0])[0] == 0.0
    if not fits:
        return {"skipped": f"needs {need / 1e9:.0f} GB per GPU, {free_b / 1e9:.0f} GB free"}
    L = StepLoop(c, n, seed=1000 + c.rank)
    steps = 6
    r = L.timed(steps, warmup=3)
    del L
    c.free()
    mps = total / (r["ms_per_step"] * 1e-3)
    out = {"workload": f"{total} matrices in total, {n} per GPU on {world} GPU(s), resident (3 arrays, "
                       f"{3 * n * 36 / 1e9:.1f} GB per GPU), step + 24-byte all-reduce",
           "steps": steps, "warmup": 3, "ms_per_step": r["ms_per_step"], "kernel_ms": r["kernel_ms"],
           "matrices_per_s": mps, "scaling": "strong",
           "frac_108B": BYTES_STEP * total / world / (r["ms_per_step"] * 1e-3) / 1e9 / c.peak,
           "frac_180B": 180 * total / world / (r["ms_per_step"] * 1e-3) / 1e9 / c.peak,
           "per_gpu_rate_vs_config2_leg": (mps / world) / weak_per_gpu_rate,
           "loss": r["loss"], "mean_angle_error_deg": r["mean_angle_error_deg"]}
    ref = os.path.join(ROOT, "profiles", "r2_config5_n1.json")
    out["efficiency_vs_n1_ref"] = None
    if world == 1:
        out["efficiency_vs_n1_ref"] = 1.0
    elif os.path.exists(ref):
        try:
            n1 = float(json.load(open(ref))["matrices_per_s"])
            out["efficiency_vs_n1_ref"] = mps / (world * n1)
            out["n1_ref"] = {"matrices_per_s": n1, "source": "profiles/r2_config5_n1.json"}
        except (KeyError, ValueError):
            pass
    return out


def run_native(args):
    c = Ctx(args)
    torch, world, rank = c.torch, c.world, c.rank
    n = args.n
    if args.total:
        if args.total % world:
            raise SystemExit("--total must be divisible by the number of GPUs")
        n = args.total // world
    if args.lean:
        args.no_e2e = args.no_cpu_baseline = args.no_configs = args.no_config5 = True

    # ---- headline: K timed steps, device-timed, with per-kernel events for the roofline --------------------------
    L = StepLoop(c, n, seed=rank)
    warm = max(args.warmup, 3)
    r = L.timed(args.steps, warm, ClockSampler(c.local) if rank == 0 else None)
    ms_per_step = r["ms_per_step"]
    value = n * world / (ms_per_step * 1e-3)
    achieved = BYTES_STEP * n / (r["kernel_ms"] * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": "stream_tma_kernel<SvdFrobStepT<true>> (pe_svd_orth_frobenius_step_metrics)",
                "achieved": achieved, "peak": c.peak, "unit": "GB/s", "frac": achieved / c.peak,
                "traffic": TRAFFIC_STEP * n / N_DEFAULT, "traffic_source": TRAFFIC_SOURCE,
                "peak_source": c.peak_src, "bytes_per_matrix": BYTES_STEP, "ms": r["kernel_ms"],
                "step_frac": BYTES_STEP * n / (ms_per_step * 1e-3) / 1e9 / c.peak,
                # informational: this HBM reads faster than it writes (tools/ubench/mix_rw.cu, profiles/r2p_mix_rw.txt:
                # read only 7111 GB/s, write only 6000 GB/s from plain streaming kernels); the additive bound for the
                # step's 72 B read + 36 B written per matrix.  `frac` above stays on the driver's 1:1 copy figure.
                "mix_aware": {"read_gbs": 7111.0, "write_gbs": 6000.0, "source": "profiles/r2p_mix_rw.txt",
                              "bound_ms": n * (72 / 7111.0 + 36 / 6000.0) * 1e-6,
                              "frac": n * (72 / 7111.0 + 36 / 6000.0) * 1e-6 / r["kernel_ms"]}}
    also = {"step_kernel_ms": r["kernel_ms"], **r["spread"], "backward_upstream_one_ms": r["bwd_ms"], "loss": r["loss"],
            "mean_angle_error_deg": r["mean_angle_error_deg"], "range_flag": r["range_flag"]}

    dropin = configs = None
    if not args.lean:
        pair, more = leg_explain(c, L)
        roofline["pair_180B"] = pair
        also.update(more)
        also["matrices_per_s_upstream_not_one"] = n / ((r["kernel_ms"] + more["backward_upstream_not_one_ms"]) * 1e-3)
    if not args.no_configs:
        dropin = leg_dropin(c, L)
    e2e = None if args.no_e2e else leg_e2e(c, L)
    del L
    c.free()
    if not args.no_configs:
        configs = {"config3": leg_config3(c, n)}
        c.free()
        configs["config4"] = leg_config4(c)
        c.free()
    config5 = None
    if not args.no_config5:
        config5 = leg_config5(c, args.config5_total, value / world)

    if rank != 0:
        if world > 1:
            c.dist.destroy_process_group()
        return

    cpu = None
    if not args.no_cpu_baseline and world == 1:
        cores = os.cpu_count() or 1
        table = cpu_reference_table(cores)
        cpu = {"value": table["n2p20"]["fwd_loss_bwd_matrices_per_s"], "unit": UNIT, "cores": cores,
               "torch_threads": torch.get_num_threads(), "kind": "port",
               "sample": "BASELINE.md 3: n = 4096 (config 1) and n = 2^20, fp32, 2 warm-up + best of 5, forward+loss and "
                         "forward+loss+backward; value = forward+loss+backward at n = 2^20; oracle port = the "
                         "reference's own torch calls",
               **table}

    if c.stdout_fd is not None:
        sys.stdout.flush()
        os.dup2(c.stdout_fd, 1)
        os.close(c.stdout_fd)
    print(json.dumps({
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if args.total else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"svd_orth+frobenius fwd+bwd with fused mean angle error, {n} random 3x3 fp32 matrices "
                               "per GPU " + (f"({n * world} in total, sharded)" if args.total else "(BASELINE config 2)"),
                   "launches_per_step": "svd_orth_frobenius_step_metrics (108 B: loss, mean angle error and gradient from "
                                        "one SVD) + scale_by_scalar_inplace (upstream gradient 1, as under loss.backward(): "
                                        "device-side no-op).  With an upstream gradient other than 1 (a weighted sum of "
                                        "losses, 3D-Pose/main2.py:94-103) the backward is a 72 B/matrix scaling pass: see "
                                        "also.backward_upstream_not_one_ms / also.matrices_per_s_upstream_not_one",
                   "matrices_per_gpu": n, "total_matrices": n * world,
                   "l2": "inputs (2 x %.0f MB) larger than the 126 MB L2" % (n * 36 / 1e6),
                   "collective": "ncclAllReduce(3 x f64: sum |.|_F, sum angle_error, range flag) per step, asynchronous "
                                 "(overlaps the next step's kernel; all K complete inside the timed region)"
                                 if world > 1 else "none"},
        "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": r["launches"], "clocks": r["clocks"],
        "dropin": dropin, "configs": configs, "config5": config5, "also": also,
    }))
    if world > 1:
        c.dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_native(a)
