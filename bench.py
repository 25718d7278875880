#!/usr/bin/env python
"""bench.py — headline benchmark of the ranking & counting metric hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Primary workload (BASELINE.json metric, configs[3]): exact BinaryAUROC over N = 2^30 synthetic
float32 scores + int64 {0,1} labels, rank-sharded over the GPUs (strong scaling: the total is
fixed, each of N ranks holds 2^30 / N).  One "step" = one full compute of the metric over the
resident scores.  Also in the line, at 1 GPU: "secondary" (configs[1]: MulticlassConfusionMatrix
update, C = 1000, 2^28 label pairs, GB/s against the HBM roofline), "secondary_binned" (configs[2]:
10k-threshold binned counting over the same 2^30 scores) and "secondary_ovr" (configs[4] shape:
one-vs-rest MulticlassAUROC, C = 1000, N = 2^20); at N GPUs: "secondary_binned" = binned update +
counter all-reduce through toolkit.sync_and_compute.

Prints ONE JSON line on rank 0 (see the task contract): value = device-timed whole-job
throughput with inputs resident in HBM; e2e = same metric through the public functional API
with pinned HOST tensors (H2D + D2H inside the timed region); roofline = dominant kernel
(one radix pass), timed with CUDA events on the launching stream during the timed steps;
cpu_baseline = the CPU port of the reference path (oracle/oracle_torch.py) on the host cores.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import statistics
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

LOG2_N_TOTAL = int(os.environ.get("TE_BENCH_LOG2_N", "30"))
CM_LOG2_N = int(os.environ.get("TE_BENCH_CM_LOG2_N", "28"))
CM_CLASSES = 1000
SEED = 20261106


def measured_peaks() -> tuple[float, str]:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json, copy r+w)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel: str, n_records: int):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture, or None when
    the capture was taken on a different record count than this run."""
    p = ROOT / "profiles" / "r01_ncu_traffic.json"
    try:
        t = json.loads(p.read_text())
        return int(t[kernel]) if int(t["n_records"]) == int(n_records) else None
    except Exception:
        return None


# ----------------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._nv = None

    def _run(self):
        nv = self._nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.05)

    def __enter__(self):
        if self._nv is not None:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()

    def summary(self) -> dict:
        return {
            "sm_mhz": statistics.median(self.samples) if self.samples else None,
            "sm_max_mhz": self.max_mhz,
            "reasons": sorted(self.reasons),
        }


# ----------------------------------------------------------------------------------- profiling
class KernelProfile:
    """Per-kernel CUDA-event timing inside libte_b200 (te_profile_enable / te_profile_read)."""

    def __init__(self):
        from torcheval_b200 import _native

        self.lib = _native.load()

    def __enter__(self):
        self.lib.te_profile_enable(1)
        return self

    def __exit__(self, *exc):
        pass

    def read(self) -> dict:
        names = ctypes.create_string_buffer(8192)
        ms = (ctypes.c_double * 64)()
        cnt = (ctypes.c_int64 * 64)()
        n = self.lib.te_profile_read(names, 8192, ctypes.cast(ms, ctypes.c_void_p), ctypes.cast(cnt, ctypes.c_void_p), 64)
        out = {}
        for i, name in enumerate(names.value.decode().split("\n")[:n]):
            out[name] = {"total_ms": ms[i], "launches": int(cnt[i])}
        self.lib.te_profile_enable(0)
        return out


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


# ----------------------------------------------------------------------------------- data
def make_scores(n: int, device, rank: int):
    g = torch.Generator(device=device).manual_seed(SEED + rank)
    x = torch.rand(n, device=device, generator=g, dtype=torch.float32)
    y = torch.randint(0, 2, (n,), device=device, generator=g, dtype=torch.int64)
    return x, y


def max_over_ranks(ms: float, world: int, device) -> float:
    if world == 1:
        return ms
    import torch.distributed as dist

    t = torch.tensor([ms], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world: int):
    if world > 1:
        import torch.distributed as dist

        dist.barrier()
    torch.cuda.synchronize()


# ----------------------------------------------------------------------------------- our arm
def run_ours(args) -> dict:
    rank, world, local = dist_env()
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node N for --gpus N > 1")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=device)

    from torcheval_b200 import ops
    from torcheval_b200.metrics.functional.classification.auroc import binary_auroc

    peak_gbs, peak_src = measured_peaks()
    n_total = 1 << LOG2_N_TOTAL
    n_local = n_total // world
    x, y = make_scores(n_local, device, rank)

    if world > 1:
        from torcheval_b200 import distributed as tdist

        def step():
            return tdist.sharded_binary_auroc(x, y)
    else:
        def step():
            return ops.binary_auroc_packed(x, y)

    for _ in range(args.warmup):
        out = step()
    barrier(world)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks, KernelProfile() as prof:
        barrier(world)
        ev0.record()
        for _ in range(args.steps):
            out = step()
        ev1.record()
        barrier(world)
        kernels = prof.read()
    ms_total = max_over_ranks(ev0.elapsed_time(ev1), world, device)
    ms_step = ms_total / args.steps
    value = n_total / (ms_step * 1e-3)
    auroc = float(out[0].item())
    launches = sum(k["launches"] for k in kernels.values())

    # roofline of the dominant kernel: one radix pass reads and writes every 4-byte record once
    dom = "onesweep_pass_keys"
    roof = None
    if dom in kernels and kernels[dom]["launches"]:
        k = kernels[dom]
        avg_ms = k["total_ms"] / k["launches"]
        sorted_n = n_local if world == 1 else None
        if sorted_n is None:
            sorted_n = int(out[5].item()) if out.numel() > 5 else n_local
        achieved = 8.0 * sorted_n / (avg_ms * 1e-3) / 1e9
        roof = {
            "kernel": dom,
            "bound": "hbm",
            "achieved": round(achieved, 1),
            "peak": peak_gbs,
            "peak_source": peak_src,
            "unit": "GB/s",
            "frac": round(achieved / peak_gbs, 4),
            "traffic": ncu_traffic(dom, sorted_n),
            "avg_launch_ms": round(avg_ms, 4),
            "algorithmic_bytes_per_launch": 8 * sorted_n,
            "step_share": round(k["total_ms"] / (ms_total), 4),
        }
    # bytes the structure that ran actually moves per score: counting-stage path
    # 16 ingest + 2 passes * 8 + 4 counting = 36; full sort 16 + 4 * 8 + 4 sweep = 52
    step_bytes = (36.0 if "leaf_count" in kernels else 52.0) * n_local
    kernel_table = {
        name: {"avg_ms": round(v["total_ms"] / max(v["launches"], 1), 4), "launches_per_step": v["launches"] / args.steps}
        for name, v in kernels.items()
    }

    # ---------------- end to end through the public functional API with pinned host tensors
    e2e = None
    cpu = None
    secondary = None
    binned = None
    ovr = None
    if world == 1:
        binned = binned_bench(args, x, y, peak_gbs, peak_src)
    else:
        binned = binned_sync_bench(args, x, y, world, device)
    if world == 1:
        xh = torch.empty(n_local, dtype=torch.float32, pin_memory=True)
        yh = torch.empty(n_local, dtype=torch.int64, pin_memory=True)
        xh.copy_(x)
        yh.copy_(y)
        torch.cuda.synchronize()
        del x, y
        torch.cuda.empty_cache()
        e2e_steps = max(1, min(args.steps, 3))
        r = binary_auroc(xh, yh)  # warm-up
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            r = binary_auroc(xh, yh)
        e2e_val = float(r.item())
        dt = (time.perf_counter() - t0) / e2e_steps
        assert abs(e2e_val - auroc) <= 1e-9 * max(1.0, abs(auroc)), (e2e_val, auroc)
        e2e = {
            "value": n_total / dt,
            "unit": "scores/s",
            "h2d_bytes_per_step": 12 * n_local,
            "d2h_bytes_per_step": 8,
            "ms_per_step": dt * 1e3,
            "steps": e2e_steps,
        }
        cpu = cpu_baseline_auroc(xh, yh)
        del xh, yh
        secondary = confusion_matrix_bench(args, device, peak_gbs, peak_src)
        ovr = one_vs_rest_bench(args, device)

    if world > 1:
        # end to end at N GPUs through the class API + toolkit: every rank feeds its pinned HOST
        # shard to BinaryAUROC.update (H2D) and all ranks call sync_and_compute (sample sort over NCCL)
        import torch.distributed as dist

        from torcheval_b200.metrics import BinaryAUROC
        from torcheval_b200.metrics.toolkit import sync_and_compute

        xh = torch.empty(n_local, dtype=torch.float32, pin_memory=True).copy_(x)
        yh = torch.empty(n_local, dtype=torch.int64, pin_memory=True).copy_(y)
        torch.cuda.synchronize()
        del x, y
        torch.cuda.empty_cache()

        def e2e_step():
            m = BinaryAUROC(device=device)
            m.update(xh, yh)
            return float(sync_and_compute(m).item())

        e2e_step()
        e2e_steps = max(1, min(args.steps, 3))
        barrier(world)
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_val = e2e_step()
        torch.cuda.synchronize()
        dt = max_over_ranks((time.perf_counter() - t0) * 1e3, world, device) / 1e3 / e2e_steps
        assert abs(e2e_val - auroc) <= 1e-9 * max(1.0, abs(auroc)), (e2e_val, auroc)
        e2e = {
            "value": n_total / dt,
            "unit": "scores/s",
            "h2d_bytes_per_step": 12 * n_local,
            "d2h_bytes_per_step": 8,
            "ms_per_step": dt * 1e3,
            "steps": e2e_steps,
            "note": "per-rank bytes; class API update() + toolkit.sync_and_compute()",
        }

    line = {
        "metric": "exact_binary_auroc_scores_per_sec",
        "value": value,
        "unit": "scores/s",
        "n_gpus": world,
        "steps": args.steps,
        "warmup": args.warmup,
        "ms_per_step": ms_step,
        "higher_is_better": True,
        "scaling": "strong",
        "vs_baseline": None,
        "dtype": "u32",
        "data": "synthetic",
        "config": {
            "workload": f"BinaryAUROC exact, N=2^{LOG2_N_TOTAL} float32 scores + int64 labels, rank-sharded",
            "n_total": n_total,
            "n_per_gpu": n_local,
            "l2": f"inputs {12 * n_local / 1e9:.1f} GB per GPU >> 126 MB L2, no reuse between steps",
            "parallelism": f"dp{world}",
            "auroc": auroc,
        },
        "clocks": clocks.summary(),
        "gpu_launches": launches,
        "kernels": kernel_table,
        "step_hbm": {
            "algorithmic_bytes_per_step": step_bytes,
            "achieved_GBps": round(step_bytes / (ms_step * 1e-3) / 1e9, 1),
            "frac": round(step_bytes / (ms_step * 1e-3) / 1e9 / peak_gbs, 4),
        },
    }
    if roof:
        line["roofline"] = roof
    if e2e:
        line["e2e"] = e2e
    if cpu:
        line["cpu_baseline"] = cpu
    if secondary:
        line["secondary"] = secondary
    if binned:
        line["secondary_binned"] = binned
    if ovr:
        line["secondary_ovr"] = ovr
    if world > 1:
        import torch.distributed as dist

        dist.destroy_process_group()
    return line if rank == 0 else {}


def cpu_baseline_auroc(xh: torch.Tensor, yh: torch.Tensor, log2_sample: int = 25) -> dict:
    from oracle import oracle_torch

    m = min(xh.numel(), 1 << log2_sample)
    xs, ys = xh[:m].clone(), yh[:m].clone()
    oracle_torch.binary_auroc(xs[: 1 << 16], ys[: 1 << 16])
    t0 = time.perf_counter()
    oracle_torch.binary_auroc(xs, ys)
    dt = time.perf_counter() - t0
    return {
        "value": m / dt,
        "unit": "scores/s",
        "cores": torch.get_num_threads(),
        "host_cpus": os.cpu_count(),
        "kind": "port",
        "sample": f"first 2^{m.bit_length() - 1} scores of the same workload, 1 run, {dt:.1f} s (oracle/oracle_torch.py)",
    }


def binned_bench(args, x, y, peak_gbs, peak_src) -> dict:
    """configs[2] on one GPU: BinaryBinnedAUROC/AUPRC counting, 10k thresholds, the same 2^30 scores."""
    from torcheval_b200 import ops

    T = 10_000
    n = x.numel()
    thr = torch.linspace(0, 1, T, device=x.device)
    hist = torch.zeros((1, T + 1, 2), dtype=torch.int64, device=x.device)
    for _ in range(3):
        ops.binned_hist_update_(hist, x, y, thr)
    torch.cuda.synchronize()
    hist.zero_()
    steps = max(args.steps, 5)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with KernelProfile() as prof:
        ev0.record()
        for _ in range(steps):
            ops.binned_hist_update_(hist, x, y, thr)
        ev1.record()
        torch.cuda.synchronize()
        kernels = prof.read()
    ms = ev0.elapsed_time(ev1) / steps
    assert int(hist.sum().item()) == steps * n
    k = kernels.get("binned_hist_smem", {"total_ms": ms * steps, "launches": steps})
    kavg = k["total_ms"] / k["launches"]
    kgbs = 12.0 * n / (kavg * 1e-3) / 1e9
    tp, fp, fn = ops.binned_counts(hist, T)
    auroc = float(ops.binned_integrate(tp, fp, fn, 0)[0].item())
    return {
        "metric": "binary_binned_auroc_update_scores_per_sec",
        "value": n / (ms * 1e-3),
        "unit": "scores/s",
        "ms_per_step": ms,
        "config": {"workload": f"BinaryBinnedAUROC/AUPRC update, T={T} thresholds, N=2^{n.bit_length() - 1} scores, 1 GPU", "binned_auroc": auroc},
        "roofline": {
            "kernel": "binned_hist_smem",
            "bound": "hbm",
            "achieved": round(kgbs, 1),
            "peak": peak_gbs,
            "peak_source": peak_src,
            "unit": "GB/s",
            "frac": round(kgbs / peak_gbs, 4),
            "traffic": None,
            "avg_launch_ms": round(kavg, 4),
        },
    }


def binned_sync_bench(args, x, y, world, device) -> dict:
    """configs[2] at N GPUs: every rank counts its shard of the 2^30 scores into the 10k-threshold
    histogram state of BinaryBinnedAUPRC (kernel), then toolkit.sync_and_compute all-reduces the
    int64 counters over NCCL and integrates.  One step = update + sync_and_compute."""
    from torcheval_b200.metrics import BinaryBinnedAUPRC
    from torcheval_b200.metrics.toolkit import sync_and_compute

    T = 10_000

    def step():
        m = BinaryBinnedAUPRC(threshold=T, device=device)
        m.update(x, y)
        return sync_and_compute(m)

    for _ in range(3):
        out = step()
    barrier(world)
    steps = max(args.steps, 5)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(steps):
        out = step()
    ev1.record()
    barrier(world)
    ms = max_over_ranks(ev0.elapsed_time(ev1), world, device) / steps
    n_total = x.numel() * world
    return {
        "metric": "binary_binned_auprc_update_sync_scores_per_sec",
        "value": n_total / (ms * 1e-3),
        "unit": "scores/s",
        "ms_per_step": ms,
        "scaling": "strong",
        "config": {
            "workload": f"BinaryBinnedAUPRC update + sync_and_compute, T={T} thresholds, N=2^{n_total.bit_length() - 1} scores over {world} GPUs, int64 counter all-reduce",
            "binned_auprc": float(out[0].item()),
        },
    }


def one_vs_rest_bench(args, device) -> dict:
    """configs[4] shape at a size one step finishes in tens of milliseconds: MulticlassAUROC
    one-vs-rest (segmented sort), C = 1000 classes, N = 2^20 samples = 1.05e9 scores on one GPU."""
    from torcheval_b200 import ops

    n, C = 1 << 20, 1000
    g = torch.Generator(device=device).manual_seed(SEED + 5)
    x = torch.rand((n, C), device=device, generator=g, dtype=torch.float32)
    y = torch.randint(0, C, (n,), device=device, generator=g, dtype=torch.int64)
    for _ in range(2):
        out = ops.ovr_rank(x, y, False, 0)
    torch.cuda.synchronize()
    steps = 3
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(steps):
        out = ops.ovr_rank(x, y, False, 0)
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1) / steps
    return {
        "metric": "multiclass_auroc_one_vs_rest_scores_per_sec",
        "value": n * C / (ms * 1e-3),
        "unit": "scores/s",
        "ms_per_step": ms,
        "config": {"workload": f"MulticlassAUROC one-vs-rest, C={C}, N=2^20 samples, float32 (N, C) scores, 1 GPU",
                   "macro_auroc": float(out[:, 0].mean().item())},
    }


def confusion_matrix_bench(args, device, peak_gbs, peak_src) -> dict:
    from oracle import oracle_torch
    from torcheval_b200 import ops

    n, C = 1 << CM_LOG2_N, CM_CLASSES
    g = torch.Generator(device=device).manual_seed(SEED)
    pred = torch.randint(0, C, (n,), device=device, generator=g, dtype=torch.int64)
    target = torch.randint(0, C, (n,), device=device, generator=g, dtype=torch.int64)
    cm = torch.zeros((C, C), dtype=torch.int64, device=device)
    for _ in range(max(3, args.warmup)):
        ops.cm_update_(cm, pred, target, C)
    torch.cuda.synchronize()
    cm.zero_()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = max(args.steps, 5)
    with KernelProfile() as prof:
        ev0.record()
        for _ in range(steps):
            ops.cm_update_(cm, pred, target, C)
        ev1.record()
        torch.cuda.synchronize()
        kernels = prof.read()
    ms = ev0.elapsed_time(ev1) / steps
    assert int(cm.sum().item()) == steps * n
    gbs = 16.0 * n / (ms * 1e-3) / 1e9
    if "cm_route" in kernels and "cm_count" in kernels:
        # bucketed path: the update IS the pair (route by matrix slice, count in shared memory);
        # the roofline is taken over both launches together, against the 16 B/sample of the update
        kname = "cm_route+cm_count"
        kavg = sum(kernels[k]["total_ms"] / kernels[k]["launches"] for k in ("cm_route", "cm_count"))
    else:
        kname = max(kernels, key=lambda k: kernels[k]["total_ms"])
        kavg = kernels[kname]["total_ms"] / kernels[kname]["launches"]
    kgbs = 16.0 * n / (kavg * 1e-3) / 1e9

    from torcheval_b200.metrics.functional.classification.confusion_matrix import multiclass_confusion_matrix

    ph = torch.empty(n, dtype=torch.int64, pin_memory=True).copy_(pred)
    th = torch.empty(n, dtype=torch.int64, pin_memory=True).copy_(target)
    torch.cuda.synchronize()
    multiclass_confusion_matrix(ph, th, C)
    t0 = time.perf_counter()
    r = multiclass_confusion_matrix(ph, th, C)
    tot = int(r.sum().item())
    dt = time.perf_counter() - t0
    assert tot == n
    m = 1 << 24
    ps, ts = ph[:m].clone(), th[:m].clone()
    oracle_torch.confusion_matrix(ps[: 1 << 16], ts[: 1 << 16], C)
    t0 = time.perf_counter()
    oracle_torch.confusion_matrix(ps, ts, C)
    cdt = time.perf_counter() - t0
    return {
        "metric": "multiclass_confusion_matrix_update_GBps",
        "value": gbs,
        "unit": "GB/s",
        "ms_per_step": ms,
        "samples_per_s": n / (ms * 1e-3),
        "config": {"workload": f"MulticlassConfusionMatrix update, C={C}, N=2^{CM_LOG2_N} int64 label pairs, 1 GPU"},
        "roofline": {
            "kernel": kname,
            "bound": "hbm",
            "achieved": round(kgbs, 1),
            "peak": peak_gbs,
            "peak_source": peak_src,
            "unit": "GB/s",
            "frac": round(kgbs / peak_gbs, 4),
            "traffic": None,
            "avg_launch_ms": round(kavg, 4),
            "kernels": {k: round(v["total_ms"] / max(v["launches"], 1), 4) for k, v in kernels.items()},
        },
        "e2e": {"value": 16.0 * n / dt / 1e9, "unit": "GB/s", "h2d_bytes_per_step": 16 * n, "d2h_bytes_per_step": 8 * C * C},
        "cpu_baseline": {
            "value": 16.0 * m / cdt / 1e9,
            "unit": "GB/s",
            "samples_per_s": m / cdt,
            "cores": torch.get_num_threads(),
            "kind": "port",
            "sample": f"first 2^24 label pairs, 1 run, {cdt:.1f} s (sparse COO coalesce, oracle/oracle_torch.py)",
        },
    }


# ----------------------------------------------------------------------------------- reference arm
def run_reference(args) -> dict:
    """The reference's CPU path (port in oracle/oracle_torch.py: the same ATen op sequence) on the
    host cores; each step is a bounded sample (2^24 scores) of the same synthetic workload."""
    rank, world, _ = dist_env()
    if rank != 0:
        return {}
    from oracle import oracle_torch

    m = 1 << int(os.environ.get("TE_BENCH_REF_LOG2_N", "24"))
    g = torch.Generator().manual_seed(SEED)
    x = torch.rand(m, generator=g, dtype=torch.float32)
    y = torch.randint(0, 2, (m,), generator=g, dtype=torch.int64)
    for _ in range(args.warmup):
        oracle_torch.binary_auroc(x, y)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        v = oracle_torch.binary_auroc(x, y)
    dt = (time.perf_counter() - t0) / args.steps
    value = m / dt
    sample = f"2^{m.bit_length() - 1} scores per step (bounded sample of the 2^{LOG2_N_TOTAL} workload)"
    return {
        "impl": "reference",
        "metric": "exact_binary_auroc_scores_per_sec",
        "value": value,
        "unit": "scores/s",
        "n_gpus": args.gpus,
        "steps": args.steps,
        "warmup": args.warmup,
        "ms_per_step": dt * 1e3,
        "higher_is_better": True,
        "scaling": "strong",
        "vs_baseline": None,
        "dtype": "f64",
        "data": "synthetic",
        "config": {
            "workload": f"BinaryAUROC exact, N=2^{LOG2_N_TOTAL} float32 scores + int64 labels, rank-sharded",
            "sample": sample,
            "auroc": float(v),
        },
        "cpu_baseline": {"value": value, "unit": "scores/s", "cores": torch.get_num_threads(), "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "scores/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    line = run_reference(args) if args.impl == "reference" else run_ours(args)
    if line:
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
