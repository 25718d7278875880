#!/usr/bin/env python
"""bench.py — headline benchmark of the ranking & counting metric hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Primary workload (BASELINE.json metric, configs[3]): exact BinaryAUROC over N = 2^30 synthetic
float32 scores + int64 {0,1} labels, rank-sharded over the GPUs (strong scaling: the total is
fixed, each of N ranks holds 2^30 / N).  One "step" = one full compute of the metric over the
resident scores.  Also in the line, at 1 GPU: "secondary" (configs[1]: MulticlassConfusionMatrix
update, C = 1000, 2^28 label pairs, GB/s against the HBM roofline), "secondary_binned" (configs[2]:
10k-threshold binned counting over the same 2^30 scores) and "secondary_ovr" (configs[4] at its
stated size: one-vs-rest MulticlassAUROC, C = 1000, N = 2^24 samples, 1.7e10 scores); at N GPUs:
"secondary_binned" = binned update + counter all-reduce through toolkit.sync_and_compute,
"secondary_ovr" = the same 2^24 x 1000 matrix sample-sharded, through MulticlassAUROC +
sync_and_compute (class-sharded transpose), and "parity": the sharded results checked, outside
the timed region, against an independent single-GPU computation on the gathered keys and
against closed forms on grid-valued scores.

Prints ONE JSON line on rank 0 (see the task contract): value = device-timed whole-job
throughput with inputs resident in HBM; e2e = same metric through the public functional API
with pinned HOST tensors (H2D + D2H inside the timed region); roofline = the dominant kernel of
the step (a cursor partition pass at 1 GPU, the exchange's sender at 8), timed with CUDA events on
the launching stream during the timed steps;
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
    """DRAM bytes per launch of `kernel` (dram__bytes_read + dram__bytes_write of one
    `ncu --set full` capture of this round's kernels, profiles/r02_ncu_traffic.json), scaled from
    the record count of the capture to this run's: the kernels stream, traffic is per record."""
    p = ROOT / "profiles" / "r02_ncu_traffic.json"
    try:
        t = json.loads(p.read_text())
        return int(round(float(t["bytes_per_record"][kernel]) * n_records))
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


# ----------------------------------------------------------------------------------- host side
def gpu_numa_cpus(local: int):
    """(numa node, cpus of that node) of CUDA device `local`, from sysfs; (None, None) when the host
    does not say (single node, a VM without topology, no sysfs)."""
    try:
        pr = torch.cuda.get_device_properties(local)
        addr = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        node = int(Path(f"/sys/bus/pci/devices/{addr}/numa_node").read_text())
        if node < 0:
            return None, None
        cpus = set()
        for part in Path(f"/sys/devices/system/node/node{node}/cpulist").read_text().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        return (node, cpus) if cpus else (None, None)
    except Exception:
        return None, None


def pinned_like(x: torch.Tensor, local: int) -> torch.Tensor:
    """A pinned host copy of x whose pages sit on the NUMA node of GPU `local` (the process is
    bound to that node only while the pages are allocated: first touch places them).  What a
    NUMA-aware data loader does; TE_BENCH_NUMA=0 leaves the pages where the launcher put the
    process."""
    node, cpus = gpu_numa_cpus(local) if os.environ.get("TE_BENCH_NUMA", "1") != "0" else (None, None)
    before = os.sched_getaffinity(0)
    if cpus:
        os.sched_setaffinity(0, cpus)
    try:
        h = torch.empty(x.shape, dtype=x.dtype, pin_memory=True)
    finally:
        if cpus:
            os.sched_setaffinity(0, before)
    h.copy_(x)
    return h


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
            # no host read inside: splitters, slab exchange, counting stage and fix-up queue up on
            # the stream; out[4] (flags, 0 = valid) is checked after the timed region
            return tdist.sharded_binary_auroc_async(x, y)
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
    assert int(out[4].item()) == 0, f"flags {out[4].item()}: the timed path did not produce a valid result"
    auroc = float(out[0].item())
    launches = sum(k["launches"] for k in kernels.values())

    # roofline of the dominant kernel (the one with the largest share of the timed region), against
    # the algorithmic HBM bytes it moves per record
    bytes_per_record = {
        "cursor_pass_top_pack": 8.125, "cursor_pass_top_pack_hot": 8.125, "cursor_pass_top": 8.0, "cursor_pass_top_hot": 8.0,
        "cursor_pass_low": 8.0, "cursor_pass_low_hot": 8.0, "partition_ranges": 8.0, "partition_ranges_pack": 8.125,
        "onesweep_pass_keys": 8.0, "onesweep_pass_keys_first": 8.0, "rank_upsweep_joint": 12.125, "rank_upsweep_bits": 12.125,
        "rank_ingest_ranges": 16.0, "rank_ingest_packed": 16.0, "leaf_count": 4.0, "pack_scatter_peer": 12.0,
    }
    roof = None
    cands = [k for k in kernels if k in bytes_per_record and kernels[k]["launches"]]
    if cands:
        dom = max(cands, key=lambda k: kernels[k]["total_ms"])
        k = kernels[dom]
        avg_ms = k["total_ms"] / k["launches"]
        # records the kernel handles per launch: this rank's shard for the sender kernel, the
        # records this rank received for everything after the exchange
        recs = n_local if (world == 1 or dom == "pack_scatter_peer") else int(out[5].item())
        bpr = bytes_per_record[dom]
        achieved = bpr * recs / (avg_ms * 1e-3) / 1e9
        roof = {
            "kernel": dom,
            "bound": "hbm",
            "achieved": round(achieved, 1),
            "peak": peak_gbs,
            "peak_source": peak_src,
            "unit": "GB/s",
            "frac": round(achieved / peak_gbs, 4),
            "traffic": ncu_traffic(dom, recs),
            "avg_launch_ms": round(avg_ms, 4),
            "algorithmic_bytes_per_launch": int(bpr * recs),
            "algorithmic_bytes_per_record": bpr,
            "step_share": round(k["total_ms"] / (ms_total), 4),
        }
    # bytes the structure that ran actually moves per score: counting-stage path
    # 16 ingest + 2 passes * 8 + 4 counting = 36; full sort 16 + 4 * 8 + 4 sweep = 52
    # (32.25 when the upsweep leaves label bits instead of packed keys: 12.125 upsweep + 8.125
    # first pass + 8 second pass + 4 counting)
    step_bytes = (32.25 if ("rank_upsweep_bits" in kernels or "rank_upsweep_joint" in kernels)
                  else 36.0 if "leaf_count" in kernels else 52.0) * n_local
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
    parity = None
    if world == 1:
        binned = guarded(binned_bench, args, x, y, peak_gbs, peak_src)
    else:
        binned = guarded(binned_sync_bench, args, x, y, world, device)
        parity = parity_checks(x, y, out, world, rank, device)  # raises: a wrong sharded result fails the run
    if world == 1:
        xh = pinned_like(x, local)
        yh = pinned_like(y, local)
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
        secondary = guarded(confusion_matrix_bench, args, device, peak_gbs, peak_src)
        ovr = guarded(one_vs_rest_bench, args, device, peak_gbs, peak_src)

    if world > 1:
        # end to end at N GPUs through the class API + toolkit: every rank feeds its pinned HOST
        # shard to BinaryAUROC.update (H2D) and all ranks call sync_and_compute (sample sort over NCCL)
        import torch.distributed as dist

        from torcheval_b200.metrics import BinaryAUROC
        from torcheval_b200.metrics.toolkit import sync_and_compute

        xh = pinned_like(x, local)
        yh = pinned_like(y, local)
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
        # where the step goes (one extra, untimed step with a device sync after each phase)
        barrier(world)
        t0 = time.perf_counter()
        m = BinaryAUROC(device=device)
        m.update(xh, yh)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        v = sync_and_compute(m)
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        float(v.item())
        t3 = time.perf_counter()
        phases = {
            "update_h2d_ms": max_over_ranks((t1 - t0) * 1e3, world, device),
            "sync_and_compute_ms": max_over_ranks((t2 - t1) * 1e3, world, device),
            "result_d2h_ms": max_over_ranks((t3 - t2) * 1e3, world, device),
        }
        e2e = {
            "value": n_total / dt,
            "unit": "scores/s",
            "h2d_bytes_per_step": 12 * n_local,
            "d2h_bytes_per_step": 8,
            "ms_per_step": dt * 1e3,
            "steps": e2e_steps,
            "phases_ms": {k: round(v_, 3) for k, v_ in phases.items()},
            "note": "per-rank bytes; class API update() + toolkit.sync_and_compute(); phases_ms: max over ranks, "
                    "one untimed step with a device sync after each phase",
        }

    if world > 1:
        del xh, yh
        torch.cuda.empty_cache()
        ovr = guarded(one_vs_rest_sharded_bench, args, world, rank, device)

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
    if parity:
        line["parity"] = parity
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


def guarded(fn, *a):
    """A secondary line must never cost the primary one: failures are reported in place."""
    try:
        return fn(*a)
    except Exception as e:  # noqa: BLE001
        torch.cuda.empty_cache()
        return {"error": f"{type(e).__name__}: {e}"[:300]}


OVR_LOG2_N = int(os.environ.get("TE_BENCH_OVR_LOG2_N", "24"))
OVR_CLASSES = 1000


def _ovr_matrix(rows: int, C: int, device, seed: int):
    """(rows, C) uniform float32 scores generated in row chunks into one allocation (67 GB at
    2^24 x 1000: the input of configs[4] itself; nothing else of that size is ever held)."""
    g = torch.Generator(device=device).manual_seed(seed)
    x = torch.empty((rows, C), device=device, dtype=torch.float32)
    step = 1 << 18
    for r0 in range(0, rows, step):
        torch.rand((min(step, rows - r0), C), device=device, generator=g, dtype=torch.float32, out=x[r0 : r0 + step])
    y = torch.randint(0, C, (rows,), device=device, generator=g, dtype=torch.int64)
    return x, y


def one_vs_rest_bench(args, device, peak_gbs, peak_src) -> dict:
    """configs[4] at its stated size on one GPU: MulticlassAUROC one-vs-rest, C = 1000 classes,
    N = 2^24 samples = 1.68e10 scores (67 GB of float32, resident; the segmented sort works through
    it eight classes at a time with ~1 GB of scratch)."""
    from torcheval_b200 import ops

    n, C = 1 << OVR_LOG2_N, OVR_CLASSES
    x, y = _ovr_matrix(n, C, device, SEED + 5)
    out = ops.ovr_rank(x, y, False, 0)
    torch.cuda.synchronize()
    steps = 2
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with KernelProfile() as prof:
        ev0.record()
        for _ in range(steps):
            out = ops.ovr_rank(x, y, False, 0)
        ev1.record()
        torch.cuda.synchronize()
        kernels = prof.read()
    ms = ev0.elapsed_time(ev1) / steps
    # closed form per class is not available for uniform scores; the macro mean of 1000 independent
    # AUROCs of random labels must sit at 0.5 +- a few 1e-5, and every class must have seen n samples
    counts_ok = bool(((out[:, 1] + out[:, 2]) == n).all().item()) and bool((out[:, 4] == 0).all().item())
    macro = float(out[:, 0].mean().item())
    assert counts_ok and abs(macro - 0.5) < 1e-3, (counts_ok, macro)
    # ingest reads N*C*4 and writes N*C*4 packed keys; then either P = 4 passes
    # of 8 B + 4 B sweep, or (counting-stage path, long columns) 4 B upsweep + 2 passes of 8 B + 4 B counting
    counting = "leaf_count" in kernels
    bytes_step = float(n) * C * ((8 + 4 + 2 * 8 + 4) if counting else (8 + 4 * 8 + 4))
    gbs = bytes_step / (ms * 1e-3) / 1e9
    dom = max(kernels, key=lambda k: kernels[k]["total_ms"]) if kernels else None
    return {
        "metric": "multiclass_auroc_one_vs_rest_scores_per_sec",
        "value": n * C / (ms * 1e-3),
        "unit": "scores/s",
        "ms_per_step": ms,
        "config": {"workload": f"MulticlassAUROC one-vs-rest, C={C}, N=2^{OVR_LOG2_N} samples, float32 (N, C) scores resident, 1 GPU",
                   "macro_auroc": macro},
        "roofline": {
            "kernel": ("whole step (ovr_ingest + upsweep + 2 cursor passes + counting stage, 8 classes per launch)" if counting
                       else "whole step (ovr_ingest + 4 radix passes + sweep, 8 classes per launch)"),
            "bound": "hbm",
            "achieved": round(gbs, 1),
            "peak": peak_gbs,
            "peak_source": peak_src,
            "unit": "GB/s",
            "frac": round(gbs / peak_gbs, 4),
            "traffic": None,
            "algorithmic_bytes_per_step": bytes_step,
            "dominant_kernel": dom,
            "kernels_ms_per_step": {k: round(v["total_ms"] / steps, 3) for k, v in kernels.items()},
        },
    }


def one_vs_rest_sharded_bench(args, world, rank, device) -> dict:
    """configs[4] at N GPUs: the 2^24 x 1000 matrix sample-sharded (2^24 / N rows per rank), through
    the class API: MulticlassAUROC.update + toolkit.sync_and_compute = class-sharded transpose
    (one all_to_all), local segmented sorts of C / N classes, all_gather of the C values."""
    from torcheval_b200.metrics import MulticlassAUROC
    from torcheval_b200.metrics.toolkit import sync_and_compute

    n, C = 1 << OVR_LOG2_N, OVR_CLASSES
    rows = n // world
    x, y = _ovr_matrix(rows, C, device, SEED + 5 + 97 * rank)

    def step():
        m = MulticlassAUROC(num_classes=C, average=None, device=device)
        m.update(x, y)
        return sync_and_compute(m)

    out = step()
    barrier(world)
    steps = 2
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(steps):
        out = step()
    ev1.record()
    barrier(world)
    ms = max_over_ranks(ev0.elapsed_time(ev1), world, device) / steps
    macro = float(out.double().mean().item())
    assert out.numel() == C and abs(macro - 0.5) < 1e-3, macro
    return {
        "metric": "multiclass_auroc_one_vs_rest_scores_per_sec",
        "value": n * C / (ms * 1e-3),
        "unit": "scores/s",
        "ms_per_step": ms,
        "scaling": "strong",
        "config": {"workload": f"MulticlassAUROC one-vs-rest, C={C}, N=2^{OVR_LOG2_N} samples sample-sharded over {world} GPUs, "
                               "update + sync_and_compute (class-sharded all_to_all transpose)",
                   "macro_auroc": macro},
    }


def parity_checks(x, y, out, world, rank, device) -> dict:
    """Outside the timed region, at N > 1: the sharded results against computations that share no
    code with the sample sort / all-reduce paths.  Any mismatch raises (the run fails).
      1. gathered keys: every rank's packed keys are all-gathered; rank 0 computes AUROC over all
         of them with the single-GPU kernel path and requires the sharded 2U, P, N digit for digit;
      2. closed forms: scores on a 2^12-level grid, sharded by rank; the (level, label) histogram
         (torch.bincount + all_reduce) gives 2U, AUROC and AUPRC in closed form, ties included;
      3. binned counters: BinaryBinnedAUPRC.update + sync_and_compute against the 10k-threshold
         histogram of the gathered scores computed on one GPU;
      4. class-sharded one-vs-rest (MulticlassAUROC, 2^16 x 64) against the single-GPU kernel on
         the gathered matrix."""
    import numpy as np
    import torch.distributed as dist

    from torcheval_b200 import distributed as tdist
    from torcheval_b200 import ops
    from torcheval_b200.metrics import BinaryBinnedAUPRC, MulticlassAUROC
    from torcheval_b200.metrics.toolkit import sync_and_compute

    checked = []
    n_local = x.numel()
    sh = out.cpu().numpy()  # [auroc, P, N, 2U, flags, ...] of the timed sharded run

    # 1 -----------------------------------------------------------------------------------------
    keys = ops.rank_pack(x, y, ())[0]
    allk = torch.empty(n_local * world, dtype=keys.dtype, device=device)
    dist.all_gather_into_tensor(allk, keys)
    del keys
    ok = torch.ones(1, dtype=torch.int32, device=device)
    if rank == 0:
        one = ops.binary_auroc_from_keys(allk).cpu().numpy()
        same = one[1] == sh[1] and one[2] == sh[2] and one[3] == sh[3] and one[4] == 0
        ok[0] = 1 if same else 0
        detail = {"single_gpu": [one[1], one[2], one[3]], "sharded": [sh[1], sh[2], sh[3]]}
    dist.broadcast(ok, 0)
    assert int(ok.item()) == 1, f"sharded AUROC differs from the single-GPU result on the gathered keys: {detail if rank == 0 else ''}"
    checked.append(f"auroc_2U_P_N_vs_single_gpu_on_gathered_keys(n=2^{(n_local * world).bit_length() - 1})")

    # 3 (uses the gathered keys) ----------------------------------------------------------------
    T = 10_000
    m = BinaryBinnedAUPRC(threshold=T, device=device)
    m.update(x, y)
    sync_and_compute(m)
    from torcheval_b200.metrics.toolkit import get_synced_metric

    synced = get_synced_metric(m)
    if rank == 0:
        xs = (allk >> 1).view(torch.float32)  # key = (float bits << 1) | label, bits < 2^31 here
        ys = (allk & 1).to(torch.int64)
        thr = torch.linspace(0, 1, T, device=device)
        hist = torch.zeros((1, T + 1, 2), dtype=torch.int64, device=device)
        ops.binned_hist_update_(hist, xs, ys, thr)
        tp, fp, fn = ops.binned_counts(hist, T)
        same = bool((synced.num_tp.to(torch.int64) == tp).all()) and bool((synced.num_fp.to(torch.int64) == fp).all())
        ok[0] = 1 if same else 0
        del xs, ys
    dist.broadcast(ok, 0)
    assert int(ok.item()) == 1, "synced binned counters differ from the single-GPU histogram of the gathered scores"
    checked.append(f"binned_counts_T{T}_vs_single_gpu")
    del allk
    torch.cuda.empty_cache()

    # 2 -----------------------------------------------------------------------------------------
    G = 1 << 12
    g = torch.Generator(device=device).manual_seed(SEED + 1000 + rank)
    lev = torch.randint(0, G, (n_local,), device=device, generator=g)
    yy = (torch.rand(n_local, device=device, generator=g) < 0.25).long()
    xx = lev.float() / G
    neg = torch.bincount(lev[yy == 0], minlength=G)
    pos = torch.bincount(lev[yy == 1], minlength=G)
    del lev
    dist.all_reduce(neg)
    dist.all_reduce(pos)
    neg, pos = neg.cpu().numpy().astype(np.int64), pos.cpu().numpy().astype(np.int64)
    below = np.concatenate([[0], np.cumsum(neg)[:-1]])
    u2 = int(np.sum(pos.astype(object) * (2 * below.astype(object) + neg.astype(object))))
    a = tdist.sharded_binary_auroc(xx, yy).cpu().numpy()
    assert a[4] == 0 and a[1] == pos.sum() and a[2] == neg.sum() and (int(a[6]) << 32) | int(a[7]) == u2, (a, u2)
    want = u2 / (2.0 * float(pos.sum()) * float(neg.sum()))
    assert abs(a[0] - want) <= 1e-12 * want
    ap = tdist.sharded_binary_auprc(xx, yy).cpu().numpy()
    tp_ge = np.cumsum(pos[::-1])[::-1].astype(np.float64)
    all_ge = np.cumsum((pos + neg)[::-1])[::-1].astype(np.float64)
    want_ap = float(np.sum(pos * (tp_ge / np.maximum(all_ge, 1))) / pos.sum())
    assert abs(ap[0] - want_ap) <= 1e-9 * want_ap, (ap[0], want_ap)
    checked.append("auroc_exact_2U_and_auprc_vs_closed_form_on_grid_scores")
    del xx, yy

    # 4 -----------------------------------------------------------------------------------------
    rows, C = (1 << 16) // world, 64
    g = torch.Generator(device=device).manual_seed(SEED + 2000 + rank)
    xm = torch.rand((rows, C), device=device, generator=g)
    ym = torch.randint(0, C, (rows,), device=device, generator=g)
    mm = MulticlassAUROC(num_classes=C, average=None, device=device)
    mm.update(xm, ym)
    got = sync_and_compute(mm).double()
    allx = torch.empty((rows * world, C), device=device)
    ally = torch.empty(rows * world, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(allx, xm)
    dist.all_gather_into_tensor(ally, ym)
    ref = ops.ovr_rank(allx, ally, False, 0)[:, 0]
    assert bool((torch.abs(got - ref) <= 1e-6 * ref).all()), "class-sharded one-vs-rest differs from the single-GPU result"
    checked.append("multiclass_auroc_class_sharded_vs_single_gpu")
    return {"checked": checked, "ok": True}


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
    """The reference's CPU path — its own ATen op sequence, restated op for op in
    oracle/oracle_torch.py because /root/reference does not exist on the GPU box — on all host
    cores; each step is a bounded sample (2^25 scores, the same sample as the in-line cpu_baseline
    of the GPU arm) of the same synthetic workload.  Also timed, once each, for BASELINE.md: the
    confusion-matrix and binned-AUPRC reference paths at 2^24."""
    rank, world, _ = dist_env()
    if rank != 0:
        return {}
    # torchrun exports OMP_NUM_THREADS=1; the reference arm is the CPU path at full strength
    ncpu = os.cpu_count() or 1
    torch.set_num_threads(ncpu)
    from oracle import oracle_torch

    m = 1 << int(os.environ.get("TE_BENCH_REF_LOG2_N", "25"))
    g = torch.Generator().manual_seed(SEED)
    x = torch.rand(m, generator=g, dtype=torch.float32)
    y = torch.randint(0, 2, (m,), generator=g, dtype=torch.int64)
    for _ in range(min(args.warmup, 1)):
        oracle_torch.binary_auroc(x, y)
    steps = max(1, min(args.steps, 3))  # ~4 s per step: keeps the run within a few minutes
    t0 = time.perf_counter()
    for _ in range(steps):
        v = oracle_torch.binary_auroc(x, y)
    dt = (time.perf_counter() - t0) / steps
    value = m / dt
    sample = f"2^{m.bit_length() - 1} scores per step (bounded sample of the 2^{LOG2_N_TOTAL} workload), {steps} timed steps"
    extra = {}
    try:
        k = 1 << 24
        for C in (10, 1000):
            p_, t_ = torch.randint(0, C, (k,), generator=g), torch.randint(0, C, (k,), generator=g)
            t0 = time.perf_counter()
            oracle_torch.confusion_matrix(p_, t_, C)
            extra[f"confusion_matrix_C{C}_2p24_samples_per_s"] = k / (time.perf_counter() - t0)
        thr = torch.linspace(0, 1, 10_000)
        t0 = time.perf_counter()
        oracle_torch.binned_counts(x[:k], y[:k], thr)
        extra["binned_counts_T10000_2p24_scores_per_s"] = k / (time.perf_counter() - t0)
    except Exception as e:  # noqa: BLE001
        extra["error"] = str(e)[:200]
    return {
        "impl": "reference",
        "metric": "exact_binary_auroc_scores_per_sec",
        "value": value,
        "unit": "scores/s",
        "n_gpus": args.gpus,
        "steps": steps,
        "warmup": min(args.warmup, 1),
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
        "cpu_baseline": {"value": value, "unit": "scores/s", "cores": torch.get_num_threads(), "host_cpus": ncpu,
                         "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "scores/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "other_reference_paths": extra,
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
