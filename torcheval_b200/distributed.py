"""Multi-GPU paths (one process per GPU, torch.distributed / NCCL over NVLink).

Exact binary AUROC over rank-sharded scores as a distributed sample sort:

  1. pack      every rank turns (score, label) into 32-bit order-preserving keys      [kernel]
  2. splitters s regular samples per rank -> all_gather -> W-1 quantile splitters; splitters
               are score values (label bit cleared), so equal scores go to ONE rank and no tie
               group ever straddles a rank boundary
  3. partition one radix-style pass groups the local keys by destination rank         [kernel]
  4. exchange  all_gather of the W bucket sizes, then one all_to_all_single of the keys
  5. local     sort + fused scan/integrate of the received range                      [kernels]
  6. fix-up    all_gather of (P_r, N_r, 2U_r) — three integers per rank:
               2U = sum_r ( 2U_r + 2 * P_r * sum_{q<r} N_q )   (rank order = score order)

The reference instead pickles every rank's whole sample set to every rank and re-sorts all N
samples everywhere (torcheval/metrics/toolkit.py:371-391, classification/auroc.py:131-136).

Every kernel-backed step is a call into ``torcheval_b200.ops``; the CPU/gloo tests of the host
protocol substitute those functions (tests/cpu_backend.py), nothing in this module computes on
the CPU.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

from . import _native as N_
from . import ops

_WIDE = N_.TE_RANK_FLAG_WIDE_SCORE  # informational flag of the pack kernel: not a reason to leave the packed path
MAX_SAMPLE_SORT_RANKS = 16  # kMaxSplitters + 1 of csrc/rank_sort.cuh
SAMPLES_PER_RANK = 4096
# All samples are sorted by ONE thread block (te_select_splitters, bitonic in shared memory:
# 0.05 ms for 8192 keys, 0.3 ms for 32768), so the pool is capped; 8192 samples put the quantile
# error of a splitter at ~0.4 % of the total, i.e. ~3 % of a rank's share at 8 ranks.
SAMPLE_POOL = 8192


class _PhaseTimer:
    """TE_B200_DIST_TIMING=1: print the wall time of every phase (with device syncs; debugging only)."""

    def __init__(self, dev):
        import os
        import time

        self.level = os.environ.get("TE_B200_DIST_TIMING", "0")
        self.on = self.level in ("1", "2")  # 2: every rank prints
        self.dev, self.t, self.time, self.rows = dev, None, time, []
        if self.on:
            torch.cuda.synchronize(dev) if dev.type == "cuda" else None
            self.t = time.perf_counter()

    def mark(self, name):
        if not self.on:
            return
        torch.cuda.synchronize(self.dev) if self.dev.type == "cuda" else None
        now = self.time.perf_counter()
        self.rows.append((name, (now - self.t) * 1e3))
        self.t = now

    def done(self, rank):
        if self.on and (rank == 0 or self.level == "2"):
            print(f"[sharded r{rank}] " + "  ".join(f"{n}={ms:.2f}ms" for n, ms in self.rows), flush=True)


class _PeerInbox:
    """Peer-mapped receive buffers (torch symmetric memory, CUDA IPC under NVLink/NVSwitch) for
    the fused partition + exchange: every rank's partition kernel stores bucket b directly into
    rank b's inbox, so there is no separate all-to-all and no local copy of the partitioned keys.
    One buffer per (process group, device), grown on demand; allocation is collective."""

    _cache: dict = {}
    _broken = False

    @classmethod
    def get(cls, group, dev: torch.device, capacity: int):
        import os

        if cls._broken or dev.type != "cuda" or os.environ.get("TE_B200_NO_PEER") == "1":
            return None  # (TE_B200_NO_PEER=1: exercise the NCCL all_to_all_single exchange)
        key = (id(group) if group is not None else 0, dev.index)
        ent = cls._cache.get(key)
        if ent is not None and ent[0] >= capacity:
            return ent
        try:
            import torch.distributed._symmetric_memory as symm_mem

            cap = ((int(capacity) + (1 << 20) - 1) >> 20) << 20
            buf = symm_mem.empty(cap, dtype=torch.int32, device=dev)
            hdl = symm_mem.rendezvous(buf, group if group is not None else dist.group.WORLD)
            ent = (cap, buf, hdl, [int(p) for p in hdl.buffer_ptrs])
            cls._cache[key] = ent
            return ent
        except Exception:  # noqa: BLE001 — no peer access on this system: use the NCCL exchange
            cls._broken = True
            return None


class _SlabInbox:
    """Inbox of the slab exchange (csrc/rank_exchange.cuh): symmetric memory holding a 1 KiB header
    and W slabs of `cap` keys, slab q written by rank q.  One per (process group, device), grown
    collectively: get() with grow=False never communicates, so the steady state of the sharded
    AUROC has no host round trip; the first call of a group and a call after an overflow do one
    MAX all-reduce of the shard sizes (host read) to agree on the capacity."""

    _cache: dict = {}
    _broken = False
    SLACK = 2.0  # slab capacity = SLACK x an even split of the largest shard

    @classmethod
    def get(cls, group, dev: torch.device, n_local: int, grow: bool):
        import os

        if cls._broken or dev.type != "cuda" or os.environ.get("TE_B200_NO_PEER") == "1" or os.environ.get("TE_B200_NO_SLABS") == "1":
            return None
        key = (id(group) if group is not None else 0, dev.index)
        ent = cls._cache.get(key)
        if ent is not None and not grow:
            return ent
        lib = N_.load()
        world = dist.get_world_size(group)
        # collective from here on: every rank of the group is in this branch together (first call /
        # all ranks saw the overflow flag of the previous exchange)
        t = torch.tensor([n_local], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        n_max = int(t.item())
        slack = float(os.environ.get("TE_B200_SLAB_SLACK", cls.SLACK))  # (tests shrink it to force an overflow)
        cap = int(lib.te_slab_capacity(int(slack / 1.3 * n_max) + 1, world))
        if ent is not None and ent["cap"] >= cap:
            cap = 2 * ent["cap"]  # the split was more skewed than the slack: double it
        if world == 1 and os.environ.get("TE_B200_SLAB_PLAIN") == "1":  # (measurement aid: ordinary device memory)
            buf = torch.empty(int(lib.te_slab_inbox_bytes(world, cap)) // 4, dtype=torch.int32, device=dev)

            class _NoBarrier:
                def barrier(self):
                    pass

            ent = {"cap": cap, "buf": buf, "hdl": _NoBarrier(), "ptrs": [buf.data_ptr()],
                   "work": torch.zeros(32, dtype=torch.int32, device=dev)}
            cls._cache[key] = ent
            return ent
        ok = 1
        buf = None
        try:
            import torch.distributed._symmetric_memory as symm_mem

            nbytes = int(lib.te_slab_inbox_bytes(world, cap))
            buf = symm_mem.empty(nbytes // 4, dtype=torch.int32, device=dev)
        except Exception:  # noqa: BLE001
            ok = 0
        agree = torch.tensor([ok], dtype=torch.int32, device=dev)
        dist.all_reduce(agree, op=dist.ReduceOp.MIN, group=group)  # one rank cannot allocate: nobody uses slabs
        if int(agree.item()) == 0:
            cls._broken = True
            return None
        hdl = symm_mem.rendezvous(buf, group if group is not None else dist.group.WORLD)
        ent = {"cap": cap, "buf": buf, "hdl": hdl, "ptrs": [int(p) for p in hdl.buffer_ptrs],
               "work": torch.zeros(32, dtype=torch.int32, device=dev)}
        cls._cache[key] = ent
        return ent


def _gather_rows(t: torch.Tensor, world: int, group) -> torch.Tensor:
    out = torch.empty((world,) + tuple(t.shape), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out.view(-1), t.contiguous().view(-1), group=group) if t.is_cuda else dist.all_gather(
        list(out.unbind(0)), t.contiguous(), group=group
    )
    return out




_SAMPLE_INDEX_CACHE: dict = {}


def _sample_index(n: int, s: int, dev: torch.device) -> torch.Tensor:
    key = (n, s, str(dev))
    idx = _SAMPLE_INDEX_CACHE.get(key)
    if idx is None:
        if len(_SAMPLE_INDEX_CACHE) > 64:
            _SAMPLE_INDEX_CACHE.clear()
        idx = torch.div(torch.arange(s, device=dev, dtype=torch.int64) * n, s, rounding_mode="floor")
        _SAMPLE_INDEX_CACHE[key] = idx
    return idx


def device_splitters(score: torch.Tensor, world: int, group):
    """(int32[W-1] splitter bit patterns for the pack kernel, int64[W-1] the same values) chosen
    from regular samples of every rank's scores; never leaves the device, identical on every rank.
    Splitters are key values with the label bit cleared.  On CUDA: one gather of the sampled bit
    patterns, one all_gather, ONE kernel (shared-memory bitonic sort + quantile pick); the torch
    op chain it replaces cost 0.6 ms of launch gaps per call (profiles/r01_dist_trace_4gpu.txt)."""
    dev = score.device
    n = score.numel()
    s = min(SAMPLES_PER_RANK, SAMPLE_POOL // world)
    if n > 0:
        smp = score.contiguous().view(torch.int32)[_sample_index(n, s, dev)]
    else:
        smp = torch.full((s,), -1, dtype=torch.int32, device=dev)
    pool = _gather_rows(smp, world, group)  # negative bit patterns (negative scores, no data) are not samples
    return ops.select_splitters(pool, world)


def _exchange_keys(score: torch.Tensor, target: torch.Tensor, group, tm: _PhaseTimer):
    """Steps 1-4: returns (inbox, flags, hist).  inbox = the packed keys of the score range this
    rank owns (rank order = ascending score order), None when flags != 0 (negative scores / soft
    labels on some rank: the packed representation does not apply; every rank sees the same
    flags).  hist = int32[1024], the radix-digit histograms of the inbox: every sender histograms
    what it sends while packing, and the rows are summed across the senders (one reduce-scatter), so
    the receiver never reads its inbox just to count digits.
    ONE host read: bucket sizes, flags and the splitters come back together."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    dev = score.device

    sp32, sp_dev = device_splitters(score, world, group)
    tm.mark("splitters")
    # --- pack (+ bucket sizes) with the splitters still on the device --------------------------
    keys, info, counts, send_hist = ops.rank_pack(score, target, sp32, want_hist=True)
    my_hist = _sum_rows_to_owner(send_hist, world, rank, group)
    tm.mark("pack")
    rows = _gather_rows(torch.cat([counts, info[1:2], sp_dev]), world, group).tolist()
    flags = 0
    for r in rows:
        flags |= int(r[world])
    if flags & ~_WIDE:
        return None, flags & ~_WIDE, None
    splitters = [int(v) for v in rows[rank][world + 1 :]]
    send = [int(c) for c in rows[rank][:world]]
    recv = [int(rows[q][rank]) for q in range(world)]

    tm.mark("sizes")
    # --- partition fused with the exchange (P2P stores into peer inboxes), or partition + NCCL ----
    n_all = [sum(int(v) for v in rows[q][:world]) for q in range(world)]
    recv_all = [sum(int(rows[q][b]) for q in range(world)) for b in range(world)]
    peer = None
    if dev.type == "cuda":
        peer = _PeerInbox.get(group, dev, max(int(1.3 * max(n_all)) + 65536, 1 << 20))
        if peer is not None and max(recv_all) > peer[0]:
            peer = None  # a skewed split does not fit the inboxes: every rank sees that and falls back
    if peer is not None:
        _, buf, hdl, ptrs = peer
        dest, excl = [], 0
        for b in range(world):
            slot = sum(int(rows[q][b]) for q in range(rank))  # my keys follow those of lower ranks
            dest.append(ptrs[b] + 4 * (slot - excl))
            excl += send[b]
        ops.partition_scatter_u32(keys, splitters, counts, dest)
        hdl.barrier()  # every rank's stores have landed before anybody reads its inbox
        inbox = buf[: recv_all[rank]]
        tm.mark("partition+p2p")
    else:
        part, _ = ops.partition_u32(keys, splitters, counts)
        tm.mark("partition")
        inbox = torch.empty(sum(recv), dtype=keys.dtype, device=dev)
        dist.all_to_all_single(inbox, part, recv, send, group=group)
        tm.mark("all_to_all")
    return inbox, flags & _WIDE, my_hist


def _sum_rows_to_owner(rows: torch.Tensor, world: int, rank: int, group) -> torch.Tensor:
    """rows: (W, k) on every rank; returns sum over ranks of row[rank] — a reduce-scatter (NCCL),
    or an all-reduce + slice where the backend has no reduce-scatter (gloo)."""
    rows = rows.contiguous()
    if rows.is_cuda:
        out = torch.empty(rows.shape[1], dtype=rows.dtype, device=rows.device)
        dist.reduce_scatter_tensor(out, rows.view(-1), op=dist.ReduceOp.SUM, group=group)
        return out
    total = rows.clone()
    dist.all_reduce(total, op=dist.ReduceOp.SUM, group=group)
    return total[rank].contiguous()


def _flags_only(flags: int, dev) -> torch.Tensor:
    out = torch.zeros(8, dtype=torch.float64, device=dev)
    out[4] = float(flags)
    return out


SLAB_OVERFLOW = 0x100  # out[4] bit of the slab exchange: a slab was too small, the result is void


def _sharded_binary_auroc_slabs(score, target, group, grow: bool = False):
    """The sample sort with the slab exchange: float64[8] as sharded_binary_auroc, or None when
    slabs are not available (no peer access, CPU tensors).  No host read anywhere: splitters,
    scatter into the peers' slabs, barrier, counting-stage path over the slabs with the fill counts
    taken from device memory, all_gather of the W result rows, exact-integer fix-up kernel."""
    dev = score.device
    if dev.type != "cuda" or target.dtype not in ops._DTYPE_CODE:
        return None
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    if world > MAX_SAMPLE_SORT_RANKS:
        return None
    box = _SlabInbox.get(group, dev, score.numel(), grow)
    if box is None:
        return None
    tm = _PhaseTimer(dev)
    sp32, _ = device_splitters(score, world, group)
    tm.mark("splitters")
    ops.rank_pack_scatter(score, target, sp32 if world > 1 else None, world, rank, box["ptrs"], box["cap"], box["work"])
    box["hdl"].barrier()  # every sender's keys, counts and flags have landed
    tm.mark("pack+scatter")
    loc = ops.binary_auroc_from_slabs(box["buf"], world, box["cap"])
    tm.mark("local")
    rows = _gather_rows(loc, world, group)
    out = ops.combine_auroc_stats(rows, 0)
    out[4] = loc[4]  # flags of all senders (identical on every rank: each header holds all of them)
    out[5] = loc[7]  # records this rank received
    tm.mark("fixup")
    tm.done(rank)
    return out


def _sharded_binary_auroc(score, target, group):
    """(float64[8] result, flags as a host int)."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    dev = score.device
    out = _sharded_binary_auroc_slabs(score, target, group)
    if out is not None:
        flags = int(out[4].item())  # the one host read, at the very end
        if flags & SLAB_OVERFLOW:  # every rank sees it: grow the slabs together, then the count-first protocol
            _SlabInbox.get(group, dev, score.numel(), grow=True)
        elif flags & _WIDE:
            pass  # scores >= 2.0: the count-first protocol below sorts all 32 key bits
        else:
            return out, flags
    tm = _PhaseTimer(dev)
    inbox, flags, hist = _exchange_keys(score, target, group, tm)
    if flags & ~_WIDE:
        return _flags_only(flags, dev), flags
    # --- local ordering + integrate, then the fix-up of three integers per rank, on the device ---
    # (scores >= 2.0 somewhere: their keys need all 32 bits, which only the full sort handles; every
    # rank takes the same branch, the flag was part of the one host read)
    loc = ops.binary_auroc_from_keys(inbox, hist, N_.TE_RANK_SORT if flags & _WIDE else N_.TE_RANK_AUTO)
    tm.mark("local_sort_sweep")
    rows = _gather_rows(loc, world, group)  # (W, 8) float64
    tm.mark("fix:gather")
    out = ops.combine_auroc_stats(rows, inbox.numel())  # one kernel, exact 64-bit integers
    tm.mark(f"fixup(n_recv={inbox.numel()})")
    tm.done(rank)
    return out, 0


def sharded_binary_auroc(
    score: torch.Tensor, target: torch.Tensor, group: Optional[dist.ProcessGroup] = None
) -> torch.Tensor:
    """AUROC of the union of all ranks' (score, target) shards; float64[8] on score.device:
    [auroc, P, N, 2U, flags, n_received_here, 2U >> 32, 2U & 0xffffffff].  flags != 0 (negative scores / soft labels
    somewhere) means the packed representation does not apply and [0] is not valid.
    One host read in total (bucket sizes); everything else stays on the device."""
    return _sharded_binary_auroc(score, target, group)[0]


def sharded_binary_auroc_async(
    score: torch.Tensor, target: torch.Tensor, group: Optional[dist.ProcessGroup] = None
) -> torch.Tensor:
    """sharded_binary_auroc without ANY host read: the caller looks at out[4] itself — 0 = valid;
    SLAB_OVERFLOW / TE_RANK_FLAG_* set = call sharded_binary_auroc instead, which falls back.
    Successive calls queue up on the stream (what a training loop that only logs the metric every
    few steps wants).  Falls back to sharded_binary_auroc where slabs are not available."""
    out = _sharded_binary_auroc_slabs(score, target, group)
    return out if out is not None else _sharded_binary_auroc(score, target, group)[0]


def _sharded_binary_auprc(score, target, group):
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    dev = score.device
    tm = _PhaseTimer(dev)
    inbox, flags, hist = _exchange_keys(score, target, group, tm)
    if flags & ~_WIDE:  # (wide keys are fine here: this path sorts all 32 bits)
        return _flags_only(flags, dev), flags
    loc, state = ops.binary_auprc_from_keys_begin(inbox, hist)
    tm.mark("local_sort_count")
    pn = _gather_rows(loc[1:3].clone(), world, group)  # (W, 2) float64 {P_r, N_r}, exact integers
    above = pn[rank + 1 :].sum(0)                      # what the ranks with higher scores hold
    above = torch.stack([above[0], above[0] + above[1]])
    part = ops.binary_auprc_from_keys_finish(state, above)
    tm.mark("integrate")
    sums = _gather_rows(part[3:4].clone(), world, group).reshape(-1)
    total = sums.sum()
    P, Nn = pn[:, 0].sum(), pn[:, 1].sum()
    ap = torch.where(P == 0, torch.full_like(total, -0.0), total / P.clamp(min=1.0))
    out = torch.zeros(8, dtype=torch.float64, device=dev)
    out[:4] = torch.stack([ap, P, Nn, total])
    out[5] = float(inbox.numel())
    tm.mark("fixup")
    tm.done(rank)
    return out, 0


def sharded_binary_auprc(
    score: torch.Tensor, target: torch.Tensor, group: Optional[dist.ProcessGroup] = None
) -> torch.Tensor:
    """Average precision (step integral, functional/classification/auprc.py:239-251) of the union
    of all ranks' shards; float64[8]: [ap, P, N, sum, flags, n_received_here, 0, 0].
    Precision is a ratio of prefix counts, so after the exchange the work is two-phase:
    sort + count locally -> all_gather (P_r, N_r) -> integrate with what the ranks above hold ->
    all_gather the per-rank sums; both gathers feed device code directly, no host read."""
    return _sharded_binary_auprc(score, target, group)[0]


def _column_bounds(num_columns: int, world: int):
    return [(q * num_columns) // world for q in range(world + 1)]


def class_sharded_one_vs_rest(
    input: torch.Tensor,
    target: torch.Tensor,
    *,
    auprc: bool,
    multilabel: bool,
    group: Optional[dist.ProcessGroup] = None,
) -> torch.Tensor:
    """Per-class (per-label) one-vs-rest AUROC / average precision over sample-sharded
    (n_r, C) scores: float64 (C,), the same on every rank.

    Classes are independent, so instead of gathering all N x C scores everywhere (the reference:
    toolkit.py:371-391 then auroc.py:206-235 / auprc.py:279-347 on the concatenation) the ranks
    transpose the sharding with ONE all-to-all: rank q ends up with all N rows of its C/W columns,
    ranks them locally with the batched one-vs-rest kernels, and the C values are all-gathered.
    No splitters, perfectly balanced, and no rank holds more than N*C/W scores."""
    from . import _ranking

    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    dev = input.device
    n_r, C = int(input.shape[0]), int(input.shape[1])
    bounds = _column_bounds(C, world)
    width = bounds[rank + 1] - bounds[rank]
    ns = [int(v) for v in _gather_rows(torch.tensor([n_r], dtype=torch.int64, device=dev), world, group).reshape(-1).tolist()]
    n_all = sum(ns)

    def transpose_sharding(t: torch.Tensor) -> torch.Tensor:
        send = torch.cat([t[:, bounds[q] : bounds[q + 1]].reshape(-1) for q in range(world)])
        send_sizes = [n_r * (bounds[q + 1] - bounds[q]) for q in range(world)]
        recv_sizes = [ns[q] * width for q in range(world)]
        recv = torch.empty(sum(recv_sizes), dtype=t.dtype, device=dev)
        dist.all_to_all_single(recv, send, recv_sizes, send_sizes, group=group)
        return recv.view(n_all, width)  # blocks arrive in rank order = row order of the union

    x_loc = transpose_sharding(input)
    if multilabel:
        # labels cross the links as one byte each (only "== 1" matters: auprc.py:330-347 calls the
        # per-label routine with pos_label 1), not as the int64 they are stored in
        t_loc = transpose_sharding((target == 1).to(torch.uint8)).to(torch.int64)
    else:
        # every rank needs all N class labels; shifted so that local column j is class j
        pad = torch.zeros(max(ns), dtype=torch.int64, device=dev)
        pad[:n_r] = target.to(torch.int64)
        rows = _gather_rows(pad, world, group)
        t_loc = torch.cat([rows[q, : ns[q]] for q in range(world)]) - bounds[rank]
    if width > 0 and n_all > 0:
        vals = _ranking._one_vs_rest(x_loc, t_loc, auprc, multilabel).to(torch.float64).to(dev)
    else:
        vals = torch.zeros(width, dtype=torch.float64, device=dev)
    wmax = max(bounds[q + 1] - bounds[q] for q in range(world))
    pad = torch.zeros(wmax, dtype=torch.float64, device=dev)
    pad[:width] = vals
    rows = _gather_rows(pad, world, group)
    return torch.cat([rows[q, : bounds[q + 1] - bounds[q]] for q in range(world)])


def _agree(ok: bool, cdev: torch.device, group) -> bool:
    flag = torch.tensor([1 if ok else 0], dtype=torch.int64, device=cdev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    return int(flag.item()) == 1


def try_sharded_compute(metric, group: Optional[dist.ProcessGroup] = None):
    """Fast paths of ``toolkit.sync_and_compute`` for the sample-keeping metrics held as rank-local
    shards: exact BinaryAUROC / BinaryAUPRC (one task, no weights) as a distributed sample sort,
    MulticlassAUROC / MulticlassAUPRC / MultilabelAUPRC as a class-sharded transpose.  Returns
    the metric value, or None if the metric / its state does not qualify on some rank (every
    rank then takes the generic path together)."""
    from .metrics.classification.auprc import BinaryAUPRC, MulticlassAUPRC, MultilabelAUPRC
    from .metrics.classification.auroc import BinaryAUROC, MulticlassAUROC, _resolve_weights
    from ._ranking import _as_f32_scores, compute_device, stage
    from .metrics.toolkit import _comm_device

    is_auroc = isinstance(metric, BinaryAUROC)
    is_auprc = isinstance(metric, BinaryAUPRC)
    is_ovr = isinstance(metric, (MulticlassAUROC, MulticlassAUPRC, MultilabelAUPRC))
    if not (is_auroc or is_auprc or is_ovr):
        return None
    cdev = _comm_device(metric.device, group)
    ok = len(metric.inputs) > 0
    if is_auroc or is_auprc:
        ok = ok and metric.num_tasks == 1 and metric.inputs[0].dim() == 1
        # the sample sort picks at most MAX_SAMPLE_SORT_RANKS - 1 splitters (te_select_splitters,
        # the partition kernels): larger groups take the generic gather path, like any other metric
        ok = ok and dist.get_world_size(group) <= MAX_SAMPLE_SORT_RANKS
        if is_auroc:
            ok = ok and _resolve_weights(metric.inputs, metric.weights) is None
    xs = None
    if ok:
        metric._prepare_for_merge_state()
        x, t = metric.inputs[0], metric.targets[0]
        dev = cdev if cdev.type == "cuda" else compute_device(x)
        try:
            xs = _as_f32_scores(stage(x, dev))
        except NotImplementedError:
            ok = False
    if not _agree(ok, cdev, group):  # one agreement round: every rank qualifies, or none takes this path
        return None
    ts = stage(t, dev)
    if is_ovr:
        multilabel = isinstance(metric, MultilabelAUPRC)
        vals = class_sharded_one_vs_rest(
            xs, ts, auprc=not isinstance(metric, MulticlassAUROC), multilabel=multilabel, group=group
        )
        vals = vals.to(torch.float32).to(metric.device)
        return vals.mean() if metric.average == "macro" else vals
    if is_auroc:
        out, flags = _sharded_binary_auroc(xs, ts, group)
        if flags:
            return None  # same answer on every rank (flags were all-gathered)
        return out[0].clone().to(metric.device)
    out, flags = _sharded_binary_auprc(xs, ts, group)
    if flags:
        return None
    return out[0].to(torch.float32).to(metric.device)
