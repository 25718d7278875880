"""Worker for tests/test_distributed_cpu.py: one rank of a gloo group on CPU.  The kernels are
replaced by the CPU stand-ins (tests/cpu_backend.py); what is under test is the host-side
protocol: splitters, partition/exchange bookkeeping, prefix fix-up, all-reduce of counters,
tensor all-gathers, and that every rank gets the same answer."""
import json
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import cpu_backend  # noqa: E402
import oracle  # noqa: E402


def shard(a, rank, world, axis=0):
    return np.array_split(a, world, axis=axis)[rank]


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    cpu_backend.install()
    from torcheval_b200 import distributed as tdist
    from torcheval_b200.metrics import (
        BinaryAUPRC, BinaryAUROC, BinaryBinnedAUPRC, Metric, MulticlassAccuracy, MulticlassAUPRC, MulticlassAUROC,
        MulticlassConfusionMatrix, MultilabelAUPRC, WindowedBinaryAUROC,
    )
    from torcheval_b200.metrics import toolkit

    res = {}
    calls = {"auprc": 0, "ovr": 0, "auroc": 0, "gather": 0}
    real_auprc, real_ovr = tdist._sharded_binary_auprc, tdist.class_sharded_one_vs_rest
    real_auroc, real_gather = tdist._sharded_binary_auroc, toolkit._gather_metrics

    def counting_auroc(*a, **k):
        calls["auroc"] += 1
        return real_auroc(*a, **k)

    def counting_gather(*a, **k):
        calls["gather"] += 1
        return real_gather(*a, **k)

    tdist._sharded_binary_auroc, toolkit._gather_metrics = counting_auroc, counting_gather

    def counting_auprc(*a, **k):
        calls["auprc"] += 1
        return real_auprc(*a, **k)

    def counting_ovr(*a, **k):
        calls["ovr"] += 1
        return real_ovr(*a, **k)

    tdist._sharded_binary_auprc, tdist.class_sharded_one_vs_rest = counting_auprc, counting_ovr
    rs = np.random.default_rng(7)
    n = 20011
    cases = {
        "uniform": (rs.random(n, dtype=np.float32), rs.integers(0, 2, n)),
        "ties": ((rs.integers(0, 50, n) / 50).astype(np.float32), rs.integers(0, 2, n)),
        "heavy_tie": (np.where(rs.random(n) < 0.6, 0.5, rs.random(n)).astype(np.float32), rs.integers(0, 2, n)),
        "constant": (np.full(n, 0.25, dtype=np.float32), rs.integers(0, 2, n)),
        "sorted_input": (np.sort(rs.random(n, dtype=np.float32)), rs.integers(0, 2, n)),
        "all_pos": (rs.random(n, dtype=np.float32), np.ones(n, dtype=np.int64)),
    }
    for name, (x, y) in cases.items():
        # uneven shards, the last rank gets most
        cut = [0] + [int(n * (i + 1) / (world * 3)) for i in range(world - 1)] + [n]
        xs, ys = x[cut[rank]:cut[rank + 1]], y[cut[rank]:cut[rank + 1]]
        out = tdist.sharded_binary_auroc(torch.from_numpy(xs), torch.from_numpy(ys))
        want, tp, fp = oracle.binary_auroc(x, y)
        res[f"sharded/{name}"] = [float(out[0]), want, float(out[1]), tp, float(out[2]), fp]
        m = BinaryAUROC()
        half = xs.size // 2
        m.update(torch.from_numpy(xs[:half]), torch.from_numpy(ys[:half]))
        m.update(torch.from_numpy(xs[half:]), torch.from_numpy(ys[half:]))
        res[f"sync/{name}"] = [float(toolkit.sync_and_compute(m)), want]
        assert len(m.inputs) == 1  # compacted by _prepare_for_merge_state, data unchanged
        # average precision over the same shards: two-phase integrate around the (P_r, N_r) exchange
        out = tdist.sharded_binary_auprc(torch.from_numpy(xs), torch.from_numpy(ys))
        res[f"sharded_ap/{name}"] = [float(out[0]), oracle.binary_auprc(x, y), float(out[1]), tp, float(out[2]), fp]
        ma = BinaryAUPRC().update(torch.from_numpy(xs), torch.from_numpy(ys))
        got = toolkit.sync_and_compute(ma)
        assert got.dtype == torch.float32 and got.dim() == 0
        res[f"sync_ap/{name}"] = [float(got), oracle.binary_auprc(x, y)]
    # negative scores: the sharded path declines on every rank, the gather path answers
    x, y = (rs.standard_normal(n) * 2).astype(np.float32), rs.integers(0, 2, n)
    m = BinaryAUROC().update(torch.from_numpy(shard(x, rank, world)), torch.from_numpy(shard(y, rank, world)))
    res["sync/logits"] = [float(toolkit.sync_and_compute(m)), oracle.binary_auroc(x, y)[0]]
    # weighted: gather path with weights
    w = rs.random(n) + 0.5
    m = BinaryAUROC().update(
        torch.from_numpy(shard(x, rank, world)), torch.from_numpy(shard(y, rank, world)), torch.from_numpy(shard(w, rank, world))
    )
    res["sync/weighted"] = [float(toolkit.sync_and_compute(m)), oracle.binary_auroc(x, y, w)[0]]
    m = BinaryAUPRC().update(torch.from_numpy(shard(x, rank, world)), torch.from_numpy(shard(y, rank, world)))
    res["sync/auprc"] = [float(toolkit.sync_and_compute(m)), oracle.binary_auprc(x, y)]
    # counters: one all-reduce
    C = 7
    p, t = rs.integers(0, C, n), rs.integers(0, C, n)
    cm = MulticlassConfusionMatrix(C).update(torch.from_numpy(shard(p, rank, world)), torch.from_numpy(shard(t, rank, world)))
    local_before = cm.confusion_matrix.clone()
    got = toolkit.sync_and_compute(cm)
    res["sync/cm_equal"] = bool(np.array_equal(got.numpy(), oracle.confusion_matrix(p, t, C)[0].astype(np.float32)))
    res["sync/cm_input_unchanged"] = bool(torch.equal(cm.confusion_matrix, local_before))
    acc = MulticlassAccuracy().update(torch.from_numpy(shard(p, rank, world)), torch.from_numpy(shard(t, rank, world)))
    xb = rs.random(n, dtype=np.float32)
    bb = BinaryBinnedAUPRC(threshold=64).update(torch.from_numpy(shard(xb, rank, world)), torch.from_numpy(shard(y, rank, world)))
    coll = toolkit.sync_and_compute_collection({"acc": acc, "cm": cm, "bb": bb})
    res["coll/acc"] = [float(coll["acc"]), float((p == t).mean())]
    tp_, fp_, fn_ = oracle.binned_update(xb, y, np.linspace(0, 1, 64, dtype=np.float32))
    res["coll/bb"] = [float(coll["bb"][0][0]), oracle.binned_auprc_from_counts(tp_, fp_, fn_)]
    sd = toolkit.get_synced_state_dict(cm)
    res["sd/cm_total"] = [int(sd["confusion_matrix"].sum()), n]
    # multiclass AUROC: tensor all-gather path (uneven shards)
    nn, CC = 901, 4
    xi, yi = rs.random((nn, CC), dtype=np.float32), rs.integers(0, CC, nn)
    cut = [0, 100 if world > 1 else nn, nn] if world == 2 else [0] + [nn * (i + 1) // world for i in range(world)]
    mm = MulticlassAUROC(num_classes=CC, average=None).update(
        torch.from_numpy(xi[cut[rank]:cut[rank + 1]]), torch.from_numpy(yi[cut[rank]:cut[rank + 1]])
    )
    res["sync/mc_auroc"] = [toolkit.sync_and_compute(mm).tolist(), oracle.multiclass_auroc(xi, yi).tolist()]
    # class-sharded transpose: more classes than ranks, uneven column blocks, macro and per-class
    nn, CC = 1203, 7
    xi, yi = rs.random((nn, CC), dtype=np.float32), rs.integers(0, CC, nn)
    xi[:, 2] = np.round(xi[:, 2] * 8) / 8  # one heavily tied column
    li = rs.integers(0, 2, (nn, CC))
    cut = [0] + [int(nn * (i + 1) / (world * 2)) for i in range(world - 1)] + [nn]
    sl = slice(cut[rank], cut[rank + 1])
    for avg in ("macro", None):
        tag = "macro" if avg else "none"
        m1 = MulticlassAUROC(num_classes=CC, average=avg).update(torch.from_numpy(xi[sl]), torch.from_numpy(yi[sl]))
        want = oracle.multiclass_auroc(xi, yi)
        res[f"ovr/mc_auroc_{tag}"] = [np.atleast_1d(toolkit.sync_and_compute(m1).numpy()).tolist(), np.atleast_1d(want.mean() if avg else want).tolist()]
        m2 = MulticlassAUPRC(num_classes=CC, average=avg).update(torch.from_numpy(xi[sl]), torch.from_numpy(yi[sl]))
        want = oracle.multiclass_auprc(xi, yi)
        res[f"ovr/mc_auprc_{tag}"] = [np.atleast_1d(toolkit.sync_and_compute(m2).numpy()).tolist(), np.atleast_1d(want.mean() if avg else want).tolist()]
        m3 = MultilabelAUPRC(num_labels=CC, average=avg).update(torch.from_numpy(xi[sl]), torch.from_numpy(li[sl]))
        want = oracle.multilabel_auprc(xi, li)
        res[f"ovr/ml_auprc_{tag}"] = [np.atleast_1d(toolkit.sync_and_compute(m3).numpy()).tolist(), np.atleast_1d(want.mean() if avg else want).tolist()]
    # fewer classes than ranks can happen (C=2, world=3): some rank owns no column
    x2, y2 = rs.random((300, 2), dtype=np.float32), rs.integers(0, 2, 300)
    m4 = MulticlassAUROC(num_classes=2, average=None).update(
        torch.from_numpy(shard(x2, rank, world)), torch.from_numpy(shard(y2, rank, world))
    )
    res["ovr/two_classes"] = [toolkit.sync_and_compute(m4).tolist(), oracle.multiclass_auroc(x2, y2).tolist()]
    res["calls"] = [calls["auprc"], calls["ovr"]]
    # a collection mixing every kind of member: the exact metrics take their sharded paths (nobody
    # gathers their samples), the counters share one all-reduce, the weighted AUROC and the windowed
    # metric — no sharded path — share ONE flat all-gather
    x, y = cases["ties"]
    xs_, ys_ = torch.from_numpy(shard(x, rank, world)), torch.from_numpy(shard(y, rank, world))
    ws_ = torch.from_numpy(shard(w, rank, world))
    before = dict(calls)
    coll = toolkit.sync_and_compute_collection({
        "auroc": BinaryAUROC().update(xs_, ys_),
        "cm": cm,
        "auprc": BinaryAUPRC().update(xs_, ys_),
        "mc": MulticlassAUROC(num_classes=CC, average=None).update(torch.from_numpy(xi[sl]), torch.from_numpy(yi[sl])),
        "bb": bb,
        "weighted": BinaryAUROC().update(xs_, ys_, ws_),
        "windowed": WindowedBinaryAUROC(max_num_samples=50).update(xs_[:50], ys_[:50]),
    })
    res["coll2/auroc"] = [float(coll["auroc"]), oracle.binary_auroc(x, y)[0]]
    res["coll2/auprc"] = [float(coll["auprc"]), oracle.binary_auprc(x, y)]
    res["coll2/weighted"] = [float(coll["weighted"]), oracle.binary_auroc(x, y, w)[0]]
    res["coll2/mc"] = [coll["mc"].tolist(), oracle.multiclass_auroc(xi, yi).tolist()]
    res["coll2/cm_equal"] = bool(np.array_equal(coll["cm"].numpy(), oracle.confusion_matrix(p, t, C)[0].astype(np.float32)))
    res["coll2/paths"] = [
        [calls["auroc"] - before["auroc"], calls["auprc"] - before["auprc"], calls["ovr"] - before["ovr"], calls["gather"] - before["gather"]],
        [1, 1, 1, 1],
    ]
    # more ranks than the sample sort has splitters for: everybody takes the generic path together
    tdist.MAX_SAMPLE_SORT_RANKS = world - 1
    before = dict(calls)
    res["sync/over_cap"] = [float(toolkit.sync_and_compute(BinaryAUROC().update(xs_, ys_))), oracle.binary_auroc(x, y)[0]]
    res["sync/over_cap_paths"] = [[calls["auroc"] - before["auroc"]], [0]]
    tdist.MAX_SAMPLE_SORT_RANKS = 16
    # a rank with no data: generic object path still works
    me = BinaryAUROC()
    if rank != 0:
        me.update(torch.from_numpy(xs), torch.from_numpy(ys))
    try:
        v = float(toolkit.sync_and_compute(me))
        res["sync/empty_rank_ok"] = bool(np.isfinite(v))
    except Exception as e:  # noqa: BLE001
        res["sync/empty_rank_ok"] = f"raised {type(e).__name__}: {e}"
    # a user-defined metric with every kind of state (tensor, list, dict, int): the generic path
    class Toy(Metric):
        def __init__(self):
            super().__init__()
            self._add_state("total", torch.zeros(3, dtype=torch.float64))
            self._add_state("chunks", [])
            self._add_state("per_key", {})
            self._add_state("count", 0)

        def update(self, v):
            self.total += v.sum(0)
            self.chunks.append(v)
            self.per_key[f"r{int(v[0, 0])}"] = v.mean()
            self.count += v.shape[0]
            return self

        def compute(self):
            return self.total, torch.cat(self.chunks).shape[0] if self.chunks else 0, dict(self.per_key), self.count

        def merge_state(self, metrics):
            for m in metrics:
                self.total += m.total
                self.chunks.extend(m.chunks)
                self.per_key.update(m.per_key)
                self.count += m.count
            return self

    toy = Toy()
    if rank != 1:  # rank 1 stays empty: empty list / dict states must travel too
        toy.update(torch.full((rank + 2, 3), float(rank), dtype=torch.float64))
    total, rows, keys, count = toolkit.sync_and_compute(toy)
    want_rows = sum(r + 2 for r in range(world) if r != 1)
    res["generic/toy"] = [
        total.tolist(), [float(sum(r * (r + 2) for r in range(world) if r != 1))] * 3,
    ]
    res["generic/toy_meta"] = [[rows, count, sorted(keys)], [want_rows, want_rows, sorted(f"r{r}" for r in range(world) if r != 1)]]
    assert toy.count == (0 if rank == 1 else rank + 2)  # the local object is left alone
    wx, wy = rs.random((world, 30), dtype=np.float32) + 0.01, rs.integers(0, 2, (world, 30))
    wm = WindowedBinaryAUROC(max_num_samples=20).update(torch.from_numpy(wx[rank]), torch.from_numpy(wy[rank]))
    wv = float(toolkit.sync_and_compute(wm))
    # merged window = the last 20 samples of every rank, concatenated
    res["generic/windowed"] = [wv, oracle.binary_auroc(np.concatenate([wx[r, 10:] for r in range(world)]), np.concatenate([wy[r, 10:] for r in range(world)]))[0]]
    # multi-task BinaryAUROC with weights (2-D states: the gather path; compute = the batched general path)
    K, nt = 3, 4001
    mx, my, mw = rs.standard_normal((K, nt)).astype(np.float32), rs.integers(0, 2, (K, nt)), rs.random((K, nt)).astype(np.float32)
    cut2 = [0] + [int(nt * (i + 1) / (world * 2)) for i in range(world - 1)] + [nt]
    sl2 = slice(cut2[rank], cut2[rank + 1])
    mt = BinaryAUROC(num_tasks=K).update(torch.from_numpy(mx[:, sl2]), torch.from_numpy(my[:, sl2]), torch.from_numpy(mw[:, sl2]))
    res["sync/multitask_weighted"] = [
        toolkit.sync_and_compute(mt).tolist(),
        [oracle.binary_auroc(mx[k], my[k], mw[k].astype(np.float64))[0] for k in range(K)],
    ]
    # float64 scores that float32 cannot tell apart: the sync must rank them as float64 (gather path)
    base = rs.integers(0, 300, nt) / 300.0
    x64 = base + rs.integers(0, 4, nt) * 2.0**-40
    y64 = (rs.random(nt) < 0.3 + 0.4 * base).astype(np.int64)
    m64 = BinaryAUROC().update(torch.from_numpy(x64[sl2]), torch.from_numpy(y64[sl2]))
    _, inv = np.unique(x64, return_inverse=True)
    res["sync/float64_scores"] = [float(toolkit.sync_and_compute(m64)), oracle.binary_auroc(inv.astype(np.float32), y64)[0]]
    Path(os.environ["TE_TEST_OUT"] + f".{rank}").write_text(json.dumps(res))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
