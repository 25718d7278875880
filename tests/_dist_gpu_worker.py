"""One rank of an NCCL group: the same checks as tests/_dist_worker.py with the real kernels."""
import json
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import oracle  # noqa: E402


def main():
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    from torcheval_b200 import distributed as tdist
    from torcheval_b200.metrics import (
        BinaryAUPRC, BinaryAUROC, BinaryBinnedAUPRC, MulticlassAUPRC, MulticlassAUROC, MulticlassConfusionMatrix,
        MultilabelAUPRC,
    )
    from torcheval_b200.metrics import toolkit

    res = {}
    rs = np.random.default_rng(11)
    n = 1 << 21
    cases = {
        "uniform": (rs.random(n, dtype=np.float32), rs.integers(0, 2, n)),
        "ties": ((rs.integers(0, 999, n) / 999).astype(np.float32), rs.integers(0, 2, n)),
        "heavy_tie": (np.where(rs.random(n) < 0.6, 0.5, rs.random(n)).astype(np.float32), rs.integers(0, 2, n)),
    }
    for name, (x, y) in cases.items():
        cut = [0] + [int(n * (i + 1) / (world * 2)) for i in range(world - 1)] + [n]
        xs = torch.from_numpy(x[cut[rank]:cut[rank + 1]]).to(dev)
        ys = torch.from_numpy(y[cut[rank]:cut[rank + 1]]).to(dev)
        out = tdist.sharded_binary_auroc(xs, ys).cpu().numpy()
        want, tp, fp = oracle.binary_auroc(x, y)
        res[f"sharded/{name}"] = [float(out[0]), want, float(out[1]), tp, float(out[2]), fp]
        m = BinaryAUROC(device=dev).update(xs, ys)
        res[f"sync/{name}"] = [float(toolkit.sync_and_compute(m)), want]
        out = tdist.sharded_binary_auprc(xs, ys).cpu().numpy()
        want_ap = oracle.binary_auprc(x, y)
        res[f"sharded_ap/{name}"] = [float(out[0]), want_ap, float(out[1]), tp, float(out[2]), fp]
        ma = BinaryAUPRC(device=dev).update(xs, ys)
        res[f"sync_ap/{name}"] = [float(toolkit.sync_and_compute(ma)), want_ap]
        mh = BinaryAUROC().update(xs.cpu(), ys.cpu())  # host-resident state, NCCL group
        res[f"sync_host/{name}"] = [float(toolkit.sync_and_compute(mh)), want]
    C = 1000
    p, t = rs.integers(0, C, n), rs.integers(0, C, n)
    sl = slice(rank * n // world, (rank + 1) * n // world)
    cm = MulticlassConfusionMatrix(C, device=dev).update(torch.from_numpy(p[sl]).to(dev), torch.from_numpy(t[sl]).to(dev))
    got = toolkit.sync_and_compute(cm)
    res["sync/cm_equal"] = bool(np.array_equal(got.cpu().numpy(), oracle.confusion_matrix(p, t, C)[0].astype(np.float32)))
    xb = rs.random(n, dtype=np.float32)
    yb = rs.integers(0, 2, n)
    bb = BinaryBinnedAUPRC(threshold=10000, device=dev).update(torch.from_numpy(xb[sl]).to(dev), torch.from_numpy(yb[sl]).to(dev))
    tp_, fp_, fn_ = oracle.binned_update(xb, yb, np.linspace(0, 1, 10000, dtype=np.float32))
    res["sync/binned_auprc"] = [float(toolkit.sync_and_compute(bb)[0][0]), oracle.binned_auprc_from_counts(tp_, fp_, fn_)]
    # class-sharded transpose (one all-to-all), uneven row shards and column blocks
    nn, CC = 50021, 13
    xi, yi = rs.random((nn, CC), dtype=np.float32), rs.integers(0, CC, nn)
    xi[:, 3] = np.round(xi[:, 3] * 16) / 16
    li = rs.integers(0, 2, (nn, CC))
    cut = [0] + [int(nn * (i + 1) / (world * 2)) for i in range(world - 1)] + [nn]
    sl = slice(cut[rank], cut[rank + 1])
    xd, yd, ld = torch.from_numpy(xi[sl]).to(dev), torch.from_numpy(yi[sl]).to(dev), torch.from_numpy(li[sl]).to(dev)
    m1 = MulticlassAUROC(num_classes=CC, average=None, device=dev).update(xd, yd)
    res["ovr/mc_auroc"] = [toolkit.sync_and_compute(m1).tolist(), oracle.multiclass_auroc(xi, yi).tolist()]
    m2 = MulticlassAUPRC(num_classes=CC, average=None, device=dev).update(xd, yd)
    res["ovr/mc_auprc"] = [toolkit.sync_and_compute(m2).tolist(), oracle.multiclass_auprc(xi, yi).tolist()]
    m3 = MultilabelAUPRC(num_labels=CC, average="macro", device=dev).update(xd, ld)
    res["ovr/ml_auprc_macro"] = [float(toolkit.sync_and_compute(m3)), float(oracle.multilabel_auprc(xi, li).mean())]
    # a collection: exact metrics take their sharded paths (no rank gathers their samples), the
    # counters share one all-reduce, what is left shares one flat all-gather
    calls = {"auroc": 0, "auprc": 0, "ovr": 0, "gather": 0}

    def counted(name, fn):
        def wrap(*a, **k):
            calls[name] += 1
            return fn(*a, **k)
        return wrap

    tdist._sharded_binary_auroc = counted("auroc", tdist._sharded_binary_auroc)
    tdist._sharded_binary_auprc = counted("auprc", tdist._sharded_binary_auprc)
    tdist.class_sharded_one_vs_rest = counted("ovr", tdist.class_sharded_one_vs_rest)
    toolkit._gather_metrics = counted("gather", toolkit._gather_metrics)
    x, y = cases["ties"]
    cut = [0] + [int(n * (i + 1) / (world * 2)) for i in range(world - 1)] + [n]
    xs = torch.from_numpy(x[cut[rank]:cut[rank + 1]]).to(dev)
    ys = torch.from_numpy(y[cut[rank]:cut[rank + 1]]).to(dev)
    wts = torch.from_numpy(rs.random(n)[cut[rank]:cut[rank + 1]] + 0.5).to(dev)
    coll = toolkit.sync_and_compute_collection({
        "auroc": BinaryAUROC(device=dev).update(xs, ys),
        "cm": cm,
        "auprc": BinaryAUPRC(device=dev).update(xs, ys),
        "mc": m1,
        "bb": bb,
        "weighted": BinaryAUROC(device=dev).update(xs[:1000], ys[:1000], wts[:1000]),
    })
    res["coll/auroc"] = [float(coll["auroc"]), oracle.binary_auroc(x, y)[0]]
    res["coll/auprc"] = [float(coll["auprc"]), oracle.binary_auprc(x, y)]
    res["coll/binned_auprc"] = [float(coll["bb"][0][0]), oracle.binned_auprc_from_counts(tp_, fp_, fn_)]
    res["coll/cm_equal"] = bool(np.array_equal(coll["cm"].cpu().numpy(), oracle.confusion_matrix(p, t, C)[0].astype(np.float32)))
    res["coll/mc"] = [coll["mc"].tolist(), oracle.multiclass_auroc(xi, yi).tolist()]
    res["coll/paths"] = [[calls["auroc"], calls["auprc"], calls["ovr"], calls["gather"]], [1, 1, 1, 1]]
    # generic state gather over NCCL (flat byte buffer): a metric without a dedicated sync path
    from torcheval_b200.metrics import WindowedBinaryAUROC

    wx, wy = rs.random((world, 300), dtype=np.float32) + 0.01, rs.integers(0, 2, (world, 300))
    wm = WindowedBinaryAUROC(max_num_samples=200, device=dev).update(torch.from_numpy(wx[rank]).to(dev), torch.from_numpy(wy[rank]).to(dev))
    res["generic/windowed"] = [
        float(toolkit.sync_and_compute(wm)),
        oracle.binary_auroc(np.concatenate([wx[r, 100:] for r in range(world)]), np.concatenate([wy[r, 100:] for r in range(world)]))[0],
    ]
    Path(os.environ["TE_TEST_OUT"] + f".{rank}").write_text(json.dumps(res))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
