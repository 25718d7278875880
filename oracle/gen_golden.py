"""Generate tests/golden/reference_vectors.npz by RUNNING THE REFERENCE (torcheval 0.0.6).

Run in the build container only (needs /root/reference, which does not exist on the GPU box):

    PYTHONPATH=/root/reference python oracle/gen_golden.py

Every case stores its inputs and the reference's outputs, so tests can check both the oracle
(oracle/te_oracle.c) and the CUDA path against the reference without importing it.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from torcheval.metrics import functional as F  # noqa: E402
from torcheval.metrics.functional.classification.binned_precision_recall_curve import (  # noqa: E402
    _binary_binned_precision_recall_curve_update,
    _multiclass_binned_precision_recall_curve_update,
    _multilabel_binned_precision_recall_curve_update,
)

OUT = Path(__file__).resolve().parent.parent / "tests" / "golden" / "reference_vectors.npz"
store: dict[str, np.ndarray] = {}


def put(case: str, **arrays) -> None:
    for k, v in arrays.items():
        if isinstance(v, torch.Tensor):
            v = v.detach().cpu().numpy()
        store[f"{case}/{k}"] = np.asarray(v)


def main() -> None:
    g = torch.Generator().manual_seed(20261106)

    def rand(*shape):
        return torch.rand(*shape, generator=g)

    def randint(lo, hi, shape):
        return torch.randint(lo, hi, shape, generator=g)

    # ---------------- exact binary AUROC / AUPRC -------------------------------------------
    n = 3000
    binary_cases = {
        "uniform": (rand(n), randint(0, 2, (n,))),
        "ties64": (randint(0, 64, (n,)).float() / 64, randint(0, 2, (n,))),
        "ties4": (randint(0, 4, (n,)).float() / 4, randint(0, 2, (n,))),
        "constant": (torch.full((n,), 0.25), randint(0, 2, (n,))),
        "all_pos": (rand(n), torch.ones(n, dtype=torch.int64)),
        "all_neg": (rand(n), torch.zeros(n, dtype=torch.int64)),
        "signed_zero": (
            torch.where(randint(0, 2, (n,)) == 0, torch.tensor(0.0), torch.tensor(-0.0))
            * (randint(0, 3, (n,)) == 0)
            + rand(n) * (randint(0, 2, (n,)) == 0),
            randint(0, 2, (n,)),
        ),
        "logits": (torch.randn(n, generator=g) * 3, randint(0, 2, (n,))),
        "skewed": (rand(n), (rand(n) < 0.01).long()),
        "int_scores": (randint(0, 2, (n,)).float(), randint(0, 2, (n,))),
        "tiny": (torch.tensor([0.1, 0.5, 0.7, 0.8]), torch.tensor([1, 0, 1, 1])),
        "single": (torch.tensor([0.3]), torch.tensor([1])),
        "n8193": (rand(8193), randint(0, 2, (8193,))),
    }
    for name, (x, y) in binary_cases.items():
        x = x.float()
        # class path of the reference = float64 weights (classification/auroc.py:112-113)
        auroc64 = F.binary_auroc(x, y, weight=torch.ones_like(x, dtype=torch.float64))
        auroc32 = F.binary_auroc(x, y)
        auprc = F.binary_auprc(x, y)
        put(f"binary/{name}", x=x, y=y, auroc_f64=auroc64, auroc_fn=auroc32, auprc=auprc)

    # weighted + soft labels + multi-task
    x, y = rand(n), randint(0, 2, (n,))
    w = rand(n).double() * 3
    put("binary_weighted/w", x=x, y=y, w=w, auroc=F.binary_auroc(x, y, weight=w))
    xs = randint(0, 32, (n,)).float() / 32
    put("binary_weighted/w_ties", x=xs, y=y, w=w, auroc=F.binary_auroc(xs, y, weight=w))
    soft = rand(n).double()
    put("binary_weighted/soft", x=xs, y=soft, w=w, auroc=F.binary_auroc(xs, soft, weight=w))
    xt, yt = rand(3, 1000), randint(0, 2, (3, 1000))
    put(
        "binary_multitask/t3",
        x=xt,
        y=yt,
        auroc=F.binary_auroc(xt, yt, num_tasks=3, weight=torch.ones_like(xt, dtype=torch.float64)),
        auprc=F.binary_auprc(xt, yt, num_tasks=3),
    )

    # ---------------- multiclass / multilabel ------------------------------------------------
    for name, (nn, c) in {"c5": (400, 5), "c17": (257, 17)}.items():
        inp = rand(nn, c)
        tgt = randint(0, c, (nn,))
        if name == "c17":
            tgt[tgt == 3] = 4  # class 3 never occurs
            inp = (inp * 8).floor() / 8  # ties
        put(
            f"multiclass/{name}",
            x=inp,
            y=tgt,
            auroc=F.multiclass_auroc(inp, tgt, num_classes=c, average=None),
            auroc_macro=F.multiclass_auroc(inp, tgt, num_classes=c),
            auprc=F.multiclass_auprc(inp, tgt, num_classes=c, average=None),
            auprc_macro=F.multiclass_auprc(inp, tgt, num_classes=c),
        )
        ml = randint(0, 2, (nn, c))
        if name == "c17":
            ml[:, 2] = 0
        put(
            f"multilabel/{name}",
            x=inp,
            y=ml,
            auprc=F.multilabel_auprc(inp, ml, num_labels=c, average=None),
            auprc_macro=F.multilabel_auprc(inp, ml, num_labels=c),
        )

    # ---------------- binned -----------------------------------------------------------------
    thr_sets = {
        "lin5": torch.linspace(0, 1, 5),
        "lin200": torch.linspace(0, 1, 200),
        "lin10000": torch.linspace(0, 1, 10000),
        "custom": torch.tensor([0.0, 0.1, 0.1, 0.35, 0.5, 0.77, 0.9, 1.0]),
        "inner": torch.tensor([0.2, 0.4, 0.6]),
    }
    xb = torch.cat([rand(1500), randint(0, 11, (500,)).float() / 10, torch.tensor([0.0, 1.0, 1.0, 0.0])])
    yb = randint(0, 2, (xb.numel(),))
    for tname, thr in thr_sets.items():
        tp, fp, fn = _binary_binned_precision_recall_curve_update(xb, yb, thr)
        rec = dict(x=xb, y=yb, thr=thr, tp=tp, fp=fp, fn=fn)
        if thr.numel() <= 200:
            auroc, _ = F.binary_binned_auroc(xb, yb, threshold=thr)
            rec["auroc"] = auroc
        if thr[0] == 0 and thr[-1] == 1:
            auprc, _ = F.binary_binned_auprc(xb, yb, threshold=thr)
            rec["auprc"] = auprc
        put(f"binned/{tname}", **rec)
    xb2, yb2 = rand(2, 700), randint(0, 2, (2, 700))
    auroc, _ = F.binary_binned_auroc(xb2, yb2, num_tasks=2, threshold=50)
    auprc, _ = F.binary_binned_auprc(xb2, yb2, num_tasks=2, threshold=50)
    put("binned_multitask/t2", x=xb2, y=yb2, thr=torch.linspace(0, 1, 50), auroc=auroc, auprc=auprc)

    # multiclass / multilabel binned PR counts (the "memory" formulation is the histogram one)
    inp = rand(300, 6)
    tgt = randint(0, 6, (300,))
    thr = torch.linspace(0, 1, 33)
    tp, fp, fn = _multiclass_binned_precision_recall_curve_update(inp, tgt, 6, thr, "memory")
    tpv, fpv, fnv = _multiclass_binned_precision_recall_curve_update(inp, tgt, 6, thr, "vectorized")
    assert torch.equal(tp, tpv) and torch.equal(fp, fpv) and torch.equal(fn, fnv)
    auprc, _ = F.multiclass_binned_auprc(inp, tgt, num_classes=6, threshold=thr, average=None)
    put("binned_multiclass/c6", x=inp, y=tgt, thr=thr, tp=tp, fp=fp, fn=fn, auprc=auprc)
    ml = randint(0, 2, (300, 6))
    tp, fp, fn = _multilabel_binned_precision_recall_curve_update(inp, ml, 6, thr, "memory")
    auprc, _ = F.multilabel_binned_auprc(inp, ml, num_labels=6, threshold=thr, average=None)
    put("binned_multilabel/l6", x=inp, y=ml, thr=thr, tp=tp, fp=fp, fn=fn, auprc=auprc)

    # ---------------- confusion matrix / accuracy / f1 ------------------------------------------
    for name, (nn, c) in {"c10": (5000, 10), "c3": (64, 3), "c100": (20000, 100)}.items():
        p, t = randint(0, c, (nn,)), randint(0, c, (nn,))
        put(f"cm/{name}", pred=p, target=t, cm=F.multiclass_confusion_matrix(p, t, num_classes=c))
    logits = torch.randn(1000, 7, generator=g)
    logits[::10, 2] = logits[::10].max(dim=1).values  # ties: first max wins
    t = randint(0, 7, (1000,))
    put("cm_logits/c7", logits=logits, target=t, cm=F.multiclass_confusion_matrix(logits, t, num_classes=7))
    logits = torch.randn(300, 130, generator=g)
    t = randint(0, 130, (300,))
    put("cm_logits/c130", logits=logits, target=t, cm=F.multiclass_confusion_matrix(logits, t, num_classes=130))
    xb3, t = rand(999), randint(0, 2, (999,))
    put("cm_binary/t05", x=xb3, target=t, cm=F.binary_confusion_matrix(xb3, t, threshold=0.5))

    p, t = randint(0, 5, (777,)), randint(0, 5, (777,))
    put(
        "acc_f1/c5",
        pred=p,
        target=t,
        acc_micro=F.multiclass_accuracy(p, t),
        acc_macro=F.multiclass_accuracy(p, t, average="macro", num_classes=5),
        acc_none=F.multiclass_accuracy(p, t, average=None, num_classes=5),
        f1_micro=F.multiclass_f1_score(p, t),
        f1_macro=F.multiclass_f1_score(p, t, num_classes=5, average="macro"),
        f1_weighted=F.multiclass_f1_score(p, t, num_classes=5, average="weighted"),
        f1_none=F.multiclass_f1_score(p, t, num_classes=5, average=None),
    )

    # ---------------- precision / recall -------------------------------
    p, t = randint(0, 6, (901,)), randint(0, 5, (901,))  # class 5 is predicted but never true
    p[p == 2] = 1                                         # class 2 is true but never predicted
    rec = dict(pred=p, target=t)
    for avg, tag in (("micro", "micro"), ("macro", "macro"), ("weighted", "weighted"), (None, "none")):
        nc = None if avg == "micro" else 7                # class 6 occurs nowhere
        rec[f"precision_{tag}"] = F.multiclass_precision(p, t, num_classes=nc, average=avg)
        # (the reference's recall masks num_tp but not num_labels for macro/weighted and raises a
        # shape error when a class is absent from both pred and target: 6 classes there)
        rec[f"recall_{tag}"] = F.multiclass_recall(p, t, num_classes=None if avg == "micro" else 6, average=avg)
    put("prec_rec/c7", **rec)
    logits, t = rand(500, 4), randint(0, 4, (500,))
    put(
        "prec_rec/logits4",
        logits=logits,
        target=t,
        precision_macro=F.multiclass_precision(logits, t, num_classes=4, average="macro"),
        recall_none=F.multiclass_recall(logits, t, num_classes=4, average=None),
    )
    xb4, t = rand(1201), randint(0, 2, (1201,))
    put(
        "prec_rec/binary",
        x=xb4,
        target=t,
        precision_05=F.binary_precision(xb4, t),
        precision_07=F.binary_precision(xb4, t, threshold=0.7),
        recall_05=F.binary_recall(xb4, t),
        recall_03=F.binary_recall(xb4, t, threshold=0.3),
    )

    # ---------------- exact precision-recall curves, recall at fixed precision (section 8f #1) ------
    for name, x, y in (
        ("ties", (randint(0, 40, (3001,)).float() / 40), randint(0, 2, (3001,))),
        ("distinct", rand(2000), randint(0, 2, (2000,))),
        ("no_pos", rand(300), torch.zeros(300, dtype=torch.int64)),
        ("logits", torch.randn(1500, generator=g) * 3, randint(0, 2, (1500,))),
        ("doc", torch.tensor([0.1, 0.5, 0.7, 0.8]), torch.tensor([0, 0, 1, 1])),
    ):
        pr, rc, th = F.binary_precision_recall_curve(x, y)
        rec = dict(x=x, y=y, precision=pr, recall=rc, thresholds=th)
        for mp in (0.0, 0.5, 0.9, 1.0):
            r, t = F.binary_recall_at_fixed_precision(x, y, min_precision=mp)
            rec[f"rafp_{int(mp * 100)}"] = torch.stack([r, t.to(r.dtype)])
        put(f"pr_curve/{name}", **rec)
    inp, tgt = rand(700, 4), randint(0, 4, (700,))
    inp[:, 1] = (inp[:, 1] * 10).round() / 10
    pr, rc, th = F.multiclass_precision_recall_curve(inp, tgt, num_classes=4)
    rec = dict(x=inp, y=tgt)
    for c in range(4):
        rec[f"precision_{c}"], rec[f"recall_{c}"], rec[f"thresholds_{c}"] = pr[c], rc[c], th[c]
    put("pr_curve_multiclass/c4", **rec)
    ml = randint(0, 2, (700, 4))
    pr, rc, th = F.multilabel_precision_recall_curve(inp, ml, num_labels=4)
    rec = dict(x=inp, y=ml)
    for c in range(4):
        rec[f"precision_{c}"], rec[f"recall_{c}"], rec[f"thresholds_{c}"] = pr[c], rc[c], th[c]
    r, t = F.multilabel_recall_at_fixed_precision(inp, ml, num_labels=4, min_precision=0.5)
    rec["rafp_50"] = torch.stack([torch.stack(r), torch.stack(t)])
    put("pr_curve_multilabel/l4", **rec)

    # ---------------- WindowedBinaryAUROC (section 8f #2) ------------------------------------------
    from torcheval.metrics import WindowedBinaryAUROC

    xs, ys = rand(2, 57) + 0.01, randint(0, 2, (2, 57))
    ws = rand(2, 57) + 0.5
    m = WindowedBinaryAUROC(num_tasks=2, max_num_samples=20)
    outs, cuts = [], [0, 7, 15, 26, 31, 57]
    for a, b in zip(cuts[:-1], cuts[1:]):
        m.update(xs[:, a:b], ys[:, a:b], ws[:, a:b] if a >= 15 else None)
        outs.append(m.compute())
    m2 = WindowedBinaryAUROC(num_tasks=2, max_num_samples=10).update(xs[:, :9], ys[:, :9])
    m3 = WindowedBinaryAUROC(num_tasks=2, max_num_samples=10).update(xs[:, 9:33], ys[:, 9:33])
    m2.merge_state([m3])
    x1, y1 = rand(33) + 0.01, randint(0, 2, (33,))
    m1 = WindowedBinaryAUROC(max_num_samples=25).update(x1, y1)
    put(
        "windowed_auroc/t2",
        x=xs, y=ys, w=ws, cuts=torch.tensor(cuts), auroc_steps=torch.stack(outs),
        merged=m2.compute(), merged_cap=torch.tensor(m2.max_num_samples), merged_next=torch.tensor(m2.next_inserted),
        x1=x1, y1=y1, single=m1.compute(),
    )

    # ---------------- appended in round 2 (after everything else: the generator sequence above is unchanged)
    # multi-task AUROC with per-sample weights on logits (the batched general path); row 1 carries a
    # weight float32 cannot hold, row 2 soft labels
    K, nt = 4, 2000
    xk = torch.randn(K, nt, generator=g)
    xk[:, : nt // 3] = torch.round(xk[:, : nt // 3])
    yk = randint(0, 2, (K, nt)).double()
    wk = rand(K, nt).double()
    wk[1, 7] = 1.0 + 2.0**-40
    yk[2] = rand(nt).double()
    put("binary_multitask_weighted/t4", x=xk, y=yk, w=wk, auroc=F.binary_auroc(xk, yk, num_tasks=K, weight=wk))
    # float64 scores that float32 cannot tell apart (the dense-rank road), and int64 beyond 2^24
    x64 = (randint(0, 50, (n,)).double() / 50) + randint(0, 5, (n,)).double() * 1e-9
    y64 = randint(0, 2, (n,))
    put("binary_f64/close", x=x64, y=y64, auroc=F.binary_auroc(x64, y64), auprc=F.binary_auprc(x64, y64))
    xi64 = (1 << 40) + randint(0, 40, (n,))
    put("binary_f64/int64", x=xi64, y=y64, auroc=F.binary_auroc(xi64, y64), auprc=F.binary_auprc(xi64, y64))

    OUT.parent.mkdir(parents=True, exist_ok=True)
    np.savez_compressed(OUT, **store)
    print(f"wrote {OUT} ({OUT.stat().st_size / 1024:.0f} KiB, {len(store)} arrays)")


if __name__ == "__main__":
    main()
