"""The drop-in surface (functional + class API) against the reference fixtures.

Every test runs twice: with the CPU stand-in kernels of tests/cpu_backend.py (host logic only,
no GPU needed) and, marked ``gpu``, with the real sm_100a kernels — same assertions."""
import copy
import pickle

import numpy as np
import pytest
import torch

import cpu_backend


@pytest.fixture(params=["cpu_standin", pytest.param("cuda", marks=pytest.mark.gpu)])
def backend(request, monkeypatch):
    if request.param == "cpu_standin":
        cpu_backend.install(monkeypatch)
    else:
        assert torch.cuda.is_available()
    return request.param


def T(a, dtype=None):
    t = torch.as_tensor(np.ascontiguousarray(a))
    return t if dtype is None else t.to(dtype)


def close(a, b, rtol=1e-6):
    a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else a
    np.testing.assert_allclose(np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64), rtol=rtol, atol=rtol * 1e-2)


# ------------------------------------------------------------------------------ functional
def test_functional_exact(backend, golden):
    from torcheval_b200.metrics import functional as F

    for name in golden.cases("binary"):
        c = golden.case(f"binary/{name}")
        a = F.binary_auroc(T(c["x"]), T(c["y"]))
        assert a.dtype == torch.float64 and a.shape == () and a.device.type == "cpu"
        close(a, c["auroc_f64"])
        p = F.binary_auprc(T(c["x"]), T(c["y"]))
        assert p.dtype == torch.float32 and p.shape == ()
        close(p, c["auprc"], 1e-4)
    c = golden.case("binary_weighted/soft")
    close(F.binary_auroc(T(c["x"]), T(c["y"]), weight=T(c["w"])), c["auroc"])
    c = golden.case("binary_multitask/t3")
    close(F.binary_auroc(T(c["x"]), T(c["y"]), num_tasks=3), c["auroc"])
    p = F.binary_auprc(T(c["x"]), T(c["y"]), num_tasks=3)
    assert p.shape == (3,) and p.dtype == torch.float32
    close(p, c["auprc"], 1e-4)
    assert F.binary_auprc(T(c["x"][:1]), T(c["y"][:1])).shape == (1,)  # 1 x N keeps its leading dim
    # reference outputs for the paths added in round 2: (tasks, n) with weights (one launch train for
    # all tasks; row 1 carries a weight float32 cannot hold, row 2 soft labels) and 64-bit scores
    c = golden.case("binary_multitask_weighted/t4")
    a = F.binary_auroc(T(c["x"]), T(c["y"]), num_tasks=4, weight=T(c["w"]))
    assert a.shape == (4,) and a.dtype == torch.float64
    close(a, c["auroc"])
    for name in ("close", "int64"):
        c = golden.case(f"binary_f64/{name}")
        close(F.binary_auroc(T(c["x"]), T(c["y"])), c["auroc"])
        close(F.binary_auprc(T(c["x"]), T(c["y"])), c["auprc"], 1e-4)
    for name in golden.cases("multiclass"):
        c = golden.case(f"multiclass/{name}")
        C = c["x"].shape[1]
        a = F.multiclass_auroc(T(c["x"]), T(c["y"]), num_classes=C, average=None)
        assert a.dtype == torch.float32 and a.shape == (C,)
        close(a, c["auroc"], 1e-4)
        close(F.multiclass_auroc(T(c["x"]), T(c["y"]), num_classes=C), c["auroc_macro"], 1e-4)
        close(F.multiclass_auprc(T(c["x"]), T(c["y"]), average=None), c["auprc"], 1e-4)
        close(F.multiclass_auprc(T(c["x"]), T(c["y"]), num_classes=C), c["auprc_macro"], 1e-4)
    for name in golden.cases("multilabel"):
        c = golden.case(f"multilabel/{name}")
        close(F.multilabel_auprc(T(c["x"]), T(c["y"]), average=None), c["auprc"], 1e-4)
        close(F.multilabel_auprc(T(c["x"]), T(c["y"])), c["auprc_macro"], 1e-4)


def test_float64_and_integer_scores(backend):
    """Scores float32 cannot hold (float64 probabilities 1e-9 apart, int64 beyond 2^24): the
    reference ranks them as they are (auroc.py:121 sorts whatever dtype it is given); here they
    go through the dense-rank transform (ranking) or are floored to float32 (threshold counts)."""
    import oracle
    from torcheval_b200.metrics import functional as F

    rs = np.random.default_rng(99)
    n = 20001
    y = rs.integers(0, 2, n)
    x64 = 0.5 + rs.integers(0, 4000, n) * 1e-9  # distinct in float64, 2 values in float32
    assert np.unique(x64.astype(np.float32)).size < 100 < np.unique(x64).size

    def ranks(a):  # the oracle takes float32 scores: hand it small integers with the same order and ties
        return np.unique(a, return_inverse=True)[1].reshape(a.shape).astype(np.float32)

    close(F.binary_auroc(T(x64), T(y)), oracle.binary_auroc(ranks(x64), y)[0])
    close(F.binary_auprc(T(x64), T(y)), oracle.binary_auprc(ranks(x64), y), 1e-4)
    w = rs.random(n) + 0.5
    close(F.binary_auroc(T(x64), T(y), weight=T(w)), oracle.binary_auroc(ranks(x64), y, w)[0])
    xi = (1 << 40) + rs.integers(0, 50, n)  # int64 scores that collapse in float32
    close(F.binary_auroc(T(xi), T(y)), oracle.binary_auroc(ranks(xi), y)[0])
    C = 5
    xm = 0.25 + rs.integers(0, 1000, (n, C)) * 1e-10
    ym = rs.integers(0, C, n)
    close(F.multiclass_auroc(T(xm), T(ym), num_classes=C, average=None), oracle.multiclass_auroc(ranks(xm), ym), 1e-4)
    # threshold counting: a float64 score just below a float32 threshold must stay below it
    thr = np.linspace(0, 1, 11, dtype=np.float32)
    xb = np.concatenate([thr.astype(np.float64) - 1e-12, thr.astype(np.float64) + 1e-12, thr.astype(np.float64)])
    xb = np.clip(xb, 0.0, 1.0)
    yb = rs.integers(0, 2, xb.size)
    from torcheval_b200.metrics.functional.classification.binned_precision_recall_curve import (
        _binary_binned_precision_recall_curve_update,
    )

    tp, fp, fn = _binary_binned_precision_recall_curve_update(T(xb), T(yb), T(thr))
    ge = xb[None, :] >= thr.astype(np.float64)[:, None]
    np.testing.assert_array_equal(tp.numpy(), (ge & (yb == 1)[None, :]).sum(1))
    np.testing.assert_array_equal(fp.numpy(), (ge & (yb == 0)[None, :]).sum(1))


def test_functional_binned(backend, golden):
    from torcheval_b200.metrics import functional as F
    from torcheval_b200.metrics.functional.classification.binned_precision_recall_curve import (
        _binary_binned_precision_recall_curve_update,
    )

    for name in golden.cases("binned"):
        c = golden.case(f"binned/{name}")
        thr = T(c["thr"])
        tp, fp, fn = _binary_binned_precision_recall_curve_update(T(c["x"]), T(c["y"]), thr)
        assert tp.dtype == torch.int64
        np.testing.assert_array_equal(tp.numpy(), c["tp"])
        np.testing.assert_array_equal(fp.numpy(), c["fp"])
        np.testing.assert_array_equal(fn.numpy(), c["fn"])
        prec, rec, t2 = F.binary_binned_precision_recall_curve(T(c["x"]), T(c["y"]), threshold=thr)
        assert prec.shape == (thr.numel() + 1,) and prec[-1] == 1 and rec[-1] == 0 and t2 is thr
        if "auroc" in c:
            a, _ = F.binary_binned_auroc(T(c["x"]), T(c["y"]), threshold=thr)
            assert a.dtype == torch.float64 and a.shape == (1,)
            close(a, c["auroc"])
        if "auprc" in c:
            p, _ = F.binary_binned_auprc(T(c["x"]), T(c["y"]), threshold=thr)
            assert p.dtype == torch.float32 and p.shape == ()
            close(p, c["auprc"], 1e-4)
    c = golden.case("binned_multitask/t2")
    a, thr = F.binary_binned_auroc(T(c["x"]), T(c["y"]), num_tasks=2, threshold=50)
    close(a, c["auroc"])
    np.testing.assert_array_equal(thr.numpy(), c["thr"])
    close(F.binary_binned_auprc(T(c["x"]), T(c["y"]), num_tasks=2, threshold=50)[0], c["auprc"], 1e-4)
    c = golden.case("binned_multiclass/c6")
    from torcheval_b200.metrics.functional.classification.binned_precision_recall_curve import (
        _multiclass_binned_precision_recall_curve_update as mc_update,
        _multilabel_binned_precision_recall_curve_update as ml_update,
    )

    for opt in ("vectorized", "memory"):
        tp, fp, fn = mc_update(T(c["x"]), T(c["y"]), 6, T(c["thr"]), opt)
        np.testing.assert_array_equal(tp.numpy(), c["tp"])
        np.testing.assert_array_equal(fp.numpy(), c["fp"])
        np.testing.assert_array_equal(fn.numpy(), c["fn"])
    close(F.multiclass_binned_auprc(T(c["x"]), T(c["y"]), num_classes=6, threshold=T(c["thr"]), average=None)[0], c["auprc"], 1e-4)
    pr, rc, _ = F.multiclass_binned_precision_recall_curve(T(c["x"]), T(c["y"]), num_classes=6, threshold=T(c["thr"]))
    assert len(pr) == 6 and pr[0].shape == (c["thr"].size + 1,)
    c = golden.case("binned_multilabel/l6")
    tp, fp, fn = ml_update(T(c["x"]), T(c["y"]), 6, T(c["thr"]))
    np.testing.assert_array_equal(tp.numpy(), c["tp"])
    np.testing.assert_array_equal(fn.numpy(), c["fn"])
    close(F.multilabel_binned_auprc(T(c["x"]), T(c["y"]), num_labels=6, threshold=T(c["thr"]), average=None)[0], c["auprc"], 1e-4)


def test_functional_counting(backend, golden):
    from torcheval_b200.metrics import functional as F

    for name in golden.cases("cm"):
        c = golden.case(f"cm/{name}")
        cm = F.multiclass_confusion_matrix(T(c["pred"]), T(c["target"]), c["cm"].shape[0])
        assert cm.dtype == torch.int64
        np.testing.assert_array_equal(cm.numpy(), c["cm"])
    for name in golden.cases("cm_logits"):
        c = golden.case(f"cm_logits/{name}")
        cm = F.multiclass_confusion_matrix(T(c["logits"]), T(c["target"]), c["cm"].shape[0])
        np.testing.assert_array_equal(cm.numpy(), c["cm"])
    c = golden.case("cm_binary/t05")
    np.testing.assert_array_equal(F.binary_confusion_matrix(T(c["x"]), T(c["target"])).numpy(), c["cm"])
    n = F.multiclass_confusion_matrix(T(c["x"] >= 0.5, torch.int64), T(c["target"]), 2, normalize="all")
    assert n.dtype == torch.float32
    close(n.sum(), 1.0)
    c = golden.case("acc_f1/c5")
    p, t = T(c["pred"]), T(c["target"])
    for avg, key in (("micro", "acc_micro"), ("macro", "acc_macro"), (None, "acc_none")):
        r = F.multiclass_accuracy(p, t, average=avg, num_classes=5)
        assert r.dtype == torch.float32
        close(r, c[key])
    for avg, key in (("micro", "f1_micro"), ("macro", "f1_macro"), ("weighted", "f1_weighted"), (None, "f1_none")):
        close(F.multiclass_f1_score(p, t, num_classes=5, average=avg), c[key])
    close(F.binary_accuracy(T([0.1, 0.9, 0.6, 0.2]), T([0, 1, 0, 0])), 0.75)
    # float targets are compared as they are (reference accuracy.py:78): 0.5 matches no prediction
    close(F.binary_accuracy(T([0.1, 0.9, 0.2, 0.3]), T([0.5, 1.0, 0.0, float("nan")])), 0.5)
    close(F.binary_f1_score(T([0.1, 0.9, 0.6, 0.2]), T([0, 1, 0, 0])), 2 / 3)
    close(F.multiclass_accuracy(T([[0.9, 0.1, 0], [0.1, 0.2, 0.7], [0.3, 0.4, 0.3]]), T([0, 1, 1]), k=2), 1.0)
    # micro accuracy on scores: a target outside [0, C) is a miss, not an error (reference accuracy.py:269-272)
    close(F.multiclass_accuracy(T([[0.9, 0.1, 0], [0.1, 0.2, 0.7], [0.3, 0.4, 0.3]]), T([0, 7, 1])), 2 / 3)


def test_functional_precision_recall(backend, golden):
    """precision / recall on the class-count kernel, against the reference."""
    from torcheval_b200.metrics import functional as F

    c = golden.case("prec_rec/c7")
    p, t = T(c["pred"]), T(c["target"])
    for avg, tag in (("micro", "micro"), ("macro", "macro"), ("weighted", "weighted"), (None, "none")):
        got = F.multiclass_precision(p, t, num_classes=None if avg == "micro" else 7, average=avg)
        assert got.dtype == torch.float32
        close(got, c[f"precision_{tag}"])
        got = F.multiclass_recall(p, t, num_classes=None if avg == "micro" else 6, average=avg)
        assert got.dtype == torch.float32
        close(got, c[f"recall_{tag}"])
    # a class absent from predictions AND targets: ignored by macro / weighted (the reference's
    # recall raises a shape error here, see functional/classification/recall.py)
    r7 = F.multiclass_recall(p, t, num_classes=7, average="macro")
    close(r7, c["recall_macro"])
    c = golden.case("prec_rec/logits4")
    close(F.multiclass_precision(T(c["logits"]), T(c["target"]), num_classes=4, average="macro"), c["precision_macro"])
    close(F.multiclass_recall(T(c["logits"]), T(c["target"]), num_classes=4, average=None), c["recall_none"])
    c = golden.case("prec_rec/binary")
    close(F.binary_precision(T(c["x"]), T(c["target"])), c["precision_05"])
    close(F.binary_precision(T(c["x"]), T(c["target"]), threshold=0.7), c["precision_07"])
    close(F.binary_recall(T(c["x"]), T(c["target"])), c["recall_05"])
    close(F.binary_recall(T(c["x"]), T(c["target"]), threshold=0.3), c["recall_03"])
    # docstring examples of the reference (precision.py:29-40, recall.py:31-38)
    close(F.binary_precision(T([0, 0, 1, 1]), T([1, 0, 1, 1])), 1.0)
    close(F.binary_precision(T([0, 0.8, 0.6, 0.7]), T([1, 0, 1, 1]), threshold=0.7), 0.5)
    close(F.multiclass_precision(T([0, 2, 1, 3]), T([0, 1, 2, 3]), average=None, num_classes=4), [1.0, 0.0, 0.0, 1.0])
    close(F.binary_recall(T([0, 0, 1, 1]), T([1, 0, 1, 1])), 2 / 3)
    close(F.binary_recall(T([0.2, 0.3]), T([0, 0])), 0.0)  # no positives: NaN -> 0
    with pytest.raises(ValueError, match="`average` was not in the allowed value of"):
        F.multiclass_precision(p, t, average="gamma")
    with pytest.raises(ValueError, match="num_classes should be a positive number when average=macro"):
        F.multiclass_precision(p, t, average="macro")
    with pytest.raises(ValueError, match="`num_classes` should be a positive number when average=None"):
        F.multiclass_recall(p, t, average=None)
    with pytest.raises(ValueError, match="should have the same first dimension"):
        F.multiclass_recall(p[:5], t, num_classes=7, average="macro")
    with pytest.raises(ValueError, match="target should be a one-dimensional tensor"):
        F.binary_precision(torch.rand(2, 2), torch.rand(2, 2))


def test_precision_recall_curves(backend, golden):
    """exact PR curves + recall at fixed precision against the reference."""
    from torcheval_b200.metrics import (
        BinaryPrecisionRecallCurve, BinaryRecallAtFixedPrecision, MulticlassPrecisionRecallCurve,
        MultilabelPrecisionRecallCurve, MultilabelRecallAtFixedPrecision,
    )
    from torcheval_b200.metrics import functional as F

    for name in golden.cases("pr_curve"):
        c = golden.case(f"pr_curve/{name}")
        pr, rc, th = F.binary_precision_recall_curve(T(c["x"]), T(c["y"]))
        assert pr.dtype == torch.float32 and rc.dtype == torch.float32 and th.dtype == torch.float32
        assert pr.shape == c["precision"].shape and th.shape == c["thresholds"].shape, name
        np.testing.assert_array_equal(th.numpy(), c["thresholds"])  # thresholds are the scores themselves
        close(pr, c["precision"], 1e-6)
        close(rc, c["recall"], 1e-6)
        for mp in (0.0, 0.5, 0.9, 1.0):
            r, t = F.binary_recall_at_fixed_precision(T(c["x"]), T(c["y"]), min_precision=mp)
            close(torch.stack([r, t]), c[f"rafp_{int(mp * 100)}"], 1e-6)
    c = golden.case("pr_curve/ties")
    feed = lambda: [(T(x), T(y)) for x, y in zip(_split(c["x"], 3), _split(c["y"], 3))]  # noqa: E731

    def curve_ok(r):
        close(r[0], c["precision"], 1e-6)
        close(r[1], c["recall"], 1e-6)
        np.testing.assert_array_equal(r[2].numpy(), c["thresholds"])

    check_lifecycle(BinaryPrecisionRecallCurve, feed, curve_ok, {"inputs", "targets"})
    check_lifecycle(lambda: BinaryRecallAtFixedPrecision(min_precision=0.5), feed,
                    lambda r: close(torch.stack(list(r)), c["rafp_50"], 1e-6), {"inputs", "targets"})
    # int64 scores and labels other than 0 / 1 (they count as negatives)
    pr, rc, th = F.binary_precision_recall_curve(T([1, 5, 7, 8]), T([0, 3, 1, 1]))
    assert th.dtype == torch.int64 and th.tolist() == [1, 5, 7, 8]
    close(pr, [0.5, 2 / 3, 1.0, 1.0, 1.0], 1e-6)
    close(rc, [1.0, 1.0, 1.0, 0.5, 0.0], 1e-6)
    c = golden.case("pr_curve_multiclass/c4")
    pr, rc, th = F.multiclass_precision_recall_curve(T(c["x"]), T(c["y"]), num_classes=4)
    m = MulticlassPrecisionRecallCurve(num_classes=4).update(T(c["x"][:300]), T(c["y"][:300])).update(T(c["x"][300:]), T(c["y"][300:]))
    mp_, mr_, mt_ = m.compute()
    for k in range(4):
        close(pr[k], c[f"precision_{k}"], 1e-6)
        close(rc[k], c[f"recall_{k}"], 1e-6)
        np.testing.assert_array_equal(th[k].numpy(), c[f"thresholds_{k}"])
        close(mp_[k], c[f"precision_{k}"], 1e-6)
        np.testing.assert_array_equal(mt_[k].numpy(), c[f"thresholds_{k}"])
    c = golden.case("pr_curve_multilabel/l4")
    pr, rc, th = F.multilabel_precision_recall_curve(T(c["x"]), T(c["y"]), num_labels=4)
    for k in range(4):
        close(pr[k], c[f"precision_{k}"], 1e-6)
        close(rc[k], c[f"recall_{k}"], 1e-6)
        np.testing.assert_array_equal(th[k].numpy(), c[f"thresholds_{k}"])
    r, t = F.multilabel_recall_at_fixed_precision(T(c["x"]), T(c["y"]), num_labels=4, min_precision=0.5)
    close(torch.stack([torch.stack(r), torch.stack(t)]), c["rafp_50"], 1e-6)
    ml = MultilabelPrecisionRecallCurve(num_labels=4).update(T(c["x"]), T(c["y"]))
    close(ml.compute()[1][2], c["recall_2"], 1e-6)
    mr = MultilabelRecallAtFixedPrecision(num_labels=4, min_precision=0.5).update(T(c["x"]), T(c["y"]))
    close(torch.stack(mr.compute()[0]), c["rafp_50"][0], 1e-6)
    with pytest.raises(ValueError, match="input should be a one-dimensional tensor"):
        F.binary_precision_recall_curve(torch.rand(2, 2), torch.rand(2, 2))
    with pytest.raises(ValueError, match="Expected min_precision to be a float in the"):
        F.binary_recall_at_fixed_precision(torch.rand(4), torch.ones(4), min_precision=2.0)
    with pytest.raises(ValueError, match="input should have shape of \\(num_sample, num_classes\\)"):
        F.multiclass_precision_recall_curve(torch.rand(4, 3), torch.zeros(4, dtype=torch.int64), num_classes=2)


def test_windowed_binary_auroc(backend, golden):
    """ring-buffer window semantics + merge against the reference."""
    from torcheval_b200.metrics import WindowedBinaryAUROC

    c = golden.case("windowed_auroc/t2")
    xs, ys, ws, cuts = T(c["x"]), T(c["y"]), T(c["w"]), c["cuts"].tolist()
    m = WindowedBinaryAUROC(num_tasks=2, max_num_samples=20)
    for k, (a, b) in enumerate(zip(cuts[:-1], cuts[1:])):
        m.update(xs[:, a:b], ys[:, a:b], ws[:, a:b] if a >= 15 else None)
        got = m.compute()
        assert got.dtype == torch.float64 and got.shape == (2,)
        close(got, c["auroc_steps"][k])
    assert m.total_samples == 57 and m.inputs.shape == (2, 20) and m.inputs.dtype == torch.float32
    m2 = WindowedBinaryAUROC(num_tasks=2, max_num_samples=10).update(xs[:, :9], ys[:, :9])
    m3 = WindowedBinaryAUROC(num_tasks=2, max_num_samples=10).update(xs[:, 9:33], ys[:, 9:33])
    m2.merge_state([m3])
    close(m2.compute(), c["merged"])
    assert m2.max_num_samples == int(c["merged_cap"]) and m2.next_inserted == int(c["merged_next"])
    m1 = WindowedBinaryAUROC(max_num_samples=25).update(T(c["x1"]), T(c["y1"]))
    assert m1.compute().shape == ()
    close(m1.compute(), c["single"])
    sd = m1.state_dict()
    assert set(sd) == {"max_num_samples", "total_samples", "inputs", "targets", "weights"}
    m4 = WindowedBinaryAUROC(max_num_samples=25)
    m4.load_state_dict(sd)
    m4.next_inserted = m1.next_inserted
    close(m4.compute(), c["single"])
    m1.reset()
    assert float(m1.inputs.abs().sum()) == 0  # (int states keep their value across reset, like the reference)
    with pytest.raises(ValueError, match="`max_num_samples` value should be greater than and equal to 1"):
        WindowedBinaryAUROC(max_num_samples=0)


def test_error_messages(backend):
    from torcheval_b200.metrics import functional as F

    x, y = torch.rand(4), torch.randint(0, 2, (4,))
    with pytest.raises(ValueError, match="The `input` and `target` should have the same shape, "):
        F.binary_auroc(torch.rand(4), torch.rand(3))
    with pytest.raises(ValueError, match="`num_tasks = 1`, `input` is expected to be one-dimensional tensor,"):
        F.binary_auroc(torch.rand(4, 5), torch.rand(4, 5))
    with pytest.raises(ValueError, match="The `weight` and `target` should have the same shape, "):
        F.binary_auroc(x, y, weight=torch.rand(3))
    with pytest.raises(ValueError, match="`average` was not in the allowed value of"):
        F.multiclass_auroc(torch.rand(4, 3), torch.randint(0, 3, (4,)), num_classes=3, average="weighted")
    with pytest.raises(ValueError, match="`num_classes` has to be at least 2."):
        F.multiclass_auroc(torch.rand(4, 1), torch.zeros(4, dtype=torch.long), num_classes=1)
    with pytest.raises(ValueError, match="input should have shape of \\(num_sample, num_classes\\)"):
        F.multiclass_auprc(torch.rand(4, 2), torch.zeros(4, dtype=torch.long), num_classes=3)
    with pytest.raises(ValueError, match="The `threshold` should be a sorted tensor."):
        F.binary_binned_auroc(x, y, threshold=torch.tensor([0.1, 0.2, 0.5, 0.7, 0.6]))
    with pytest.raises(ValueError, match="The values in `threshold` should be in the range of \\[0, 1\\]."):
        F.binary_binned_auroc(x, y, threshold=torch.tensor([-0.1, 0.2, 0.5, 0.7]))
    with pytest.raises(ValueError, match="First value in `threshold` should be 0."):
        F.binary_binned_auprc(x, y, threshold=torch.tensor([0.1, 0.2, 1.0]))
    with pytest.raises(ValueError, match="Last value in `threshold` should be 1."):
        F.binary_binned_auprc(x, y, threshold=torch.tensor([0.0, 0.2, 0.9]))
    with pytest.raises(ValueError, match="Unknown memory approach: expected 'vectorized' or 'memory'"):
        F.multiclass_binned_auprc(torch.rand(4, 3), torch.randint(0, 3, (4,)), optimization="cpu")
    with pytest.raises(ValueError, match="Must be at least two classes for confusion matrix"):
        F.multiclass_confusion_matrix(y, y, 1)
    with pytest.raises(ValueError, match="normalize must be one of 'all', 'pred', 'true', or 'none'."):
        F.multiclass_confusion_matrix(y, y, 2, normalize="rows")
    with pytest.raises(ValueError, match="num_classes: 3 must be strictly greater than max class predicted: 4."):
        F.multiclass_confusion_matrix(T([1, 4, 0]), T([0, 1, 2]), 3)
    with pytest.raises(ValueError, match="num_classes: 3 must be strictly greater than max target: 5."):
        F.multiclass_confusion_matrix(T([1, 2, 0]), T([0, 5, 2]), 3)
    with pytest.raises(ValueError, match="num_classes should be a positive number when average=macro"):
        F.multiclass_accuracy(y, y, average="macro")
    with pytest.raises(TypeError, match="Expected `k` to be an integer"):
        F.multiclass_accuracy(y, y, k=1.5)
    with pytest.raises(ValueError, match="input should be a one-dimensional tensor for binary f1 score"):
        F.binary_f1_score(torch.rand(2, 2), torch.rand(2, 2))


# ------------------------------------------------------------------------------ classes
def _split(a, k):
    return np.array_split(a, k, axis=-1 if a.ndim == 1 or a.shape[0] < a.shape[-1] else 0)


def check_lifecycle(make, feed, expect_fn, state_names):
    """update in 4 batches == one shot; 2-way merge == one shot; merge leaves its arguments alone;
    pickle / deepcopy / hash / state_dict round trips; reset (reference tester :148-290)."""
    m = make()
    assert set(m._state_name_to_default) == state_names
    pickle.loads(pickle.dumps(m))
    hash(m)
    batches = feed()
    for b in batches:
        m.update(*b)
    r1 = m.compute()
    r2 = m.compute()  # idempotent
    expect_fn(r1)
    expect_fn(r2)
    a, b = make(), make()
    for i, bt in enumerate(batches):
        (a if i % 2 == 0 else b).update(*bt)
    b_state = copy.deepcopy(b.state_dict())
    a.merge_state([b])
    expect_fn(a.compute())
    for k, v in b.state_dict().items():  # untouched
        if isinstance(v, torch.Tensor):
            assert torch.equal(v, b_state[k])
        else:
            assert all(torch.equal(x, y) for x, y in zip(v, b_state[k]))
    a.update(*batches[0])  # still usable after a merge
    c = make()
    c.load_state_dict(m.state_dict())
    expect_fn(c.compute())
    expect_fn(pickle.loads(pickle.dumps(m)).compute())
    m.reset()
    for name, default in m._state_name_to_default.items():
        v = getattr(m, name)
        assert (v == default).all() if isinstance(v, torch.Tensor) else v == []
    empty = make()
    empty.merge_state([make()])  # merging empties is fine


def test_class_binary_auroc(backend, golden):
    from torcheval_b200.metrics import BinaryAUROC

    c = golden.case("binary/ties64")
    xs, ys = _split(c["x"], 4), _split(c["y"], 4)
    check_lifecycle(
        BinaryAUROC, lambda: [(T(x), T(y)) for x, y in zip(xs, ys)],
        lambda r: (close(r, c["auroc_f64"]), r.dtype == torch.float64), {"inputs", "targets", "weights"},
    )
    m = BinaryAUROC()
    m.update(T(c["x"]), T(c["y"]))
    assert m.weights[0].numel() == 0  # no 8 B/sample of ones for unweighted updates
    w = golden.case("binary_weighted/w")
    m = BinaryAUROC()
    half = w["x"].size // 2
    m.update(T(w["x"][:half]), T(w["y"][:half]), T(w["w"][:half]))
    m.update(T(w["x"][half:]), T(w["y"][half:]), T(w["w"][half:]))
    close(m.compute(), w["auroc"])
    m2 = BinaryAUROC()  # mixing weighted and unweighted updates
    m2.update(T(w["x"][:half]), T(w["y"][:half]), T(w["w"][:half]))
    m2.update(T(w["x"][half:]), T(w["y"][half:]))
    ww = w["w"].copy()
    ww[half:] = 1.0
    import oracle

    close(m2.compute(), oracle.binary_auroc(w["x"], w["y"], ww)[0])
    m2._prepare_for_merge_state()
    close(m2.compute(), oracle.binary_auroc(w["x"], w["y"], ww)[0])
    t3 = golden.case("binary_multitask/t3")
    m = BinaryAUROC(num_tasks=3)
    m.update(T(t3["x"][:, :500]), T(t3["y"][:, :500])).update(T(t3["x"][:, 500:]), T(t3["y"][:, 500:]))
    close(m.compute(), t3["auroc"])
    with pytest.raises(ValueError, match="`num_tasks` value should be greater than or equal to 1"):
        BinaryAUROC(num_tasks=0)
    with pytest.raises(ValueError, match="`use_fbgemm` is enabled but `fbgemm_gpu` is not found"):
        BinaryAUROC(use_fbgemm=True)
    with pytest.raises(Exception):
        BinaryAUROC().compute()  # like the reference: cat of an empty list


def test_class_exact_others(backend, golden):
    from torcheval_b200.metrics import BinaryAUPRC, MulticlassAUPRC, MulticlassAUROC, MultilabelAUPRC

    c = golden.case("binary/uniform")
    check_lifecycle(
        BinaryAUPRC, lambda: [(T(x), T(y)) for x, y in zip(_split(c["x"], 4), _split(c["y"], 4))],
        lambda r: close(r, c["auprc"], 1e-4), {"inputs", "targets"},
    )
    mc = golden.case("multiclass/c5")
    rows = np.array_split(np.arange(mc["x"].shape[0]), 4)
    feed = lambda: [(T(mc["x"][r]), T(mc["y"][r])) for r in rows]  # noqa: E731
    check_lifecycle(lambda: MulticlassAUROC(num_classes=5, average=None), feed, lambda r: close(r, mc["auroc"], 1e-4), {"inputs", "targets"})
    check_lifecycle(lambda: MulticlassAUPRC(num_classes=5), feed, lambda r: close(r, mc["auprc_macro"], 1e-4), {"inputs", "targets"})
    ml = golden.case("multilabel/c5")
    check_lifecycle(
        lambda: MultilabelAUPRC(num_labels=5, average=None), lambda: [(T(ml["x"][r]), T(ml["y"][r])) for r in rows],
        lambda r: close(r, ml["auprc"], 1e-4), {"inputs", "targets"},
    )


def test_class_binned(backend, golden):
    from torcheval_b200.metrics import (
        BinaryBinnedAUPRC, BinaryBinnedAUROC, BinaryBinnedPrecisionRecallCurve, MulticlassBinnedAUPRC,
        MulticlassBinnedPrecisionRecallCurve, MultilabelBinnedAUPRC,
    )

    c = golden.case("binned/lin200")
    feed = lambda: [(T(x), T(y)) for x, y in zip(_split(c["x"], 4), _split(c["y"], 4))]  # noqa: E731
    check_lifecycle(lambda: BinaryBinnedAUROC(threshold=200), feed, lambda r: close(r[0], c["auroc"]), {"num_tp", "num_fp"})
    check_lifecycle(lambda: BinaryBinnedAUPRC(threshold=200), feed, lambda r: close(r[0], c["auprc"], 1e-4), {"num_tp", "num_fp", "num_fn"})

    def pr_ok(r):
        np.testing.assert_array_equal(r[0].shape, (201,))
        tp, fp = c["tp"].astype(np.float64), c["fp"].astype(np.float64)
        want = np.where(tp + fp > 0, tp / np.maximum(tp + fp, 1), 1.0)
        close(r[0][:-1], want, 1e-6)

    check_lifecycle(lambda: BinaryBinnedPrecisionRecallCurve(threshold=200), feed, pr_ok, {"num_tp", "num_fp", "num_fn"})
    m = BinaryBinnedAUROC(threshold=200)
    assert m.num_tp.dtype == torch.int64 and m.num_tp.shape == (1, 200)
    t2 = golden.case("binned_multitask/t2")
    m = BinaryBinnedAUROC(num_tasks=2, threshold=50)
    m.update(T(t2["x"][:, :300]), T(t2["y"][:, :300])).update(T(t2["x"][:, 300:]), T(t2["y"][:, 300:]))
    close(m.compute()[0], t2["auroc"])
    mc = golden.case("binned_multiclass/c6")
    rows = np.array_split(np.arange(300), 3)
    feed = lambda: [(T(mc["x"][r]), T(mc["y"][r])) for r in rows]  # noqa: E731
    check_lifecycle(
        lambda: MulticlassBinnedAUPRC(num_classes=6, threshold=T(mc["thr"]), average=None), feed,
        lambda r: close(r, mc["auprc"], 1e-4), {"num_tp", "num_fp", "num_fn"},
    )
    m = MulticlassBinnedPrecisionRecallCurve(num_classes=6, threshold=33)
    for b in feed():
        m.update(*b)
    np.testing.assert_array_equal(m.num_tp.numpy(), mc["tp"])
    ml = golden.case("binned_multilabel/l6")
    check_lifecycle(
        lambda: MultilabelBinnedAUPRC(num_labels=6, threshold=T(ml["thr"]), average=None),
        lambda: [(T(ml["x"][r]), T(ml["y"][r])) for r in rows], lambda r: close(r, ml["auprc"], 1e-4),
        {"num_tp", "num_fp", "num_fn"},
    )


def test_class_counting(backend, golden):
    from torcheval_b200.metrics import (
        BinaryAccuracy, BinaryConfusionMatrix, BinaryF1Score, MulticlassAccuracy, MulticlassConfusionMatrix, MulticlassF1Score,
    )

    c = golden.case("cm/c10")
    feed = lambda: [(T(p), T(t)) for p, t in zip(_split(c["pred"], 4), _split(c["target"], 4))]  # noqa: E731

    def cm_ok(r):
        assert r.dtype == torch.float32  # class API returns float32 like the reference
        np.testing.assert_array_equal(r.numpy(), c["cm"].astype(np.float32))

    check_lifecycle(lambda: MulticlassConfusionMatrix(10), feed, cm_ok, {"confusion_matrix"})
    m = MulticlassConfusionMatrix(10)
    for b in feed():
        m.update(*b)
    close(m.normalized("true").sum(dim=1), np.ones(10))
    before = m.confusion_matrix.clone()
    with pytest.raises(ValueError, match="too large for the number of classes"):
        m.update(T([3, 11]), T([1, 2]))
    assert torch.equal(m.confusion_matrix, before)  # a rejected update leaves the state alone
    b = golden.case("cm_binary/t05")
    mb = BinaryConfusionMatrix()
    mb.update(T(b["x"]), T(b["target"]))
    np.testing.assert_array_equal(mb.compute().numpy(), b["cm"].astype(np.float32))
    a = golden.case("acc_f1/c5")
    feed = lambda: [(T(p), T(t)) for p, t in zip(_split(a["pred"], 3), _split(a["target"], 3))]  # noqa: E731
    check_lifecycle(MulticlassAccuracy, feed, lambda r: close(r, a["acc_micro"]), {"num_correct", "num_total"})
    check_lifecycle(lambda: MulticlassAccuracy(average="macro", num_classes=5), feed, lambda r: close(r, a["acc_macro"]), {"num_correct", "num_total"})
    check_lifecycle(lambda: MulticlassF1Score(num_classes=5, average="weighted"), feed, lambda r: close(r, a["f1_weighted"]), {"num_tp", "num_label", "num_prediction"})
    check_lifecycle(MulticlassF1Score, feed, lambda r: close(r, a["f1_micro"]), {"num_tp", "num_label", "num_prediction"})
    ba = BinaryAccuracy().update(T([0.1, 0.9, 0.6, 0.2]), T([0, 1, 0, 0]))
    close(ba.compute(), 0.75)
    from torcheval_b200.metrics import BinaryPrecision, BinaryRecall, MulticlassPrecision, MulticlassRecall

    pr = golden.case("prec_rec/c7")
    feed = lambda: [(T(p), T(t)) for p, t in zip(_split(pr["pred"], 3), _split(pr["target"], 3))]  # noqa: E731
    check_lifecycle(MulticlassPrecision, feed, lambda r: close(r, pr["precision_micro"]), {"num_tp", "num_fp", "num_label"})
    check_lifecycle(lambda: MulticlassPrecision(num_classes=7, average="weighted"), feed, lambda r: close(r, pr["precision_weighted"]), {"num_tp", "num_fp", "num_label"})
    check_lifecycle(lambda: MulticlassRecall(num_classes=6, average="macro"), feed, lambda r: close(r, pr["recall_macro"]), {"num_tp", "num_labels", "num_predictions"})
    check_lifecycle(lambda: MulticlassRecall(num_classes=6, average=None), feed, lambda r: close(r, pr["recall_none"]), {"num_tp", "num_labels", "num_predictions"})
    pb = golden.case("prec_rec/binary")
    feed = lambda: [(T(x), T(t)) for x, t in zip(_split(pb["x"], 4), _split(pb["target"], 4))]  # noqa: E731
    check_lifecycle(lambda: BinaryPrecision(threshold=0.7), feed, lambda r: close(r, pb["precision_07"]), {"num_tp", "num_fp", "num_label"})
    check_lifecycle(lambda: BinaryRecall(threshold=0.3), feed, lambda r: close(r, pb["recall_03"]), {"num_tp", "num_true_labels"})
    bf = BinaryF1Score().update(T([0.1, 0.9, 0.6, 0.2]), T([0, 1, 0, 0]))
    close(bf.compute(), 2 / 3)
    assert torch.isnan(MulticlassAccuracy().compute())


def test_metric_base_and_toolkit_single_process(backend, golden):
    from torcheval_b200.metrics import BinaryAUROC, Metric, MulticlassConfusionMatrix
    from torcheval_b200.metrics import toolkit

    with pytest.raises(TypeError, match="The value of state variable must be"):
        MulticlassConfusionMatrix(3)._add_state("bad", "a string")
    m = MulticlassConfusionMatrix(3)
    with pytest.raises(RuntimeError, match="Error\\(s\\) in loading state_dict for MulticlassConfusionMatrix"):
        m.load_state_dict({"nope": torch.zeros(1)})
    m.load_state_dict({"nope": torch.zeros(1)}, strict=False)
    m.load_state_dict({"confusion_matrix": torch.ones(3, 3)})  # a reference float32 checkpoint loads
    m.update(T([0, 1]), T([0, 1]))
    assert m.confusion_matrix.dtype == torch.float32 and m.confusion_matrix[0, 0] == 2
    assert m.to("cpu").device == torch.device("cpu")
    assert issubclass(BinaryAUROC, Metric)
    c = golden.case("binary/uniform")
    a = BinaryAUROC().update(T(c["x"]), T(c["y"]))
    close(toolkit.sync_and_compute(a), c["auroc_f64"])  # world size 1: plain compute
    assert toolkit.get_synced_metric(a) is a
    out = toolkit.sync_and_compute_collection({"auroc": a, "cm": m})
    assert set(out) == {"auroc", "cm"}
    cl = toolkit.clone_metric(a)
    assert cl is not a and torch.equal(cl.inputs[0], a.inputs[0])
    toolkit.reset_metrics([a])
    assert a.inputs == []
    d = toolkit.classwise_converter(torch.tensor([1.0, 2.0]), "acc", ["a", "b"])
    assert set(d) == {"acc_a", "acc_b"}
    with pytest.raises(ValueError, match="Number of labels 1 must be equal to the number of classes 2!"):
        toolkit.classwise_converter(torch.tensor([1.0, 2.0]), "acc", ["a"])
