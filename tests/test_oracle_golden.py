"""Pin the oracle (oracle/te_oracle.c) to the reference: fixtures produced by running torcheval
itself (oracle/gen_golden.py) and the hard-coded goldens of the reference's own unit tests."""
import numpy as np
import pytest

import oracle

RTOL64 = 1e-9   # vs the reference's float64 path
RTOL32 = 2e-5   # vs outputs the reference itself rounds to / accumulates in float32


def close(a, b, rtol):
    np.testing.assert_allclose(np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64), rtol=rtol, atol=rtol * 1e-2)


def test_binary_auroc_auprc_vs_reference(golden):
    for name in golden.cases("binary"):
        c = golden.case(f"binary/{name}")
        v, tp, fp = oracle.binary_auroc(c["x"], c["y"])
        close(v, c["auroc_f64"], RTOL64)
        close(v, c["auroc_fn"], RTOL32)
        assert tp == c["y"].sum() and fp == (1 - c["y"]).sum()
        ap = oracle.binary_auprc(c["x"], c["y"])
        close(ap, c["auprc"], RTOL32)
        if c["y"].sum() == 0:  # the reference returns -0.0 here
            assert np.signbit(ap) and np.signbit(c["auprc"])


def test_binary_auroc_weighted_soft(golden):
    for name in golden.cases("binary_weighted"):
        c = golden.case(f"binary_weighted/{name}")
        v, _, _ = oracle.binary_auroc(c["x"], c["y"], c["w"])
        close(v, c["auroc"], RTOL64)


def test_binary_multitask(golden):
    c = golden.case("binary_multitask/t3")
    for k in range(3):
        close(oracle.binary_auroc(c["x"][k], c["y"][k])[0], c["auroc"][k], RTOL64)
        close(oracle.binary_auprc(c["x"][k], c["y"][k]), c["auprc"][k], RTOL32)


def test_round2_fixtures(golden):
    """Fixtures appended in round 2: multi-task weighted AUROC; 64-bit scores (the C oracle takes
    float32 scores, so it is handed their dense ranks: same order, same ties)."""
    c = golden.case("binary_multitask_weighted/t4")
    for k in range(4):
        close(oracle.binary_auroc(c["x"][k], c["y"][k], c["w"][k])[0], c["auroc"][k], RTOL64)
    for name in ("close", "int64"):
        c = golden.case(f"binary_f64/{name}")
        ranks = np.unique(c["x"], return_inverse=True)[1].astype(np.float32)
        close(oracle.binary_auroc(ranks, c["y"])[0], c["auroc"], RTOL64)
        close(oracle.binary_auprc(ranks, c["y"]), c["auprc"], RTOL32)


def test_multiclass_multilabel(golden):
    for name in golden.cases("multiclass"):
        c = golden.case(f"multiclass/{name}")
        a = oracle.multiclass_auroc(c["x"], c["y"])
        close(a, c["auroc"], RTOL32)
        close(a.mean(), c["auroc_macro"], RTOL32)
        p = oracle.multiclass_auprc(c["x"], c["y"])
        close(p, c["auprc"], RTOL32)
        close(p.mean(), c["auprc_macro"], RTOL32)
    for name in golden.cases("multilabel"):
        c = golden.case(f"multilabel/{name}")
        p = oracle.multilabel_auprc(c["x"], c["y"])
        close(p, c["auprc"], RTOL32)


def test_binned_counts_bit_exact(golden):
    for name in golden.cases("binned"):
        c = golden.case(f"binned/{name}")
        tp, fp, fn = oracle.binned_update(c["x"], c["y"], c["thr"])
        np.testing.assert_array_equal(tp, c["tp"])
        np.testing.assert_array_equal(fp, c["fp"])
        np.testing.assert_array_equal(fn, c["fn"])
        if "auroc" in c:
            close(oracle.binned_auroc_from_counts(tp, fp), c["auroc"][0], 1e-6)
            close(oracle.binned_auroc_direct(c["x"], c["y"], c["thr"]), c["auroc"][0], 1e-6)
        if "auprc" in c:
            close(oracle.binned_auprc_from_counts(tp, fp, fn), c["auprc"], RTOL32)


def test_binned_multitask(golden):
    c = golden.case("binned_multitask/t2")
    for k in range(2):
        tp, fp, fn = oracle.binned_update(c["x"][k], c["y"][k], c["thr"])
        close(oracle.binned_auroc_from_counts(tp, fp), c["auroc"][k], 1e-6)
        close(oracle.binned_auprc_from_counts(tp, fp, fn), c["auprc"][k], RTOL32)


def test_confusion_matrix_bit_exact(golden):
    for name in golden.cases("cm"):
        c = golden.case(f"cm/{name}")
        cm, bad, _ = oracle.confusion_matrix(c["pred"], c["target"], c["cm"].shape[0])
        assert not bad
        np.testing.assert_array_equal(cm, c["cm"])
    for name in golden.cases("cm_logits"):
        c = golden.case(f"cm_logits/{name}")
        cm, bad, _ = oracle.confusion_matrix(oracle.argmax_rows(c["logits"]), c["target"], c["cm"].shape[0])
        np.testing.assert_array_equal(cm, c["cm"])
    c = golden.case("cm_binary/t05")
    cm, _, _ = oracle.confusion_matrix((c["x"] >= 0.5).astype(np.int64), c["target"], 2)
    np.testing.assert_array_equal(cm, c["cm"])


def test_confusion_matrix_range_report():
    cm, bad, rng = oracle.confusion_matrix([0, 5, 1], [1, 1, -2], 3)
    assert bad and list(rng) == [0, 5, -2, 1]
    assert cm.sum() == 1 and cm[1, 0] == 1


def test_reference_unit_test_goldens():
    """Hard-coded values from the reference's own tests and docstrings (cited); random inputs are
    re-derived with the same torch.manual_seed streams (reproducible under torch 2.11)."""
    import torch

    # functional/classification/auprc.py docstring + tests/.../functional/classification/test_auprc.py:76-107
    close(oracle.binary_auprc([0.1, 0.5, 0.7, 0.8], [1, 0, 1, 1]), 0.9167, 1e-4)
    close(oracle.binary_auroc([0.1, 0.5, 0.7, 0.8], [1, 0, 1, 1])[0], 2.0 / 3.0, 1e-12)
    close(oracle.binary_auroc([1, 1, 1, 0], [1, 1, 1, 0])[0], 1.0, 1e-12)
    close(oracle.binary_auroc([1, 1, 1, 0], [1, 0, 1, 0])[0], 0.75, 1e-12)
    # tests/metrics/functional/classification/test_auroc.py:137-167 (average=None golden)
    inp = [[0.16, 0.04, 0.8], [0.1, 0.7, 0.2], [0.16, 0.8, 0.04], [0.16, 0.04, 0.8]]
    close(oracle.multiclass_auroc(inp, [0, 0, 1, 2]), [0.25, 1.0, 5 / 6], 1e-12)
    # tests/metrics/classification/test_auroc.py:99-108: seed 123, 2 tasks
    torch.manual_seed(123)
    x = torch.rand(8, 2, 16)
    y = torch.randint(high=2, size=(8, 2, 16))
    xs = x.permute(1, 0, 2).reshape(2, -1).numpy()
    ys = y.permute(1, 0, 2).reshape(2, -1).numpy()
    close([oracle.binary_auroc(xs[k], ys[k])[0] for k in range(2)], [0.549048678033111, 0.512218963831867], 1e-12)
    # tests/metrics/classification/test_binned_auroc.py:36-71: seed 123, threshold=5
    thr5 = np.linspace(0, 1, 5, dtype=np.float32)
    torch.manual_seed(123)
    xi = torch.randint(high=2, size=(8, 16)).reshape(-1).numpy()
    yi = torch.randint(high=2, size=(8, 16)).reshape(-1).numpy()
    tp, fp, _ = oracle.binned_update(xi, yi, thr5)
    close(oracle.binned_auroc_from_counts(tp, fp), 0.5078144078144078, 1e-7)
    torch.manual_seed(123)
    x = torch.rand(8, 2, 16)
    y = torch.randint(high=2, size=(8, 2, 16))
    xs = x.permute(1, 0, 2).reshape(2, -1).numpy()
    ys = y.permute(1, 0, 2).reshape(2, -1).numpy()
    got = []
    for k in range(2):
        tp, fp, _ = oracle.binned_update(xs[k], ys[k], thr5)
        got.append(oracle.binned_auroc_from_counts(tp, fp))
    close(got, [0.5406473931307141, 0.49572336265884653], 1e-7)
    # functional/classification/binned_precision_recall_curve.py:46-63 (docstring) and
    # tests/.../functional/classification/test_binned_precision_recall_curve.py:39-48
    tp, fp, fn = oracle.binned_update([0.2, 0.8, 0.5, 0.9], [0, 1, 0, 1], thr5)
    prec = np.where(tp + fp > 0, tp / np.maximum(tp + fp, 1), 1.0)
    np.testing.assert_allclose(prec, [0.5, 2 / 3, 2 / 3, 1.0, 1.0], rtol=1e-6)
    np.testing.assert_allclose(tp / (tp + fn), [1, 1, 1, 1, 0])
    tp, fp, fn = oracle.binned_update([0.2, 0.3, 0.4, 0.5], [0, 0, 1, 1], [0.0, 0.25, 0.75, 1.0])
    prec = np.where(tp + fp > 0, tp / np.maximum(tp + fp, 1), 1.0)
    np.testing.assert_allclose(prec, [0.5, 2 / 3, 1.0, 1.0], rtol=1e-6)
    np.testing.assert_allclose(tp / (tp + fn), [1, 1, 0, 0])


def test_level_hist_closed_form_matches_sort():
    rng = np.random.default_rng(5)
    G = 37
    lev = rng.integers(0, G, 5000)
    y = rng.integers(0, 2, 5000)
    neg = np.bincount(lev[y == 0], minlength=G)
    pos = np.bincount(lev[y == 1], minlength=G)
    a = oracle.auroc_from_level_hist(neg, pos)
    b = oracle.binary_auroc(lev.astype(np.float32) / G, y)[0]
    close(a, b, 1e-12)
