"""Parity oracle (TEST INFRASTRUCTURE ONLY — see the header of te_oracle.c).

ctypes wrappers over the plain-C restatement.  Importers: tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  Nothing under torcheval_b200/ may import this.
"""
from __future__ import annotations

import ctypes as C
import hashlib
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_SRC = _HERE / "te_oracle.c"
_OUT = _HERE / "_build"
_LIB = _OUT / "libte_oracle.so"


def build(force: bool = False) -> Path:
    """gcc -O2 -shared te_oracle.c (a second or two); rebuilt when the source changes."""
    digest = hashlib.sha256(_SRC.read_bytes()).hexdigest()
    stamp = _OUT / "digest"
    if not force and _LIB.exists() and stamp.exists() and stamp.read_text() == digest:
        return _LIB
    _OUT.mkdir(exist_ok=True)
    subprocess.run(
        ["gcc", "-O2", "-std=c11", "-fPIC", "-shared", "-fvisibility=hidden", "-o", str(_LIB), str(_SRC), "-lm"],
        check=True,
    )
    stamp.write_text(digest)
    return _LIB


_lib = None


def _l() -> C.CDLL:
    global _lib
    if _lib is None:
        _lib = C.CDLL(str(build()))
        _lib.oracle_binned_auroc_direct.restype = C.c_double
        _lib.oracle_binned_auroc_from_counts.restype = C.c_double
        _lib.oracle_binned_auprc_from_counts.restype = C.c_double
        _lib.oracle_binary_auroc.restype = C.c_double
        _lib.oracle_binary_auprc.restype = C.c_double
        _lib.oracle_auroc_from_level_hist.restype = C.c_double
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def confusion_matrix(pred, target, num_classes: int):
    """(cm int64 [C, C], bad flag, range[4]) — confusion_matrix.py:219-234."""
    pred, target = _i64(pred), _i64(target)
    cm = np.zeros((num_classes, num_classes), dtype=np.int64)
    rng = np.zeros(4, dtype=np.int64)
    bad = _l().oracle_confusion_matrix(_p(pred), _p(target), C.c_int64(pred.size), C.c_int32(num_classes), _p(cm), _p(rng))
    return cm, bool(bad), rng


def argmax_rows(logits):
    logits = _f32(logits)
    n, c = logits.shape
    out = np.zeros(n, dtype=np.int64)
    _l().oracle_argmax_rows(_p(logits), C.c_int64(c), C.c_int64(n), C.c_int32(c), _p(out))
    return out


def binned_hist(x, target, threshold):
    """(T+1, 2) int64 histogram of (bucket, label)."""
    x, target, threshold = _f32(x), _i64(target), _f32(threshold)
    T = threshold.size
    hist = np.zeros((T + 1, 2), dtype=np.int64)
    _l().oracle_binned_hist(_p(x), _p(target), C.c_int64(x.size), _p(threshold), C.c_int32(T), _p(hist))
    return hist


def binned_counts(hist):
    """(num_tp, num_fp, num_fn), each (T,) int64 — binned_precision_recall_curve.py:102-110."""
    hist = _i64(hist)
    T = hist.shape[0] - 1
    tp, fp, fn = (np.zeros(T, dtype=np.int64) for _ in range(3))
    _l().oracle_binned_counts(_p(hist), C.c_int32(T), _p(tp), _p(fp), _p(fn))
    return tp, fp, fn


def binned_update(x, target, threshold):
    return binned_counts(binned_hist(x, target, threshold))


def binned_auroc_direct(x, target, threshold) -> float:
    x, target, threshold = _f32(x), _i64(target), _f32(threshold)
    return _l().oracle_binned_auroc_direct(_p(x), _p(target), C.c_int64(x.size), _p(threshold), C.c_int32(threshold.size))


def binned_auroc_from_counts(tp, fp) -> float:
    tp, fp = _i64(tp), _i64(fp)
    return _l().oracle_binned_auroc_from_counts(_p(tp), _p(fp), C.c_int32(tp.size))


def binned_auprc_from_counts(tp, fp, fn) -> float:
    tp, fp, fn = _i64(tp), _i64(fp), _i64(fn)
    return _l().oracle_binned_auprc_from_counts(_p(tp), _p(fp), _p(fn), C.c_int32(tp.size))


def binary_auroc(score, target, weight=None):
    """(auroc, TP_total, FP_total) — auroc.py:115-152; target may be soft, weight optional."""
    score, target = _f32(score), _f64(target)
    weight = None if weight is None else _f64(weight)
    tp, fp = C.c_double(), C.c_double()
    v = _l().oracle_binary_auroc(_p(score), _p(target), _p(weight), C.c_int64(score.size), C.byref(tp), C.byref(fp))
    return v, tp.value, fp.value


def binary_auprc(score, target, pos_label: int = 1) -> float:
    score, target = _f32(score), _i64(target)
    return _l().oracle_binary_auprc(_p(score), _p(target), C.c_int64(pos_label), C.c_int64(score.size))


def multiclass_auroc(inp, target):
    inp, target = _f32(inp), _i64(target)
    n, c = inp.shape
    out = np.zeros(c, dtype=np.float64)
    _l().oracle_multiclass_auroc(_p(inp), _p(target), C.c_int64(n), C.c_int32(c), _p(out))
    return out


def multiclass_auprc(inp, target):
    inp, target = _f32(inp), _i64(target)
    n, c = inp.shape
    out = np.zeros(c, dtype=np.float64)
    _l().oracle_multiclass_auprc(_p(inp), _p(target), C.c_int64(n), C.c_int32(c), _p(out))
    return out


def multilabel_auprc(inp, target):
    inp, target = _f32(inp), _i64(target)
    n, c = inp.shape
    out = np.zeros(c, dtype=np.float64)
    _l().oracle_multilabel_auprc(_p(inp), _p(target), C.c_int64(n), C.c_int32(c), _p(out))
    return out


def auroc_from_level_hist(neg, pos) -> float:
    neg, pos = _i64(neg), _i64(pos)
    return _l().oracle_auroc_from_level_hist(_p(neg), _p(pos), C.c_int64(neg.size))
