"""TEST-ONLY stand-in for the CUDA kernels: every function of ``torcheval_b200.ops`` re-implemented
on CPU tensors with the oracle (oracle/te_oracle.c) and numpy, so that the host-side logic —
argument checks, state handling, merge/sync, the distributed sample-sort protocol — can be tested
without a GPU.  Installed by monkeypatching (``install()``); the product never imports this and
has no CPU path of its own."""
from __future__ import annotations

import numpy as np
import torch

import oracle

I64_MAX = np.iinfo(np.int64).max
I64_MIN = np.iinfo(np.int64).min


def _np(t):
    return t.detach().cpu().numpy()


def _range_report(p, t, C):
    bad_p = p[(p < 0) | (p >= C)]
    bad_t = t[(t < 0) | (t >= C)]
    r = [I64_MAX, I64_MIN, I64_MAX, I64_MIN]
    if bad_p.size:
        r[0], r[1] = int(bad_p.min()), int(bad_p.max())
    if bad_t.size:
        r[2], r[3] = int(bad_t.min()), int(bad_t.max())
    return torch.tensor(r, dtype=torch.int64)


def cm_update_(cm, pred, target, num_classes, variant=0):
    p, t = _np(pred), _np(target)
    add, _, _ = oracle.confusion_matrix(p, t, num_classes)
    cm += torch.from_numpy(add)
    return _range_report(p, t, num_classes)


def cm_update_logits_(cm, logits, target, num_classes):
    return cm_update_(cm, torch.from_numpy(oracle.argmax_rows(_np(logits))), target, num_classes)


def class_counts_(counts3, input, target, num_classes):
    p = oracle.argmax_rows(_np(input)) if input.dim() == 2 else _np(input)
    t = _np(target)
    ok = (p >= 0) & (p < num_classes) & (t >= 0) & (t < num_classes)
    po, to = p[ok], t[ok]
    counts3[0] += torch.from_numpy(np.bincount(to[po == to], minlength=num_classes))
    counts3[1] += torch.from_numpy(np.bincount(to, minlength=num_classes))
    counts3[2] += torch.from_numpy(np.bincount(po, minlength=num_classes))
    return _range_report(p, t, num_classes)


def match_count(pred, target):
    return (pred == target).sum().to(torch.int64)


def binned_hist_update_(hist, x, target, threshold):
    xs, ts = _np(x), _np(target)
    thr = _np(threshold)
    if xs.ndim == 1:
        xs, ts = xs[None], ts[None]
    for k in range(xs.shape[0]):
        keep = ~np.isnan(xs[k])
        h = oracle.binned_hist(xs[k][keep], ts[k][keep], thr)
        h[0, 0] += int(((~keep) & (ts[k] == 0)).sum())
        h[0, 1] += int(((~keep) & (ts[k] != 0)).sum())
        hist.view(-1, thr.size + 1, 2)[k] += torch.from_numpy(h)


def binned_hist2d_update_(hist, x, target, threshold, multilabel):
    xs, ts, thr = _np(x), _np(target), _np(threshold)
    for c in range(xs.shape[1]):
        lab = ts[:, c] != 0 if multilabel else ts == c
        hist[c] += torch.from_numpy(oracle.binned_hist(xs[:, c], lab.astype(np.int64), thr))


def binned_counts(hist, T, want_fn=True):
    h = _np(hist).reshape(-1, T + 1, 2)
    tp, fp, fn = zip(*(oracle.binned_counts(h[k]) for k in range(h.shape[0])))
    f = lambda a: torch.from_numpy(np.stack(a))  # noqa: E731
    return f(tp), f(fp), (f(fn) if want_fn else None)


def binned_integrate(tp, fp, fn, mode):
    out = []
    for k in range(tp.shape[0]):
        if mode == 0:
            out.append(oracle.binned_auroc_from_counts(_np(tp[k]), _np(fp[k])))
        else:
            out.append(oracle.binned_auprc_from_counts(_np(tp[k]), _np(fp[k]), _np(fn[k])))
    return torch.tensor(out, dtype=torch.float64)


def _flags(score, target):
    x, t = _np(score), _np(target).astype(np.float64)
    f = 0
    if ((t != 0) & (t != 1)).any():
        f |= 1
    if (np.signbit(x) & (x != 0)).any():
        f |= 2
    return f


def _exact_2u(x, y):
    """2U as an exact Python int from the (score, label) pairs, ties included."""
    order = np.argsort(x, kind="stable")
    xs, ys = x[order], y[order].astype(np.int64)
    _, start, cnt = np.unique(xs, return_index=True, return_counts=True)
    neg = np.add.reduceat(1 - ys, start)
    pos = np.add.reduceat(ys, start)
    below = np.concatenate([[0], np.cumsum(neg)[:-1]])
    return int(np.sum(pos.astype(object) * (2 * below.astype(object) + neg.astype(object))))


def binary_auroc_packed(score, target, variant=0):
    f = _flags(score, target)
    out = torch.zeros(8, dtype=torch.float64)
    out[4] = f
    if f:
        return out
    x, y = _np(score), _np(target).astype(np.int64)
    v, tp, fp = oracle.binary_auroc(x, y)
    u2 = _exact_2u(x, y) if x.size else 0
    out[:4] = torch.tensor([v, tp, fp, float(u2)], dtype=torch.float64)
    out[5], out[6], out[7] = float(u2 >> 32), float(u2 & 0xFFFFFFFF), float(x.size)
    return out


def binary_auprc_packed(score, target, variant=0):
    f = _flags(score, target)
    out = torch.zeros(8, dtype=torch.float64)
    out[4] = f
    if not f:
        y = _np(target).astype(np.int64)
        out[0] = oracle.binary_auprc(_np(score), y)
        out[1], out[2] = y.sum(), (1 - y).sum()
    return out


def binary_auroc_general(score, target, weight=None):
    v, tp, fp = oracle.binary_auroc(_np(score), _np(target).astype(np.float64), None if weight is None else _np(weight))
    out = torch.zeros(8, dtype=torch.float64)
    out[:3] = torch.tensor([v, tp, fp], dtype=torch.float64)
    return out


def binary_auroc_general_batched(score, target, weight=None):
    rows = [binary_auroc_general(score[k], target[k], None if weight is None else weight[k]) for k in range(score.shape[0])]
    return torch.stack(rows) if rows else torch.zeros((0, 8), dtype=torch.float64)


def binary_auprc_general(score, target):
    t = _np(target)
    out = torch.zeros(8, dtype=torch.float64)
    out[0] = oracle.binary_auprc(_np(score), (t == 1).astype(np.int64))
    return out


def ovr_rank(input, target, multilabel, mode):
    xs, ts = _np(input), _np(target)
    rows = []
    for c in range(xs.shape[1]):
        lab = torch.from_numpy(((ts[:, c] == 1) if multilabel else (ts == c)).astype(np.int64))
        col = torch.from_numpy(xs[:, c].copy())
        rows.append(binary_auprc_packed(col, lab) if mode == 1 else binary_auroc_packed(col, lab))
    return torch.stack(rows)


def sort_u32(keys, vals=None):
    k = _np(keys).view(np.uint32)
    order = np.argsort(k, kind="stable")
    ok = torch.from_numpy(k[order].view(np.int32).copy())
    return ok, (None if vals is None else vals[torch.from_numpy(order)])


def dense_rank(x, with_values=False):
    vals, inv = np.unique(_np(x), return_inverse=True)
    rank = torch.from_numpy(inv.reshape(tuple(x.shape)).astype(np.int32))
    return rank, (torch.from_numpy(vals) if with_values else None)


def _digit_hist(k):
    """int32[1024]: the four 8-bit digit histograms of uint32 keys (pass-major)."""
    return np.concatenate([np.bincount((k >> np.uint32(8 * p)) & np.uint32(255), minlength=256) for p in range(4)]).astype(np.int32)


def rank_pack(score, target, splitters=(), want_hist=False):
    if isinstance(splitters, torch.Tensor):  # "device" splitters: int32 bit patterns
        splitters = [int(v) & 0xFFFFFFFF for v in splitters.tolist()]
    x = _np(score).astype(np.float32).copy()
    bits = x.view(np.uint32).copy()
    bits[bits == 0x80000000] = 0
    y = _np(target)
    lab = (y == 1).astype(np.uint32)
    keys = ((bits << np.uint32(1)) | lab).astype(np.uint32)
    info = torch.tensor([int(lab.sum()), _flags(score, target)], dtype=torch.int64)
    kt = torch.from_numpy(keys.view(np.int32).copy())
    part, counts = partition_u32(kt, splitters)
    if not want_hist:
        return kt, info, counts
    pk, off, rows = part.numpy().view(np.uint32), 0, []
    for c in counts.tolist():
        rows.append(_digit_hist(pk[off : off + c]))
        off += c
    return kt, info, counts, torch.from_numpy(np.stack(rows))


def partition_u32(keys, splitters, counts_in=None):
    k = _np(keys).view(np.uint32)
    sp = np.asarray([int(v) & 0xFFFFFFFF for v in splitters], dtype=np.uint64)
    bucket = np.searchsorted(sp, k.astype(np.uint64), side="right") if sp.size else np.zeros(k.size, dtype=np.int64)
    order = np.argsort(bucket, kind="stable")
    counts = np.bincount(bucket, minlength=len(splitters) + 1).astype(np.int64)
    return torch.from_numpy(k[order].view(np.int32).copy()), torch.from_numpy(counts)


def binary_auroc_from_keys(keys, hist=None, variant=0):
    k = _np(keys).view(np.uint32)
    if hist is not None:  # histograms that travelled with the keys must be THE histograms of the keys
        assert np.array_equal(_np(hist), _digit_hist(k)), "digit histograms do not match the received keys"
    x = (k >> np.uint32(1)).astype(np.uint32).view(np.float32)
    y = (k & np.uint32(1)).astype(np.int64)
    return binary_auroc_packed(torch.from_numpy(x.copy()), torch.from_numpy(y))


def binary_auprc_from_keys_begin(keys, hist=None):
    k = np.sort(_np(keys).view(np.uint32))
    return binary_auroc_from_keys(keys, hist), k


def binary_auprc_from_keys_finish(state, positives_above, samples_above=None):
    """sum over the local positives of TP/(TP+FP) at their threshold (ascending sorted keys)."""
    if isinstance(positives_above, torch.Tensor):
        positives_above, samples_above = (int(v) for v in positives_above.tolist())
    k = state
    out = torch.zeros(8, dtype=torch.float64)
    if k.size == 0:
        return out
    s, y = k >> np.uint32(1), (k & np.uint32(1)).astype(np.int64)
    n, P = k.size, int(y.sum())
    start = np.concatenate([[True], s[1:] != s[:-1]])
    gp = np.maximum.accumulate(np.where(start, np.arange(n), 0))  # index of my tie group's start
    pos_before = np.concatenate([[0], np.cumsum(y)[:-1]])[gp]      # positives strictly below my score
    tp = (P - pos_before + positives_above).astype(np.float64)
    den = (n - gp + samples_above).astype(np.float64)
    out[3] = float(np.sum((tp / den)[y == 1]))
    out[1], out[2] = P, n - P
    return out


def binary_pr_curve(score, target):
    """Ascending thresholds; one point per distinct score + the closing (1, 0)."""
    f = _flags(score, target)
    if f:
        return None, f
    x, y = _np(score).astype(np.float32), _np(target).astype(np.int64)
    x = np.where(x == 0, np.float32(0.0), x)  # -0.0 ties with +0.0
    order = np.argsort(x, kind="stable")
    xs, ys = x[order], y[order]
    n, P = xs.size, int(ys.sum())
    if n == 0:
        return (torch.ones(1), torch.zeros(1), torch.zeros(0)), 0
    start = np.concatenate([[True], xs[1:] != xs[:-1]])
    gp = np.nonzero(start)[0]
    pos_before = np.concatenate([[0], np.cumsum(ys)[:-1]])[gp]
    tp = (P - pos_before).astype(np.float64)
    prec = (tp / (n - gp)).astype(np.float32)
    rec = (tp / P).astype(np.float32) if P else np.ones(gp.size, dtype=np.float32)
    prec = np.concatenate([prec, np.ones(1, dtype=np.float32)])
    rec = np.concatenate([rec, np.zeros(1, dtype=np.float32)])
    return (torch.from_numpy(prec), torch.from_numpy(rec), torch.from_numpy(xs[gp].copy())), 0



# ---- host restatements of the two one-CTA helpers of the sample sort (te_select_splitters,
# ---- te_combine_auroc_stats): what the gloo tests run in their place, and what the GPU tests
# ---- check the kernels against
_NO_SAMPLE = 1 << 33  # above every 32-bit key: what a rank without data contributes to the pool


def choose_splitters(samples_sorted, world: int):
    """W-1 ascending splitters from the sorted pool of sampled keys; label bit cleared so that
    both labels of one score compare alike.  (Host restatement of what te_select_splitters does
    on the device; the tests check the kernel against it.)"""
    m = len(samples_sorted)
    out = []
    for i in range(1, world):
        v = int(samples_sorted[min(m - 1, (i * m) // world)]) if m else 0
        out.append(v & 0xFFFFFFFE)
    return out


def combine_rank_stats(stats):
    """stats: list over ranks (ascending score ranges) of (P_r, N_r, 2U_r) Python ints.
    Returns (auroc, P, N, 2U) with exact integer arithmetic.  (Host restatement of
    te_combine_auroc_stats / combine_rank_stats_tensor; the tests check both against it.)"""
    u2 = 0
    neg_below = 0
    P = N = 0
    for p_r, n_r, u2_r in stats:
        u2 += u2_r + 2 * p_r * neg_below
        neg_below += n_r
        P += p_r
        N += n_r
    auroc = 0.5 if P == 0 or N == 0 else u2 / (2 * P * N)
    return auroc, P, N, u2


def combine_rank_stats_tensor(st: torch.Tensor):
    """st: int64 (W, 4) rows {P_r, N_r, 2U_r >> 32, 2U_r & 0xffffffff}, ranks in ascending score
    order.  Returns (auroc float64, P, N, 2U int64) as 0-d tensors on st.device, no host read:
        2U = sum_r ( 2U_r + 2 * P_r * sum_{q<r} N_q )        (fits int64 for N_total < 2^31)."""
    p, nn = st[:, 0], st[:, 1]
    u2r = (st[:, 2] << 32) | st[:, 3]
    neg_below = torch.cumsum(nn, 0) - nn
    u2 = (u2r + 2 * p * neg_below).sum()
    P, Nn = p.sum(), nn.sum()
    den = 2.0 * P.to(torch.float64) * Nn.to(torch.float64)
    auroc = torch.where(den == 0, torch.full_like(den, 0.5), u2.to(torch.float64) / den.clamp(min=1.0))
    return auroc, P, Nn, u2



def select_splitters(samples, world):
    """(int32[W-1], int64[W-1]) quantile splitters in key space, like ops.select_splitters."""
    pool = samples.reshape(-1)
    pool, _ = torch.sort(torch.where(pool >= 0, pool.to(torch.int64), torch.full_like(pool, _NO_SAMPLE, dtype=torch.int64)))
    m = (pool < _NO_SAMPLE).sum()
    i = torch.arange(1, world, dtype=torch.int64)
    pick = torch.minimum(torch.div(i * m, world, rounding_mode="floor"), (m - 1).clamp(min=0))
    sp = torch.where(m > 0, (pool[pick] << 1) & 0xFFFFFFFE, torch.zeros_like(pick))
    return sp.to(torch.int32), sp


def combine_auroc_stats(rows, n_received):
    st = torch.stack([rows[:, 1], rows[:, 2], rows[:, 5], rows[:, 6]], dim=1).to(torch.int64)
    auroc, P, Nn, u2 = combine_rank_stats_tensor(st)
    out = torch.zeros(8, dtype=torch.float64)
    out[:4] = torch.stack([auroc, P.to(torch.float64), Nn.to(torch.float64), u2.to(torch.float64)])
    out[5] = float(n_received)
    out[6], out[7] = (u2 >> 32).to(torch.float64), (u2 & 0xFFFFFFFF).to(torch.float64)
    return out


_NAMES = [
    "cm_update_", "cm_update_logits_", "class_counts_", "match_count", "binned_hist_update_",
    "binned_hist2d_update_", "binned_counts", "binned_integrate", "binary_auroc_packed",
    "binary_auprc_packed", "binary_auroc_general", "binary_auprc_general", "sort_u32", "rank_pack", "ovr_rank",
    "partition_u32", "binary_auroc_from_keys", "binary_auprc_from_keys_begin", "binary_auprc_from_keys_finish", "binary_pr_curve",
    "select_splitters", "combine_auroc_stats", "dense_rank", "binary_auroc_general_batched",
]


def install(monkeypatch=None):
    """Route torcheval_b200.ops to the CPU stand-ins and make the CPU the 'compute device'."""
    import torcheval_b200._binned as _binned
    import torcheval_b200._counting as _counting
    import torcheval_b200._ranking as _ranking
    import torcheval_b200.ops as ops

    def setter(obj, name, val):
        if monkeypatch is not None:
            monkeypatch.setattr(obj, name, val)
        else:
            setattr(obj, name, val)

    g = globals()
    for name in _NAMES:
        setter(ops, name, g[name])
    cpu = lambda t: torch.device("cpu")  # noqa: E731
    for mod in (_ranking, _binned, _counting):
        setter(mod, "compute_device", cpu)
    import torcheval_b200.metrics.functional.classification.accuracy as acc
    import torcheval_b200.metrics.functional.classification.binned_auroc as bauroc
    import torcheval_b200.metrics.functional.classification.f1_score as f1

    import torcheval_b200.metrics.functional.classification.precision as prec
    import torcheval_b200.metrics.functional.classification.recall as rec

    for mod in (acc, bauroc, f1, prec, rec):  # (modules that bound the name at import time)
        setter(mod, "compute_device", cpu)
    import torcheval_b200.distributed as tdist

    setter(tdist, "compute_device", cpu) if hasattr(tdist, "compute_device") else None
