"""CPU port of the reference hot path in torch CPU ops (TEST / BASELINE INFRASTRUCTURE ONLY).

The reference *is* a sequence of ATen ops, so its CPU performance is the performance of those ops
on the host cores.  This module restates the op sequences (sort -> gather -> cumsum -> tie mask ->
trapezoid; sparse-COO coalesce for the confusion matrix; searchsorted + histc for the binned
counts) so that bench.py can time "the reference's CPU path" on the GPU box, where
/root/reference does not exist.  Every function is checked against the C oracle and the reference
fixtures in tests/test_oracle_golden.py.  Only bench.py (cpu_baseline / --impl reference) and
tests/ import it.
"""
from __future__ import annotations

import torch


@torch.inference_mode()
def binary_auroc(score: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """The op sequence of the reference's class path, op for op: BinaryAUROC.update stores a float64
    weight of ones when none is given (classification/auroc.py:112-113), and compute() runs
    _binary_auroc_compute_jit (functional/classification/auroc.py:115-152) on it — descending sort,
    `diff != 0` padded with a trailing 1, two gathers, two cumsums, the right-aligned
    masked_scatter_ of the points at tie boundaries into zero-filled arrays, trapz, where.
    This is what bench.py --impl reference times: nothing is skipped or shortened."""
    n = score.numel()
    weight = torch.ones_like(score, dtype=torch.float64)
    thr, order = score.sort(descending=True)
    boundary = torch.nn.functional.pad(thr.diff(dim=-1) != 0, [0, 1], value=1.0)
    st = torch.gather(target, -1, order)
    sw = torch.gather(weight, -1, order)
    ctp = (sw * st).cumsum(-1)
    cfp = (sw * (1 - st)).cumsum(-1)
    right = boundary.sum(-1, keepdim=True) >= torch.arange(n, 0, -1)
    tp = torch.zeros_like(ctp)
    fp = torch.zeros_like(cfp)
    tp.masked_scatter_(right, ctp[boundary])
    fp.masked_scatter_(right, cfp[boundary])
    factor = tp[-1] * fp[-1]
    return torch.where(factor == 0, 0.5, torch.trapz(tp, fp).double() / factor)


@torch.inference_mode()
def binary_auprc(score: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """precision_recall_curve.py:208-231 + tensor_utils.py:12-16."""
    s, idx = torch.sort(score, descending=True)
    pos = target[idx] == 1
    last = torch.ones_like(s, dtype=torch.bool)
    last[:-1] = s[1:] != s[:-1]
    tp = torch.cumsum(pos, 0)[last].to(torch.float64)
    fp = torch.cumsum(~pos, 0)[last].to(torch.float64)
    if tp[-1] == 0:
        return torch.tensor(-0.0, dtype=torch.float64)
    prec = tp / (tp + fp)
    rec = tp / tp[-1]
    prev = torch.cat([rec.new_zeros(1), rec[:-1]])
    return torch.sum((rec - prev) * prec)


@torch.inference_mode()
def confusion_matrix(pred: torch.Tensor, target: torch.Tensor, num_classes: int) -> torch.Tensor:
    """confusion_matrix.py:219-234: COO entries at (target, pred), coalesced by to_dense()."""
    coords = torch.vstack((target, pred))
    return torch.sparse_coo_tensor(coords, torch.ones_like(target), (num_classes, num_classes)).to_dense()


@torch.inference_mode()
def binned_counts(score: torch.Tensor, target: torch.Tensor, threshold: torch.Tensor):
    """binned_precision_recall_curve.py:84-110."""
    T = threshold.numel()
    key = (2 * (torch.searchsorted(threshold, score, right=True) - 1) + target).to(torch.float64)
    hist = torch.histc(key, bins=2 * T, min=0, max=2 * T).reshape(T, 2).T.to(torch.int64)
    suffix = hist.flip(1).cumsum(1).flip(1)
    num_fp, num_tp = suffix[0], suffix[1]
    return num_tp, num_fp, target.sum() - num_tp
