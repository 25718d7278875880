"""Host-side dispatch for the exact ranking metrics: staging, dtype handling, choice between the
packed fast path and the general path, and the one device->host read of the status word."""
from __future__ import annotations

from typing import Optional

import torch

from . import _native as N
from . import ops


def compute_device(t: torch.Tensor) -> torch.device:
    """Where the kernels run for this input: its own device if CUDA, else the current CUDA device."""
    if t.is_cuda:
        return t.device
    if not torch.cuda.is_available():
        raise RuntimeError(
            "torcheval_b200 needs a CUDA device: its metrics are computed by sm_100a kernels and "
            "there is no CPU fallback."
        )
    return torch.device("cuda", torch.cuda.current_device())


def stage(t: torch.Tensor, dev: torch.device, dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    if t.device != dev:
        t = t.to(dev, non_blocking=True)
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    return t


def _as_f32_scores(x: torch.Tensor) -> torch.Tensor:
    """`x` as float32 if that is exact, else NotImplementedError (callers decide what to do:
    `_rank_scores` for ranking, `_floor_f32_scores` for threshold counting, the generic gather
    path for rank-sharded scores, whose ranks would have to agree across ranks)."""
    if x.dtype == torch.float32:
        return x
    if x.dtype in (torch.float16, torch.bfloat16):
        return x.float()  # exact
    y = x.float()
    if not bool((y.to(x.dtype) == x).all()):
        raise NotImplementedError(
            f"scores of dtype {x.dtype} that are not exactly representable in float32"
        )
    return y


def _floor_f32_scores(x: torch.Tensor) -> torch.Tensor:
    """Largest float32 <= x.  For float32 thresholds t: t <= x  <=>  t <= floor32(x), so counting
    scores against thresholds is unchanged by this rounding (unlike round-to-nearest)."""
    try:
        return _as_f32_scores(x)
    except NotImplementedError:
        y = x.float()
        above = y.to(x.dtype) > x
        return torch.where(above, torch.nextafter(y, torch.full_like(y, float("-inf"))), y)


def _dense_rank_scores(x: torch.Tensor, with_values: bool = False):
    """(dense ranks of `x` as float32 BIT PATTERNS, distinct values ascending or None): the native
    rank transform (te_dense_rank: two pair sorts on the 64-bit image of the scores + a scan).
    Ranks order and tie exactly like the scores, non-negative floats order like their bit patterns,
    and the kernels only ever compare and sort the bits."""
    if x.numel() > 0x7F800000:  # a rank would read as inf / NaN
        raise NotImplementedError(f"{x.numel()} scores of dtype {x.dtype}")
    if x.is_floating_point():
        wide = x if x.dtype == torch.float64 else x.double()  # exact
    else:
        wide = x if x.dtype == torch.int64 else x.to(torch.int64)
    rank, values = ops.dense_rank(wide, with_values)
    if values is not None and values.dtype != x.dtype:
        values = values.to(x.dtype)
    return rank.view(torch.float32), values


def _rank_scores(x: torch.Tensor) -> torch.Tensor:
    """A float32 tensor with the same order and the same ties as `x` — all the ranking kernels
    look at.  Scores that float32 holds exactly are passed through; anything else (float64
    probabilities, large integers) is replaced by its dense rank (the reference sorts any dtype in
    place, auroc.py:121; here the 64-bit sort is its own kernel train and the 32-bit kernels stay
    as they are)."""
    try:
        return _as_f32_scores(x)
    except NotImplementedError:
        return _dense_rank_scores(x)[0]


def binary_auroc(input: torch.Tensor, target: torch.Tensor, weight: Optional[torch.Tensor] = None) -> torch.Tensor:
    out_dev = input.device
    dev = compute_device(input)
    x = _rank_scores(stage(input, dev))
    t = stage(target, dev)
    w = None if weight is None else stage(weight, dev, torch.float64)
    if x.dim() == 1:
        return _binary_auroc_row(x, t, w).to(out_dev)
    if x.shape[0] == 1:
        return _binary_auroc_row(x[0], t[0], None if w is None else w[0]).reshape(1).to(out_dev)
    if w is None and not t.is_floating_point() and bool(((t == 0) | (t == 1)).all()):
        # (tasks, n) without weights: the tasks are independent rankings of equal length, i.e. the
        # multilabel one-vs-rest problem on the transposed arrays — eight tasks per launch
        out = ops.ovr_rank(x.t().contiguous(), t.t().contiguous().to(torch.int64), True, 0)
        vals = out[:, 0].clone()
        bad = torch.nonzero(out[:, 4] != 0).flatten()  # rows with negative scores
        if bad.numel():
            vals[bad] = ops.binary_auroc_general_batched(x[bad], t[bad], None)[:, 0]
        return vals.to(out_dev)
    # weights, soft labels: the general path, all tasks in one launch train (grid.y = task)
    return ops.binary_auroc_general_batched(x, t, w)[:, 0].clone().to(out_dev)


def _binary_auroc_row(x: torch.Tensor, t: torch.Tensor, w: Optional[torch.Tensor]) -> torch.Tensor:
    if w is None and t.dtype in ops._DTYPE_CODE:
        out = ops.binary_auroc_packed(x, t)
        flags = int(out[4].item())  # the one host read of this path
        if flags == N.TE_RANK_FLAG_WIDE_SCORE:  # scores >= 2.0: the counting stage does not apply, the full sort does
            out = ops.binary_auroc_packed(x, t, N.TE_RANK_SORT)
            flags = int(out[4].item())
        if flags == 0:
            return out[0].clone()
    return ops.binary_auroc_general(x, t, w)[0].clone()


def _binary_auprc_row(x: torch.Tensor, t: torch.Tensor) -> torch.Tensor:
    if t.dtype in ops._DTYPE_CODE:
        out = ops.binary_auprc_packed(x, t)
        flags = int(out[4].item())
        if flags == N.TE_RANK_FLAG_WIDE_SCORE:  # scores >= 2.0: the full sort handles all 32 key bits
            out = ops.binary_auprc_packed(x, t, N.TE_RANK_SORT)
            flags = int(out[4].item())
        if flags == 0:
            return out[0].clone()
    return ops.binary_auprc_general(x, t)[0].clone()


def binary_auprc(input: torch.Tensor, target: torch.Tensor, num_tasks: int = 1) -> torch.Tensor:
    """float32 like the reference (auprc.py:239-251): () for 1-D input with num_tasks == 1,
    otherwise (num_tasks,)."""
    out_dev = input.device
    dev = compute_device(input)
    x = _rank_scores(stage(input, dev))
    t = stage(target, dev)
    if num_tasks == 1 and x.dim() == 1:
        return _binary_auprc_row(x, t).to(torch.float32).to(out_dev)
    vals = [_binary_auprc_row(x[k], t[k]) for k in range(num_tasks)]
    return torch.stack(vals).to(torch.float32).to(out_dev)


def _one_vs_rest(input: torch.Tensor, target: torch.Tensor, auprc: bool, multilabel: bool) -> torch.Tensor:
    """Per class / per label values (float64, on the compute device): the batched one-vs-rest
    kernels do all columns; a column the packed representation cannot take (negative scores) is
    redone on its own through the general path."""
    dev = compute_device(input)
    x = _rank_scores(stage(input, dev))
    t = stage(target, dev)
    if t.dtype != torch.int64:
        t = t.to(torch.int64)
    out = ops.ovr_rank(x, t, multilabel, 1 if auprc else 0)
    vals = out[:, 0].clone()
    flagged = torch.nonzero(out[:, 4] != 0).flatten().tolist()  # the one host read of this path
    for c in flagged:
        col = x[:, c].contiguous()
        lab = (t[:, c] == 1) if multilabel else (t == c)
        vals[c] = ops.binary_auprc_general(col, lab)[0] if auprc else ops.binary_auroc_general(col, lab, None)[0]
    return vals


def multiclass_auroc(input: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    return _one_vs_rest(input, target, auprc=False, multilabel=False)


def multiclass_auprc(input: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    return _one_vs_rest(input, target, auprc=True, multilabel=False)


def multilabel_auprc(input: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    return _one_vs_rest(input, target, auprc=True, multilabel=True)


def binary_precision_recall_curve(input: torch.Tensor, target: torch.Tensor, pos_label: int = 1):
    """(precision, recall, thresholds) of the reference's ``_compute_for_each_class``
    (precision_recall_curve.py:208-231): thresholds = distinct scores, ascending, in the input's
    dtype; precision / recall float32 with the closing (1, 0) appended; on input.device.
    Positives are ``target == pos_label``; anything else counts as a negative."""
    out_dev = input.device
    dev = compute_device(input)
    x = stage(input, dev)
    t = stage(target, dev)
    lab = t if (pos_label == 1 and t.dtype in ops._DTYPE_CODE) else (t == pos_label).to(torch.uint8)
    curve, flags = None, N.TE_RANK_FLAG_NEGATIVE_SCORE
    values = None
    if x.dtype in (torch.float32, torch.float16, torch.bfloat16):
        curve, flags = ops.binary_pr_curve(x.float(), lab)
        if flags & N.TE_RANK_FLAG_BAD_LABEL:  # labels other than 0 / 1 are negatives here, not an error
            lab = (t == pos_label).to(torch.uint8)
            curve, flags = ops.binary_pr_curve(x.float(), lab)
    if flags:
        # negative scores, or a dtype float32 cannot hold: dense ranks go through the kernel (they
        # order and tie like the scores), the thresholds are the distinct score values themselves
        ranks, values = _dense_rank_scores(x, with_values=True)
        if lab.dtype != torch.uint8:
            lab = (t == pos_label).to(torch.uint8)
        curve, flags = ops.binary_pr_curve(ranks, lab)
        assert flags == 0
    precision, recall, thresholds = curve
    thresholds = values if values is not None else thresholds.to(input.dtype)
    return precision.to(out_dev), recall.to(out_dev), thresholds.to(out_dev)
