"""Host-side dispatch for the counting metrics (confusion matrix and what derives from it)."""
from __future__ import annotations

import torch

from . import ops
from ._ranking import compute_device, stage

_I64_MAX = torch.iinfo(torch.int64).max
_I64_MIN = torch.iinfo(torch.int64).min


def confusion_matrix_counts(input: torch.Tensor, target: torch.Tensor, num_classes: int) -> torch.Tensor:
    """int64 (C, C) counts on the compute device; raises the reference's ValueErrors for labels
    outside [0, C) (functional/classification/confusion_matrix.py:271-282), from the fused range
    report of the kernel — one small device->host read instead of two torch.max syncs."""
    dev = compute_device(input)
    t = stage(target, dev)
    if t.dtype != torch.int64:
        # sparse_coo_tensor in the reference rejects anything else (confusion_matrix.py:228-232)
        raise RuntimeError("indices must be an int64 tensor")
    cm = torch.zeros((num_classes, num_classes), dtype=torch.int64, device=dev)
    if input.ndim == 2:
        x = stage(input, dev)
        if x.dtype != torch.float32:
            x = x.float()
        rng = ops.cm_update_logits_(cm, x, t, num_classes)
    else:
        p = stage(input, dev)
        if p.dtype != torch.int64:
            raise RuntimeError("indices must be an int64 tensor")
        rng = ops.cm_update_(cm, p, t, num_classes)
    min_p, max_p, min_t, max_t = rng.tolist()
    if max_p != _I64_MIN and max_p >= num_classes:
        raise ValueError(
            "Got `input` prediction class which is too large for the number of classes, "
            f"num_classes: {num_classes} must be strictly greater than max class predicted: {max_p}."
        )
    if max_t != _I64_MIN and max_t >= num_classes:
        raise ValueError(
            "Got `target` class which is larger than the number of classes, "
            f"num_classes: {num_classes} must be strictly greater than max target: {max_t}."
        )
    if min_p != _I64_MAX and min_p < 0:
        raise ValueError(f"Got negative `input` prediction class {min_p}; classes must be in [0, {num_classes}).")
    if min_t != _I64_MAX and min_t < 0:
        raise ValueError(f"Got negative `target` class {min_t}; classes must be in [0, {num_classes}).")
    return cm


def _raise_for_range(rng: torch.Tensor, num_classes: int) -> None:
    min_p, max_p, min_t, max_t = rng.tolist()
    if max_p != _I64_MIN and max_p >= num_classes:
        raise ValueError(
            f"Got `input` class {max_p} which is too large for num_classes={num_classes}."
        )
    if max_t != _I64_MIN and max_t >= num_classes:
        raise ValueError(
            f"Got `target` class {max_t} which is too large for num_classes={num_classes}."
        )
    if (min_p != _I64_MAX and min_p < 0) or (min_t != _I64_MAX and min_t < 0):
        raise ValueError(f"Got a negative class label; classes must be in [0, {num_classes}).")


def class_counts(input: torch.Tensor, target: torch.Tensor, num_classes: int):
    """(num_tp, num_label, num_pred): int64 (C,) each, on the compute device — the diagonal, row
    sums and column sums of the confusion matrix.  input: (n,) classes or (n, C) scores."""
    dev = compute_device(input)
    t = stage(target, dev, torch.int64)
    x = stage(input, dev)
    x = x.float() if x.dim() == 2 and x.dtype != torch.float32 else x
    if x.dim() == 1 and x.dtype != torch.int64:
        x = x.to(torch.int64)
    c3 = torch.zeros((3, num_classes), dtype=torch.int64, device=dev)
    rng = ops.class_counts_(c3, x, t, num_classes)
    _raise_for_range(rng, num_classes)
    return c3[0], c3[1], c3[2]


def match_count(input: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """int64 scalar on the compute device: #{pred == target}; (n, C) scores are argmax'ed."""
    dev = compute_device(input)
    t = stage(target, dev, torch.int64)
    x = stage(input, dev)
    if x.dim() == 2:
        C = x.shape[1]
        x = x.float() if x.dtype != torch.float32 else x
        c3 = torch.zeros((3, C), dtype=torch.int64, device=dev)
        ops.class_counts_(c3, x, t, C)
        # (a target outside [0, C) is skipped by the kernel = counted as a miss, which is what the
        # reference's `(argmax == target).sum()` does with it, accuracy.py:269-272: no error here)
        return c3[0].sum()
    return ops.match_count(x.to(torch.int64), t)
