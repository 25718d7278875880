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
