"""Binned precision-recall curves with the reference's signatures
(torcheval/metrics/functional/classification/binned_precision_recall_curve.py:20-548).  The
per-threshold TP/FP/FN counts come from the shared-memory histogram kernels (csrc/binned.cu);
the `optimization` argument is validated and otherwise ignored — there is one kernel."""
from __future__ import annotations

from typing import List, Optional, Tuple, Union

import torch

from .... import _binned
from ..tensor_utils import THR_RANGE, THR_UNSORTED, _create_threshold_tensor, _threshold_faults

DEFAULT_NUM_THRESHOLD = 100


@torch.inference_mode()
def binary_binned_precision_recall_curve(
    input: torch.Tensor,
    target: torch.Tensor,
    *,
    threshold: Union[int, List[float], torch.Tensor] = DEFAULT_NUM_THRESHOLD,
) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    threshold = _create_threshold_tensor(threshold, target.device)
    _binned_precision_recall_curve_param_check(threshold)
    num_tp, num_fp, num_fn = _binary_binned_precision_recall_curve_update(input, target, threshold)
    return _binary_binned_precision_recall_curve_compute(num_tp, num_fp, num_fn, threshold)


@torch.inference_mode()
def multiclass_binned_precision_recall_curve(
    input: torch.Tensor,
    target: torch.Tensor,
    num_classes: Optional[int] = None,
    threshold: Union[int, List[float], torch.Tensor] = DEFAULT_NUM_THRESHOLD,
    optimization: str = "vectorized",
) -> Tuple[List[torch.Tensor], List[torch.Tensor], torch.Tensor]:
    _optimization_param_check(optimization)
    threshold = _create_threshold_tensor(threshold, target.device)
    _binned_precision_recall_curve_param_check(threshold)
    if num_classes is None and input.ndim == 2:
        num_classes = input.shape[1]
    num_tp, num_fp, num_fn = _multiclass_binned_precision_recall_curve_update(
        input, target, num_classes, threshold, optimization
    )
    return _multiclass_binned_precision_recall_curve_compute(num_tp, num_fp, num_fn, num_classes, threshold)


@torch.inference_mode()
def multilabel_binned_precision_recall_curve(
    input: torch.Tensor,
    target: torch.Tensor,
    num_labels: Optional[int] = None,
    threshold: Union[int, List[float], torch.Tensor] = DEFAULT_NUM_THRESHOLD,
    optimization: str = "vectorized",
) -> Tuple[List[torch.Tensor], List[torch.Tensor], torch.Tensor]:
    _optimization_param_check(optimization)
    threshold = _create_threshold_tensor(threshold, target.device)
    _binned_precision_recall_curve_param_check(threshold)
    if num_labels is None:
        num_labels = input.shape[1]
    num_tp, num_fp, num_fn = _multilabel_binned_precision_recall_curve_update(
        input, target, num_labels, threshold, optimization
    )
    return _multilabel_binned_precision_recall_curve_compute(num_tp, num_fp, num_fn, num_labels, threshold)


# ------------------------------------------------------------------------------ updates
def _binary_binned_precision_recall_curve_update(
    input: torch.Tensor, target: torch.Tensor, threshold: torch.Tensor
) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(num_tp, num_fp, num_fn), each (T,) int64 on input's device (reference :75-110)."""
    _binary_precision_recall_curve_update_input_check(input, target)
    tp, fp, fn = _binned.binary_counts(input, target, threshold)
    d = input.device
    return tp[0].to(d), fp[0].to(d), fn[0].to(d)


def _multiclass_binned_precision_recall_curve_update(
    input: torch.Tensor,
    target: torch.Tensor,
    num_classes: Optional[int],
    threshold: torch.Tensor,
    optimization: str = "vectorized",
) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(num_tp, num_fp, num_fn), each (T, C) int64 (reference :214-309)."""
    _optimization_param_check(optimization)
    _multiclass_precision_recall_curve_update_input_check(input, target, num_classes)
    tp, fp, fn = _binned.counts_2d(input, target, threshold, multilabel=False)
    d = input.device
    return tp.T.contiguous().to(d), fp.T.contiguous().to(d), fn.T.contiguous().to(d)


def _multilabel_binned_precision_recall_curve_update(
    input: torch.Tensor,
    target: torch.Tensor,
    num_labels: int,
    threshold: torch.Tensor,
    optimization: str = "vectorized",
) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(num_tp, num_fp, num_fn), each (T, L) int64 (reference :406-504)."""
    _optimization_param_check(optimization)
    _multilabel_precision_recall_curve_update_input_check(input, target, num_labels)
    tp, fp, fn = _binned.counts_2d(input, target, threshold, multilabel=True)
    d = input.device
    return tp.T.contiguous().to(d), fp.T.contiguous().to(d), fn.T.contiguous().to(d)


# ------------------------------------------------------------------------------ computes
def _binary_binned_precision_recall_curve_compute(
    num_tp: torch.Tensor, num_fp: torch.Tensor, num_fn: torch.Tensor, threshold: torch.Tensor
) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    precision, recall = _binned.pr_curve_from_counts(num_tp, num_fp, num_fn, dim=0)
    return precision, recall, threshold


def _multiclass_binned_precision_recall_curve_compute(
    num_tp: torch.Tensor,
    num_fp: torch.Tensor,
    num_fn: torch.Tensor,
    num_classes: Optional[int],
    threshold: torch.Tensor,
) -> Tuple[List[torch.Tensor], List[torch.Tensor], torch.Tensor]:
    precision, recall = _binned.pr_curve_from_counts(num_tp, num_fp, num_fn, dim=0)
    return list(precision.T), list(recall.T), threshold


def _multilabel_binned_precision_recall_curve_compute(
    num_tp: torch.Tensor,
    num_fp: torch.Tensor,
    num_fn: torch.Tensor,
    num_labels: int,
    threshold: torch.Tensor,
) -> Tuple[List[torch.Tensor], List[torch.Tensor], torch.Tensor]:
    precision, recall = _binned.pr_curve_from_counts(num_tp, num_fp, num_fn, dim=0)
    return list(precision.T), list(recall.T), threshold


# ------------------------------------------------------------------------------ checks
def _binned_precision_recall_curve_param_check(threshold: torch.Tensor) -> None:
    faults = _threshold_faults(threshold)
    if faults is None:
        faults = (THR_UNSORTED if (torch.diff(threshold) < 0.0).any() else 0) | (
            THR_RANGE if (threshold < 0.0).any() or (threshold > 1.0).any() else 0
        )
    if faults & THR_UNSORTED:
        raise ValueError("The `threshold` should be a sorted tensor.")
    if faults & THR_RANGE:
        raise ValueError("The values in `threshold` should be in the range of [0, 1].")


def _optimization_param_check(optimization: str) -> None:
    if optimization not in ("vectorized", "memory"):
        raise ValueError(
            f"Unknown memory approach: expected 'vectorized' or 'memory', but got {optimization}."
        )


def _binary_precision_recall_curve_update_input_check(input: torch.Tensor, target: torch.Tensor) -> None:
    if input.ndim != 1:
        raise ValueError(f"input should be a one-dimensional tensor, got shape {input.shape}.")
    if target.ndim != 1:
        raise ValueError(f"target should be a one-dimensional tensor, got shape {target.shape}.")
    if input.shape != target.shape:
        raise ValueError(
            "The `input` and `target` should have the same shape, "
            f"got shapes {input.shape} and {target.shape}."
        )


def _multiclass_precision_recall_curve_update_input_check(
    input: torch.Tensor, target: torch.Tensor, num_classes: Optional[int]
) -> None:
    if input.size(0) != target.size(0):
        raise ValueError(
            "The `input` and `target` should have the same first dimension, "
            f"got shapes {input.shape} and {target.shape}."
        )
    if target.ndim != 1:
        raise ValueError("target should be a one-dimensional tensor, " f"got shape {target.shape}.")
    if not (input.ndim == 2 and (num_classes is None or input.shape[1] == num_classes)):
        raise ValueError(
            f"input should have shape of (num_sample, num_classes), "
            f"got {input.shape} and num_classes={num_classes}."
        )


def _multilabel_precision_recall_curve_update_input_check(
    input: torch.Tensor, target: torch.Tensor, num_labels: int
) -> None:
    if input.shape != target.shape:
        raise ValueError(
            "Expected both input.shape and target.shape to have the same shape"
            f" but got {input.shape} and {target.shape}."
        )
    if input.ndim != 2:
        raise ValueError(f"input should be a two-dimensional tensor, got shape {input.shape}.")
    if input.shape[1] != num_labels:
        raise ValueError(
            f"input should have shape of (num_sample, num_labels), "
            f"got {input.shape} and num_labels={num_labels}."
        )
