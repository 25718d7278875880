"""binary / multiclass / multilabel_precision_recall_curve with the reference's signatures
(torcheval/metrics/functional/classification/precision_recall_curve.py:19-285).  The sort, the
tie-boundary compaction and the cumulative counts run in the ranking kernels (csrc/rank.cu:
radix passes, count sweep, `rank_pr_curve_kernel`); SURVEY.md section 8f #1."""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

from .... import _ranking


@torch.inference_mode()
def binary_precision_recall_curve(
    input: torch.Tensor, target: torch.Tensor
) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(precision, recall, thresholds): thresholds ascending, one per distinct score; precision and
    recall have one more element, the closing point (1, 0)."""
    _binary_precision_recall_curve_update(input, target)
    return _binary_precision_recall_curve_compute(input, target)


def _binary_precision_recall_curve_update(input: torch.Tensor, target: torch.Tensor) -> None:
    _binary_precision_recall_curve_update_input_check(input, target)


def _binary_precision_recall_curve_compute(
    input: torch.Tensor, target: torch.Tensor
) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    return _compute_for_each_class(input, target, 1)


def _compute_for_each_class(
    input: torch.Tensor, target: torch.Tensor, pos_label: int
) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    return _ranking.binary_precision_recall_curve(input, target, pos_label)


def _binary_precision_recall_curve_update_input_check(input: torch.Tensor, target: torch.Tensor) -> None:
    if input.ndim != 1:
        raise ValueError("input should be a one-dimensional tensor, " f"got shape {input.shape}.")
    if target.ndim != 1:
        raise ValueError("target should be a one-dimensional tensor, " f"got shape {target.shape}.")
    if input.shape != target.shape:
        raise ValueError(
            "The `input` and `target` should have the same shape, "
            f"got shapes {input.shape} and {target.shape}."
        )


@torch.inference_mode()
def multiclass_precision_recall_curve(
    input: torch.Tensor, target: torch.Tensor, *, num_classes: Optional[int] = None
) -> Tuple[List[torch.Tensor], List[torch.Tensor], List[torch.Tensor]]:
    """One-vs-rest curves: three lists of num_classes tensors."""
    if num_classes is None and input.ndim == 2:
        num_classes = input.shape[1]
    _multiclass_precision_recall_curve_update(input, target, num_classes)
    return _multiclass_precision_recall_curve_compute(input, target, num_classes)


def _multiclass_precision_recall_curve_update(
    input: torch.Tensor, target: torch.Tensor, num_classes: Optional[int]
) -> None:
    _multiclass_precision_recall_curve_update_input_check(input, target, num_classes)


def _multiclass_precision_recall_curve_compute(
    input: torch.Tensor, target: torch.Tensor, num_classes: Optional[int]
) -> Tuple[List[torch.Tensor], List[torch.Tensor], List[torch.Tensor]]:
    if num_classes is None:
        num_classes = input.shape[1]
    precisions, recalls, thresholds = [], [], []
    for c in range(num_classes):
        p, r, t = _compute_for_each_class(input[:, c], target, c)
        precisions.append(p)
        recalls.append(r)
        thresholds.append(t)
    return precisions, recalls, thresholds


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


@torch.inference_mode()
def multilabel_precision_recall_curve(
    input: torch.Tensor, target: torch.Tensor, *, num_labels: Optional[int] = None
) -> Tuple[List[torch.Tensor], List[torch.Tensor], List[torch.Tensor]]:
    if input.ndim != 2:
        raise ValueError(f"input should be a two-dimensional tensor, got shape {input.shape}.")
    if num_labels is None:
        num_labels = input.shape[1]
    _multilabel_precision_recall_curve_update(input, target, num_labels)
    return _multilabel_precision_recall_curve_compute(input, target, num_labels)


def _multilabel_precision_recall_curve_update(input: torch.Tensor, target: torch.Tensor, num_labels: int) -> None:
    _multilabel_precision_recall_curve_update_input_check(input, target, num_labels)


def _multilabel_precision_recall_curve_compute(
    input: torch.Tensor, target: torch.Tensor, num_labels: int
) -> Tuple[List[torch.Tensor], List[torch.Tensor], List[torch.Tensor]]:
    precisions, recalls, thresholds = [], [], []
    for i in range(num_labels):
        p, r, t = _compute_for_each_class(input[:, i], target[:, i], 1)
        precisions.append(p)
        recalls.append(r)
        thresholds.append(t)
    return precisions, recalls, thresholds


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
