"""binary_precision / multiclass_precision with the reference's signatures
(torcheval/metrics/functional/classification/precision.py:15-250).  The counts are the diagonal
and the column sums of the confusion matrix and come from the class-count kernel (csrc/cm.cu:
3*C shared-memory bins, one streaming pass) instead of three scatter_ calls (:128-134)."""
from __future__ import annotations

import logging
from typing import Optional, Tuple

import torch

from .... import _counting
from ...._ranking import compute_device, stage


@torch.inference_mode()
def binary_precision(input: torch.Tensor, target: torch.Tensor, *, threshold: float = 0.5) -> torch.Tensor:
    """TP / (TP + FP) after ``torch.where(input < threshold, 0, 1)``."""
    num_tp, num_fp, num_label = _binary_precision_update(input, target, threshold)
    return _precision_compute(num_tp, num_fp, num_label, "micro")


@torch.inference_mode()
def multiclass_precision(
    input: torch.Tensor,
    target: torch.Tensor,
    *,
    num_classes: Optional[int] = None,
    average: Optional[str] = "micro",
) -> torch.Tensor:
    _precision_param_check(num_classes, average)
    num_tp, num_fp, num_label = _precision_update(input, target, num_classes, average)
    return _precision_compute(num_tp, num_fp, num_label, average)


def _precision_update(
    input: torch.Tensor, target: torch.Tensor, num_classes: Optional[int], average: Optional[str]
) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(num_tp, num_fp, num_label): scalars for micro (num_label unused, 0.0), else (C,) int64."""
    _precision_update_input_check(input, target, num_classes)
    out_dev = input.device
    if isinstance(average, str) and average == "micro":
        num_tp = _counting.match_count(input, target).to(out_dev)
        num_fp = target.shape[0] - num_tp
        return num_tp, num_fp, torch.tensor(0.0, device=out_dev)
    tp, label, npred = _counting.class_counts(input, target, num_classes)
    return tp.to(out_dev), (npred - tp).to(out_dev), label.to(out_dev)


def _precision_compute(
    num_tp: torch.Tensor, num_fp: torch.Tensor, num_label: torch.Tensor, average: Optional[str]
) -> torch.Tensor:
    if average in ("macro", "weighted"):
        # classes absent from both the predictions and the targets are ignored (reference :146-148)
        mask = (num_label != 0) | ((num_tp + num_fp) != 0)
        num_tp, num_fp = num_tp[mask], num_fp[mask]
    precision = num_tp / (num_tp + num_fp)
    if average in (None, "None") and torch.isnan(precision).sum():
        bad_class = torch.nonzero(torch.isnan(precision))
        logging.warning(
            f"{bad_class} classes have zero instances in both the "
            "predictions and the ground truth labels. Precision is still logged "
            "as zero."
        )
    precision = torch.nan_to_num(precision)
    if average == "micro":
        return precision
    elif average == "macro":
        return precision.mean()
    elif average == "weighted":
        return torch.inner(precision, (num_label[mask] / num_label.sum()))
    return precision


def _precision_param_check(num_classes: Optional[int], average: Optional[str]) -> None:
    average_options = ("micro", "macro", "weighted", "None", None)
    if average not in average_options:
        raise ValueError(f"`average` was not in the allowed value of {average_options}, got {average}.")
    if average != "micro" and (num_classes is None or num_classes <= 0):
        raise ValueError(
            f"num_classes should be a positive number when average={average}."
            f" Got num_classes={num_classes}."
        )


def _precision_update_input_check(input: torch.Tensor, target: torch.Tensor, num_classes: Optional[int]) -> None:
    if input.size(0) != target.size(0):
        raise ValueError(
            "The `input` and `target` should have the same first dimension, "
            f"got shapes {input.shape} and {target.shape}."
        )
    if target.ndim != 1:
        raise ValueError(f"target should be a one-dimensional tensor, got shape {target.shape}.")
    if not input.ndim == 1 and not (
        input.ndim == 2 and (num_classes is None or input.shape[1] == num_classes)
    ):
        raise ValueError(
            "input should have shape of (num_sample,) or (num_sample, num_classes), "
            f"got {input.shape}."
        )


def _binary_precision_update(
    input: torch.Tensor, target: torch.Tensor, threshold: float = 0.5
) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(sum(pred * target), sum(pred) - that, 0.0) for 0/1 targets (reference :206-221)."""
    _binary_precision_update_input_check(input, target)
    out_dev = input.device
    dev = compute_device(input)
    pred = torch.where(stage(input, dev) < threshold, 0, 1)
    tp, _, npred = _counting.class_counts(pred, stage(target, dev).to(torch.int64), 2)
    return tp[1].to(out_dev), (npred[1] - tp[1]).to(out_dev), torch.tensor(0.0, device=out_dev)


def _binary_precision_update_input_check(input: torch.Tensor, target: torch.Tensor) -> None:
    if input.shape != target.shape:
        raise ValueError(
            "The `input` and `target` should have the same dimensions, "
            f"got shapes {input.shape} and {target.shape}."
        )
    if target.ndim != 1:
        raise ValueError(f"target should be a one-dimensional tensor, got shape {target.shape}.")
