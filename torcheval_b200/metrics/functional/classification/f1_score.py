"""binary_f1_score / multiclass_f1_score with the reference's signatures
(torcheval/metrics/functional/classification/f1_score.py:16-273).  num_tp / num_label /
num_prediction are the diagonal, row sums and column sums of the confusion matrix and come from
the class-count kernel (csrc/cm.cu)."""
from __future__ import annotations

import logging
from typing import Optional, Tuple

import torch

from .... import _counting
from ...._ranking import compute_device, stage


@torch.inference_mode()
def binary_f1_score(input: torch.Tensor, target: torch.Tensor, *, threshold: float = 0.5) -> torch.Tensor:
    num_tp, num_label, num_prediction = _binary_f1_score_update(input, target, threshold)
    return _f1_score_compute(num_tp, num_label, num_prediction, "micro")


@torch.inference_mode()
def multiclass_f1_score(
    input: torch.Tensor,
    target: torch.Tensor,
    *,
    num_classes: Optional[int] = None,
    average: Optional[str] = "micro",
) -> torch.Tensor:
    _f1_score_param_check(num_classes, average)
    num_tp, num_label, num_prediction = _f1_score_update(input, target, num_classes, average)
    return _f1_score_compute(num_tp, num_label, num_prediction, average)


def _binary_f1_score_update(
    input: torch.Tensor, target: torch.Tensor, threshold: float = 0.5
) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(sum(pred * target), sum(target), sum(pred)) for 0/1 targets (reference :120-134)."""
    _binary_f1_score_update_input_check(input, target)
    out_dev = input.device
    dev = compute_device(input)
    pred = torch.where(stage(input, dev) < threshold, 0, 1)
    tp, label, npred = _counting.class_counts(pred, stage(target, dev).to(torch.int64), 2)
    return tp[1].to(out_dev), label[1].to(out_dev), npred[1].to(out_dev)


def _f1_score_update(
    input: torch.Tensor, target: torch.Tensor, num_classes: Optional[int], average: Optional[str]
) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    _f1_score_update_input_check(input, target, num_classes)
    out_dev = input.device
    if isinstance(average, str) and average == "micro":
        num_tp = _counting.match_count(input, target).to(out_dev)
        num_label = torch.tensor(target.shape[0], device=out_dev)
        return num_tp, num_label, num_label
    tp, label, npred = _counting.class_counts(input, target, num_classes)
    return tp.to(out_dev), label.to(out_dev), npred.to(out_dev)


def _f1_score_compute(
    num_tp: torch.Tensor, num_label: torch.Tensor, num_prediction: torch.Tensor, average: Optional[str]
) -> torch.Tensor:
    num_label_is_zero = num_label == 0
    if num_label_is_zero.any():
        logging.warning(
            "Warning: Some classes do not exist in the target. F1 scores for these classes will be cast to zeros."
        )
    if average in ("macro", "weighted"):
        # classes absent from both input and target are ignored
        mask = (~num_label_is_zero) | (num_prediction != 0)
        num_tp, num_label, num_prediction = num_tp[mask], num_label[mask], num_prediction[mask]
    precision = num_tp / num_prediction
    recall = num_tp / num_label
    f1 = torch.nan_to_num(2 * precision * recall / (precision + recall))
    if average == "micro":
        return f1
    elif average == "macro":
        return f1.mean()
    elif average == "weighted":
        return (f1 * (num_label / num_label.sum())).sum()
    return f1


def _f1_score_param_check(num_classes: Optional[int], average: Optional[str]) -> None:
    average_options = ("micro", "macro", "weighted", None)
    if average not in average_options:
        raise ValueError(f"`average` was not in the allowed value of {average_options}, got {average}.")
    if average != "micro" and (num_classes is None or num_classes <= 0):
        raise ValueError(
            f"num_classes should be a positive number when average={average}, "
            f"got num_classes={num_classes}."
        )


def _binary_f1_score_update_input_check(input: torch.Tensor, target: torch.Tensor) -> None:
    if input.ndim != 1:
        raise ValueError(f"input should be a one-dimensional tensor for binary f1 score, got shape {input.shape}.")
    if target.ndim != 1:
        raise ValueError(f"target should be a one-dimensional tensor for binary f1 score, got shape {target.shape}.")
    if input.shape != target.shape:
        raise ValueError(
            "The `input` and `target` should have the same dimensions, "
            f"got shapes {input.shape} and {target.shape}."
        )


def _f1_score_update_input_check(input: torch.Tensor, target: torch.Tensor, num_classes: Optional[int]) -> None:
    if input.size(0) != target.size(0):
        raise ValueError(
            "The `input` and `target` should have the same first dimension, "
            f"got shapes {input.shape} and {target.shape}."
        )
    if target.ndim != 1:
        raise ValueError("target should be a one-dimensional tensor, " f"got shape {target.shape}.")
    if not input.ndim == 1 and not (
        input.ndim == 2 and (num_classes is None or input.shape[1] == num_classes)
    ):
        raise ValueError(
            f"input should have shape of (num_sample,) or (num_sample, num_classes), "
            f"got {input.shape}."
        )
