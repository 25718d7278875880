"""binary_recall / multiclass_recall with the reference's signatures
(torcheval/metrics/functional/classification/recall.py:15-250): diagonal and row sums of the
confusion matrix from the class-count kernel (csrc/cm.cu)."""
from __future__ import annotations

import logging
from typing import Optional, Tuple

import torch

from .... import _counting
from ...._ranking import compute_device, stage


@torch.inference_mode()
def binary_recall(input: torch.Tensor, target: torch.Tensor, *, threshold: float = 0.5) -> torch.Tensor:
    """TP / (TP + FN) after ``torch.where(input < threshold, 0, 1)``."""
    num_tp, num_true_labels = _binary_recall_update(input, target, threshold)
    return _binary_recall_compute(num_tp, num_true_labels)


def _binary_recall_update(
    input: torch.Tensor, target: torch.Tensor, threshold: float = 0.5
) -> Tuple[torch.Tensor, torch.Tensor]:
    _binary_recall_update_input_check(input, target)
    if target.is_floating_point():
        # the reference computes `input & target` (:71) and fails the same way on float targets
        raise RuntimeError('"bitwise_and" is not implemented for floating point targets')
    out_dev = input.device
    dev = compute_device(input)
    pred = torch.where(stage(input, dev) < threshold, 0, 1)
    tp, label, _ = _counting.class_counts(pred, stage(target, dev).to(torch.int64), 2)
    return tp[1].to(out_dev), label[1].to(out_dev)


def _binary_recall_compute(num_tp: torch.Tensor, num_true_labels: torch.Tensor) -> torch.Tensor:
    recall = num_tp / num_true_labels
    if torch.isnan(recall):
        logging.warning("No positive instances have been seen in target. Recall is converted from NaN to 0s.")
        return torch.nan_to_num(recall)
    return recall


def _binary_recall_update_input_check(input: torch.Tensor, target: torch.Tensor) -> None:
    if input.shape != target.shape:
        raise ValueError(
            "The `input` and `target` should have the same dimensions, "
            f"got shapes {input.shape} and {target.shape}."
        )
    if target.ndim != 1:
        raise ValueError(f"target should be a one-dimensional tensor, got shape {target.shape}.")


@torch.inference_mode()
def multiclass_recall(
    input: torch.Tensor,
    target: torch.Tensor,
    *,
    num_classes: Optional[int] = None,
    average: Optional[str] = "micro",
) -> torch.Tensor:
    _recall_param_check(num_classes, average)
    num_tp, num_labels, num_predictions = _recall_update(input, target, num_classes, average)
    return _recall_compute(num_tp, num_labels, num_predictions, average)


def _recall_update(
    input: torch.Tensor, target: torch.Tensor, num_classes: Optional[int], average: Optional[str]
) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    _recall_update_input_check(input, target, num_classes)
    out_dev = input.device
    if isinstance(average, str) and average == "micro":
        num_tp = _counting.match_count(input, target).to(out_dev)
        num_labels = torch.tensor(target.numel(), device=out_dev)
        return num_tp, num_labels, num_labels
    assert isinstance(num_classes, int), f"`num_classes` must be an integer, but received {num_classes}."
    tp, label, npred = _counting.class_counts(input, target, num_classes)
    return tp.to(out_dev), label.to(out_dev), npred.to(out_dev)


def _recall_compute(
    num_tp: torch.Tensor, num_labels: torch.Tensor, num_predictions: torch.Tensor, average: Optional[str]
) -> torch.Tensor:
    if average in ("macro", "weighted"):
        # Classes with no samples in `target` and `input` are ignored.  The reference masks num_tp
        # only (:168-170) and then fails with a shape error whenever such a class exists; both
        # operands are masked here, which is what its docstring describes.
        mask = (num_labels != 0) | (num_predictions != 0)
        num_tp, num_labels = num_tp[mask], num_labels[mask]
    recall = num_tp / num_labels
    isnan_class = torch.isnan(recall)
    if isnan_class.any():
        nan_classes = isnan_class.nonzero(as_tuple=True)[0]
        logging.warning(
            f"One or more NaNs identified, as no ground-truth instances of "
            f"{nan_classes.tolist()} have been seen. These have been converted to zero."
        )
        recall = torch.nan_to_num(recall)
    if average == "micro":
        return recall
    elif average == "macro":
        return recall.mean()
    elif average == "weighted":
        return (recall * (num_labels / num_labels.sum())).sum()
    return recall


def _recall_param_check(num_classes: Optional[int], average: Optional[str]) -> None:
    average_options = ("micro", "macro", "weighted", None)
    if average not in average_options:
        raise ValueError(f"`average` was not in the allowed values of {average_options}, got {average}.")
    if average != "micro" and (num_classes is None or num_classes <= 0):
        raise ValueError(
            f"`num_classes` should be a positive number when average={average}, "
            f"got num_classes={num_classes}."
        )


def _recall_update_input_check(input: torch.Tensor, target: torch.Tensor, num_classes: Optional[int]) -> None:
    if input.size(0) != target.size(0):
        raise ValueError(
            f"The `input` and `target` should have the same first dimension, "
            f"got shapes {input.shape} and {target.shape}."
        )
    if target.ndim != 1:
        raise ValueError(f"`target` should be a one-dimensional tensor, got shape {target.shape}.")
    if input.ndim != 1 and not (input.ndim == 2 and (num_classes is None or input.shape[1] == num_classes)):
        raise ValueError(
            f"`input` should have shape (num_samples,) or (num_samples, num_classes), "
            f"got {input.shape}."
        )
