"""binary_confusion_matrix / multiclass_confusion_matrix with the reference's signatures
(torcheval/metrics/functional/classification/confusion_matrix.py:15-149); counting is done by the
sm_100a kernels of csrc/cm.cu."""
from __future__ import annotations

from typing import Optional

import torch

from .... import _counting


@torch.inference_mode()
def binary_confusion_matrix(
    input: torch.Tensor,
    target: torch.Tensor,
    *,
    threshold: float = 0.5,
    normalize: Optional[str] = None,
) -> torch.Tensor:
    """2x2 confusion matrix; rows are true classes, columns predicted (reference :15-65)."""
    _confusion_matrix_param_check(2, normalize)
    matrix = _binary_confusion_matrix_update(input, target, threshold)
    return _confusion_matrix_compute(matrix, normalize)


@torch.inference_mode()
def multiclass_confusion_matrix(
    input: torch.Tensor,
    target: torch.Tensor,
    num_classes: int,
    *,
    normalize: Optional[str] = None,
) -> torch.Tensor:
    """(C, C) confusion matrix of `input` ((n,) int64 classes or (n, C) scores -> argmax) against
    `target` (n,) int64 (reference :69-149).  int64 counts, or float32 when normalised."""
    _confusion_matrix_param_check(num_classes, normalize)
    matrix = _confusion_matrix_update(input, target, num_classes)
    return _confusion_matrix_compute(matrix, normalize)


def _binary_confusion_matrix_update(input: torch.Tensor, target: torch.Tensor, threshold: float = 0.5) -> torch.Tensor:
    _binary_confusion_matrix_update_input_check(input, target)
    pred = torch.where(input < threshold, 0, 1)  # reference :172
    return _confusion_matrix_update(pred, target, 2, _checked=True)


def _confusion_matrix_update(input: torch.Tensor, target: torch.Tensor, num_classes: int, _checked: bool = False) -> torch.Tensor:
    if not _checked:
        _confusion_matrix_update_input_check(input, target, num_classes)
    out_dev = input.device
    return _counting.confusion_matrix_counts(input, target, num_classes).to(out_dev)


def _confusion_matrix_compute(confusion_matrix: torch.Tensor, normalize: Optional[str]) -> torch.Tensor:
    # reference :197-209
    if normalize == "pred":
        return torch.nn.functional.normalize(confusion_matrix.to(torch.float), p=1, dim=0)
    elif normalize == "true":
        return torch.nn.functional.normalize(confusion_matrix.to(torch.float), p=1, dim=1)
    elif normalize == "all":
        return confusion_matrix.to(torch.float) / torch.sum(confusion_matrix)
    else:
        return confusion_matrix


def _confusion_matrix_param_check(num_classes: int, normalize: Optional[str]) -> None:
    if num_classes < 2:
        raise ValueError("Must be at least two classes for confusion matrix")
    if (normalize is not None) and (normalize not in ["all", "pred", "true", "none"]):
        raise ValueError("normalize must be one of 'all', 'pred', 'true', or 'none'.")


def _binary_confusion_matrix_update_input_check(input: torch.Tensor, target: torch.Tensor) -> None:
    if input.ndim != 1:
        raise ValueError(
            f"input should be a one-dimensional tensor for binary confusion matrix, got shape {input.shape}."
        )
    if target.ndim != 1:
        raise ValueError(
            f"target should be a one-dimensional tensor for binary confusion matrix, got shape {target.shape}."
        )
    if input.shape != target.shape:
        raise ValueError(
            "The `input` and `target` should have the same dimensions, "
            f"got shapes {input.shape} and {target.shape}."
        )


def _confusion_matrix_update_input_check(input: torch.Tensor, target: torch.Tensor, num_classes: Optional[int]) -> None:
    # shape checks as the reference (:247-268); the value-range checks (:271-282) are fused into
    # the counting kernel and raised by _counting.confusion_matrix_counts
    if input.size(0) != target.size(0):
        raise ValueError(
            "The `input` and `target` should have the same first dimension, "
            f"got shapes {input.shape} and {target.shape}."
        )
    if target.ndim != 1:
        raise ValueError("target should be a one-dimensional tensor, " f"got shape {target.shape}.")
    if not input.ndim == 1:
        if not (input.ndim == 2 and (input.shape[1] == num_classes)):
            raise ValueError(
                f"input should have shape of (num_sample,) or (num_sample, num_classes), "
                f"got {input.shape}."
            )
