"""binary_auprc / multiclass_auprc / multilabel_auprc with the reference's signatures
(torcheval/metrics/functional/classification/auprc.py:19-236).  Average precision = step
integration of the precision-recall curve over the distinct score thresholds, computed by the
ranking kernels (csrc/rank.cu); results are float32 like the reference's."""
from __future__ import annotations

from typing import Optional

import torch

from .... import _ranking


@torch.inference_mode()
def binary_auprc(input: torch.Tensor, target: torch.Tensor, *, num_tasks: int = 1) -> torch.Tensor:
    """1-D input -> shape (); 2-D input (even 1 x N) -> (num_tasks,) (reference auprc.py:239-251)."""
    _binary_auprc_update_input_check(input, target, num_tasks)
    return _binary_auprc_compute(input, target, num_tasks)


@torch.inference_mode()
def multiclass_auprc(
    input: torch.Tensor,
    target: torch.Tensor,
    num_classes: Optional[int] = None,
    *,
    average: Optional[str] = "macro",
) -> torch.Tensor:
    if num_classes is None:
        num_classes = input.shape[1]
    _multiclass_auprc_param_check(num_classes, average)
    _multiclass_auprc_update_input_check(input, target, num_classes)
    return _multiclass_auprc_compute(input, target, average)


@torch.inference_mode()
def multilabel_auprc(
    input: torch.Tensor,
    target: torch.Tensor,
    num_labels: Optional[int] = None,
    *,
    average: Optional[str] = "macro",
) -> torch.Tensor:
    if input.ndim != 2:
        raise ValueError(f"input should be a two-dimensional tensor, got shape {input.shape}.")
    if num_labels is None:
        num_labels = input.shape[1]
    _multilabel_auprc_param_check(num_labels, average)
    _multilabel_auprc_update_input_check(input, target, num_labels)
    return _multilabel_auprc_compute(input, target, num_labels, average)


def _binary_auprc_compute(input: torch.Tensor, target: torch.Tensor, num_tasks: int = 1) -> torch.Tensor:
    return _ranking.binary_auprc(input, target, num_tasks)


def _multiclass_auprc_compute(input: torch.Tensor, target: torch.Tensor, average: Optional[str] = "macro") -> torch.Tensor:
    auprcs = _ranking.multiclass_auprc(input, target).to(torch.float32).to(input.device)
    return torch.mean(auprcs) if average == "macro" else auprcs


def _multilabel_auprc_compute(
    input: torch.Tensor, target: torch.Tensor, num_labels: int, average: Optional[str] = "macro"
) -> torch.Tensor:
    auprcs = _ranking.multilabel_auprc(input, target).to(torch.float32).to(input.device)
    return torch.mean(auprcs) if average == "macro" else auprcs


def _binary_auprc_update_input_check(input: torch.Tensor, target: torch.Tensor, num_tasks: int) -> None:
    if input.shape != target.shape:
        raise ValueError(
            "The `input` and `target` should have the same shape, "
            f"got shapes {input.shape} and {target.shape}."
        )
    if num_tasks == 1:
        if (input.ndim == 2 and input.shape[0] > 1) or input.ndim > 2:
            raise ValueError(
                f"`num_tasks = 1`, `input` and `target` are expected to be one-dimensional tensors or 1xN tensors, but got shape input: {input.shape}, target: {target.shape}."
            )
    elif input.shape[0] != num_tasks:
        raise ValueError(
            f"`num_tasks = {num_tasks}`, `input` and `target` shape is expected to be ({num_tasks}, num_samples), but got shape input: {input.shape}, target: {target.shape}."
        )


def _average_check(average: Optional[str]) -> None:
    average_options = ("macro", "none", None)
    if average not in average_options:
        raise ValueError(f"`average` was not in the allowed value of {average_options}, got {average}.")


def _multiclass_auprc_param_check(num_classes: int, average: Optional[str]) -> None:
    _average_check(average)
    if num_classes < 2:
        raise ValueError("`num_classes` has to be at least 2.")


def _multiclass_auprc_update_input_check(input: torch.Tensor, target: torch.Tensor, num_classes: int) -> None:
    if input.size(0) != target.size(0):
        raise ValueError(
            "The `input` and `target` should have the same first dimension, "
            f"got shapes {input.shape} and {target.shape}."
        )
    if target.ndim != 1:
        raise ValueError("target should be a one-dimensional tensor, " f"got shape {target.shape}.")
    if not (input.ndim == 2 and input.shape[1] == num_classes):
        raise ValueError(
            f"input should have shape of (num_sample, num_classes), "
            f"got {input.shape} and num_classes={num_classes}."
        )


def _multilabel_auprc_param_check(num_labels: int, average: Optional[str]) -> None:
    _average_check(average)
    if num_labels < 2:
        raise ValueError("`num_labels` has to be at least 2.")


def _multilabel_auprc_update_input_check(input: torch.Tensor, target: torch.Tensor, num_labels: int) -> None:
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
