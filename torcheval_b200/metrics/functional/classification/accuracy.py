"""binary_accuracy / multiclass_accuracy with the reference's signatures
(torcheval/metrics/functional/classification/accuracy.py:13-106, 250-386).  The k = 1 counting
(correct / total, overall or per class) runs on the sm_100a counting kernels (csrc/cm.cu:
match_count, class_counts); the top-k (k > 1) variant is a per-row rank, not a counting kernel,
and stays on torch ops."""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from .... import _counting
from ...._ranking import compute_device, stage


@torch.inference_mode()
def binary_accuracy(input: torch.Tensor, target: torch.Tensor, *, threshold: float = 0.5) -> torch.Tensor:
    num_correct, num_total = _binary_accuracy_update(input, target, threshold)
    return _accuracy_compute(num_correct, num_total, "micro")


@torch.inference_mode()
def multiclass_accuracy(
    input: torch.Tensor,
    target: torch.Tensor,
    *,
    average: Optional[str] = "micro",
    num_classes: Optional[int] = None,
    k: int = 1,
) -> torch.Tensor:
    _accuracy_param_check(average, num_classes, k)
    num_correct, num_total = _multiclass_accuracy_update(input, target, average, num_classes, k)
    return _accuracy_compute(num_correct, num_total, average)


def _multiclass_accuracy_update(
    input: torch.Tensor,
    target: torch.Tensor,
    average: Optional[str],
    num_classes: Optional[int],
    k: int,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """(num_correct, num_total): int64 scalars for micro, int64 (C,) otherwise (reference :250-278)."""
    _accuracy_update_input_check(input, target, num_classes, k)
    out_dev = input.device
    if k > 1:
        dev = compute_device(input)
        x, t = stage(input, dev), stage(target, dev)
        y_score = torch.gather(x, dim=-1, index=t.unsqueeze(dim=-1))
        rank = torch.gt(x, y_score).sum(dim=-1)
        mask = (rank < k).float()
        if average == "micro":
            return mask.sum().to(out_dev), torch.tensor(target.shape[0], device=out_dev)
        num_correct = mask.new_zeros(num_classes).scatter_add_(0, t, mask)
        num_total = t.new_zeros(num_classes).scatter_add_(0, t, torch.ones_like(t))
        return num_correct.to(out_dev), num_total.to(out_dev)
    if average == "micro":
        num_correct = _counting.match_count(input, target).to(out_dev)
        return num_correct, torch.tensor(target.shape[0], device=out_dev)
    num_tp, num_label, _ = _counting.class_counts(input, target, num_classes)
    return num_tp.to(out_dev), num_label.to(out_dev)


def _accuracy_compute(num_correct: torch.Tensor, num_total: torch.Tensor, average: Optional[str]) -> torch.Tensor:
    if isinstance(average, str) and average == "macro":
        mask = num_total != 0
        return (num_correct[mask] / num_total[mask]).mean()
    return num_correct / num_total


def _binary_accuracy_update(
    input: torch.Tensor, target: torch.Tensor, threshold: float = 0.5
) -> Tuple[torch.Tensor, torch.Tensor]:
    _binary_accuracy_update_input_check(input, target)
    out_dev = input.device
    dev = compute_device(input)
    pred = torch.where(stage(input, dev) < threshold, 0, 1)
    t = stage(target, dev)
    if t.is_floating_point():
        # the reference compares in the target's own dtype (accuracy.py:78: `input == target`): a
        # target that is not an integer value never equals a prediction
        ti = t.to(torch.int64)
        t = torch.where(ti.to(t.dtype) == t, ti, torch.full_like(ti, -1))
    num_correct = _counting.match_count(pred, t.to(torch.int64)).to(out_dev)
    return num_correct, torch.tensor(target.shape[0], dtype=torch.int64, device=out_dev)


def _accuracy_param_check(average: Optional[str], num_classes: Optional[int], k: int) -> None:
    average_options = ("micro", "macro", "none", None)
    if average not in average_options:
        raise ValueError(f"`average` was not in the allowed value of {average_options}, got {average}.")
    if average != "micro" and (num_classes is None or num_classes <= 0):
        raise ValueError(
            f"num_classes should be a positive number when average={average}."
            f" Got num_classes={num_classes}."
        )
    if type(k) != int:
        raise TypeError(f"Expected `k` to be an integer, but {type(k)} was provided.")
    if k < 1:
        raise ValueError(f"Expected `k` to be an integer greater than 0, but {k} was provided.")


def _accuracy_update_input_check(
    input: torch.Tensor, target: torch.Tensor, num_classes: Optional[int], k: int
) -> None:
    if input.size(0) != target.size(0):
        raise ValueError(
            "The `input` and `target` should have the same first dimension, "
            f"got shapes {input.shape} and {target.shape}."
        )
    if target.ndim != 1:
        raise ValueError(f"target should be a one-dimensional tensor, got shape {target.shape}.")
    if k > 1 and input.ndim != 2:
        raise ValueError(
            "input should have shape (num_sample, num_classes) for k > 1, "
            f"got shape {input.shape}."
        )
    if not input.ndim == 1 and not (
        input.ndim == 2 and (num_classes is None or input.shape[1] == num_classes)
    ):
        raise ValueError(
            "input should have shape of (num_sample,) or (num_sample, num_classes), "
            f"got {input.shape}."
        )


def _binary_accuracy_update_input_check(input: torch.Tensor, target: torch.Tensor) -> None:
    if input.shape != target.shape:
        raise ValueError(
            "The `input` and `target` should have the same dimensions, "
            f"got shapes {input.shape} and {target.shape}."
        )
    if target.ndim != 1:
        raise ValueError(f"target should be a one-dimensional tensor, got shape {target.shape}.")
