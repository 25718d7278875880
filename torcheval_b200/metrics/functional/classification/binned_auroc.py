"""binary_binned_auroc / multiclass_binned_auroc with the reference's signatures
(torcheval/metrics/functional/classification/binned_auroc.py:17-256)."""
from __future__ import annotations

from typing import List, Optional, Tuple, Union

import torch

from .... import _binned
from ...._ranking import compute_device, stage
from ..tensor_utils import THR_RANGE, THR_UNSORTED, _create_threshold_tensor, _threshold_faults

DEFAULT_NUM_THRESHOLD = 200


@torch.inference_mode()
def binary_binned_auroc(
    input: torch.Tensor,
    target: torch.Tensor,
    *,
    num_tasks: int = 1,
    threshold: Union[int, List[float], torch.Tensor] = DEFAULT_NUM_THRESHOLD,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """(auroc float64 of shape (num_tasks,) — (1,) for 1-D input —, thresholds).  The reference
    materialises a (T, tasks, N) boolean (binned_auroc.py:117-125); the count-identical
    (bucket, label) histogram is used here, so T = 10^4 with N = 2^30 is an ordinary call."""
    threshold = _create_threshold_tensor(threshold, target.device)
    _binary_binned_auroc_param_check(num_tasks, threshold)
    _binary_binned_auroc_update_input_check(input, target, num_tasks, threshold)
    return _binary_binned_auroc_compute(input, target, threshold)


def _binary_binned_auroc_compute(
    input: torch.Tensor, target: torch.Tensor, threshold: torch.Tensor
) -> Tuple[torch.Tensor, torch.Tensor]:
    tp, fp, _ = _binned.binary_counts(input, target, threshold)
    return _binned.auroc_from_counts(tp, fp).to(input.device), threshold


@torch.inference_mode()
def multiclass_binned_auroc(
    input: torch.Tensor,
    target: torch.Tensor,
    *,
    num_classes: int,
    threshold: Union[int, List[float], torch.Tensor] = DEFAULT_NUM_THRESHOLD,
    average: Optional[str] = "macro",
) -> Tuple[torch.Tensor, torch.Tensor]:
    threshold = _create_threshold_tensor(threshold, target.device)
    _multiclass_binned_auroc_param_check(num_classes, threshold, average)
    _multiclass_binned_auroc_update_input_check(input, target, num_classes)
    return _multiclass_binned_auroc_compute(input, target, num_classes, threshold, average)


def _multiclass_binned_auroc_compute(
    input: torch.Tensor,
    target: torch.Tensor,
    num_classes: int,
    threshold: torch.Tensor,
    average: Optional[str] = "macro",
) -> Tuple[torch.Tensor, torch.Tensor]:
    """Kept bug-compatible on purpose: the reference reduces over the
    CLASS axis (binned_auroc.py:196-205), so ``average=None`` yields one value per *sample*.
    Not a kernel target: a handful of torch ops on the compute device, O(T*N*C) memory like the
    reference."""
    dev = compute_device(input)
    x, t, thr = stage(input, dev), stage(target, dev), stage(threshold, dev)
    pred_label = x >= thr[:, None, None]
    onehot = torch.nn.functional.one_hot(t, num_classes)
    input_target = pred_label * onehot
    cum_tp = torch.nn.functional.pad(input_target.sum(dim=-1).rot90(1, [1, 0]), (1, 0), value=0.0)
    cum_fp = torch.nn.functional.pad(
        (pred_label.sum(dim=-1) - input_target.sum(dim=-1)).rot90(1, [1, 0]), (1, 0), value=0.0
    )
    factor = cum_tp[:, -1] * cum_fp[:, -1]
    auroc = torch.where(factor == 0, 0.5, torch.trapezoid(cum_tp, cum_fp, dim=1) / factor).to(input.device)
    if isinstance(average, str) and average == "macro":
        return auroc.mean(), threshold
    return auroc, threshold


def _threshold_range_check(threshold: torch.Tensor) -> None:
    faults = _threshold_faults(threshold)
    if faults is None:
        faults = (THR_UNSORTED if (torch.diff(threshold) < 0.0).any() else 0) | (
            THR_RANGE if (threshold < 0.0).any() or (threshold > 1.0).any() else 0
        )
    if faults & THR_UNSORTED:
        raise ValueError("The `threshold` should be a sorted tensor.")
    if faults & THR_RANGE:
        raise ValueError("The values in `threshold` should be in the range of [0, 1].")


def _binary_binned_auroc_param_check(num_tasks: int, threshold: torch.Tensor) -> None:
    if num_tasks < 1:
        raise ValueError("`num_tasks` has to be at least 1.")
    _threshold_range_check(threshold)


def _binary_binned_auroc_update_input_check(
    input: torch.Tensor, target: torch.Tensor, num_tasks: int, threshold: torch.Tensor
) -> None:
    if input.shape != target.shape:
        raise ValueError(
            "The `input` and `target` should have the same shape, "
            f"got shapes {input.shape} and {target.shape}."
        )
    if len(input.shape) > 2:
        raise ValueError(
            f"`input` is expected to be two dimensions or less, but got {len(input.shape)}D tensor."
        )
    if num_tasks == 1:
        if len(input.shape) > 1:
            raise ValueError(
                f"`num_tasks = 1`, `input` is expected to be one-dimensional tensor, but got shape {input.shape}."
            )
    elif len(input.shape) == 1 or input.shape[0] != num_tasks:
        raise ValueError(
            f"`num_tasks = {num_tasks}`, `input`'s shape is expected to be ({num_tasks}, num_samples), but got shape ({input.shape})."
        )


def _multiclass_binned_auroc_param_check(num_classes: int, threshold: torch.Tensor, average: Optional[str]) -> None:
    average_options = ("macro", "none", None)
    if average not in average_options:
        raise ValueError(f"`average` was not in the allowed value of {average_options}, got {average}.")
    if num_classes < 2:
        raise ValueError("`num_classes` has to be at least 2.")
    _threshold_range_check(threshold)


def _multiclass_binned_auroc_update_input_check(input: torch.Tensor, target: torch.Tensor, num_classes: int) -> None:
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
