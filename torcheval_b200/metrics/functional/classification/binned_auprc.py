"""binary / multiclass / multilabel binned AUPRC with the reference's signatures
(torcheval/metrics/functional/classification/binned_auprc.py:28-470): per-threshold counts from
the histogram kernels, step integration on the device, float32 results, NaN -> 0."""
from __future__ import annotations

from typing import List, Optional, Tuple, Union

import torch

from .... import _binned
from ..tensor_utils import THR_FIRST, THR_LAST, THR_RANGE, THR_UNSORTED, _create_threshold_tensor, _threshold_faults
from .binned_precision_recall_curve import _optimization_param_check

DEFAULT_NUM_THRESHOLD = 100


@torch.inference_mode()
def binary_binned_auprc(
    input: torch.Tensor,
    target: torch.Tensor,
    *,
    num_tasks: int = 1,
    threshold: Union[int, List[float], torch.Tensor] = DEFAULT_NUM_THRESHOLD,
) -> Tuple[torch.Tensor, torch.Tensor]:
    threshold = _create_threshold_tensor(threshold, target.device)
    _binary_binned_auprc_param_check(num_tasks, threshold)
    _binary_binned_auprc_update_input_check(input, target, num_tasks, threshold)
    return _binary_binned_auprc_compute(input, target, num_tasks, threshold), threshold


@torch.inference_mode()
def multiclass_binned_auprc(
    input: torch.Tensor,
    target: torch.Tensor,
    num_classes: Optional[int] = None,
    *,
    threshold: Union[int, List[float], torch.Tensor] = DEFAULT_NUM_THRESHOLD,
    average: Optional[str] = "macro",
    optimization: str = "vectorized",
) -> Tuple[torch.Tensor, torch.Tensor]:
    _optimization_param_check(optimization)
    threshold = _create_threshold_tensor(threshold, target.device)
    if num_classes is None and input.ndim == 2:
        num_classes = input.shape[1]
    _multiclass_binned_auprc_param_check(num_classes, threshold, average)
    _multiclass_binned_auprc_update_input_check(input, target, num_classes)
    return _multiclass_binned_auprc_compute(input, target, num_classes, threshold, average, optimization), threshold


@torch.inference_mode()
def multilabel_binned_auprc(
    input: torch.Tensor,
    target: torch.Tensor,
    num_labels: Optional[int] = None,
    *,
    threshold: Union[int, List[float], torch.Tensor] = DEFAULT_NUM_THRESHOLD,
    average: Optional[str] = "macro",
    optimization: str = "vectorized",
) -> Tuple[torch.Tensor, torch.Tensor]:
    _optimization_param_check(optimization)
    threshold = _create_threshold_tensor(threshold, target.device)
    if input.ndim != 2:
        raise ValueError(f"input should be a two-dimensional tensor, got shape {input.shape}.")
    if num_labels is None:
        num_labels = input.shape[1]
    _multilabel_binned_auprc_param_check(num_labels, threshold, average)
    _multilabel_binned_auprc_update_input_check(input, target, num_labels)
    return _multilabel_binned_auprc_compute(input, target, num_labels, threshold, average, optimization), threshold


# ------------------------------------------------------------------------------ computes
def _binary_binned_auprc_compute(
    input: torch.Tensor, target: torch.Tensor, num_tasks: int, threshold: torch.Tensor
) -> torch.Tensor:
    tp, fp, fn = _binned.binary_counts(input, target, threshold)
    res = _binned.auprc_from_counts(tp, fp, fn).to(torch.float32).to(input.device)
    if num_tasks == 1 and input.ndim == 1:
        return res[0]
    return res


def _average(auprcs: torch.Tensor, average: Optional[str]) -> torch.Tensor:
    return torch.mean(auprcs) if average == "macro" else auprcs


def _multiclass_binned_auprc_compute(
    input: torch.Tensor,
    target: torch.Tensor,
    num_classes: int,
    threshold: torch.Tensor,
    average: Optional[str],
    optimization: str = "vectorized",
) -> torch.Tensor:
    tp, fp, fn = _binned.counts_2d(input, target, threshold, multilabel=False)
    return _average(_binned.auprc_from_counts(tp, fp, fn).to(torch.float32).to(input.device), average)


def _multilabel_binned_auprc_compute(
    input: torch.Tensor,
    target: torch.Tensor,
    num_labels: int,
    threshold: torch.Tensor,
    average: Optional[str],
    optimization: str = "vectorized",
) -> torch.Tensor:
    tp, fp, fn = _binned.counts_2d(input, target, threshold, multilabel=True)
    return _average(_binned.auprc_from_counts(tp, fp, fn).to(torch.float32).to(input.device), average)


def _compute_riemann_integrals_from_counts(
    num_tp: torch.Tensor, num_fp: torch.Tensor, num_fn: torch.Tensor, average: Optional[str], device: torch.device
) -> torch.Tensor:
    """Class-state epilogue: counts are (T, C) like the reference's states
    (classification/binned_auprc.py); returns float32 () or (C,)."""
    res = _binned.auprc_from_counts(
        num_tp.T.contiguous().to(torch.int64), num_fp.T.contiguous().to(torch.int64), num_fn.T.contiguous().to(torch.int64)
    )
    return _average(res.to(torch.float32).to(device), average)


# ------------------------------------------------------------------------------ checks
def _auprc_threshold_check(threshold: torch.Tensor) -> None:
    if threshold.ndim != 1:
        raise ValueError(f"`threshold` should be 1-dimensional, but got {threshold.ndim}D tensor.")
    faults = _threshold_faults(threshold)
    if faults is None:  # empty: the reference's statements and their errors, one by one
        faults = (
            (THR_UNSORTED if (torch.diff(threshold) < 0.0).any() else 0)
            | (THR_RANGE if (threshold < 0.0).any() or (threshold > 1.0).any() else 0)
            | (THR_FIRST if threshold[0] != 0 else 0)
            | (THR_LAST if threshold[-1] != 1 else 0)
        )
    if faults & THR_UNSORTED:
        raise ValueError("The `threshold` should be a sorted tensor.")
    if faults & THR_RANGE:
        raise ValueError("The values in `threshold` should be in the range of [0, 1].")
    if faults & THR_FIRST:
        raise ValueError("First value in `threshold` should be 0.")
    if faults & THR_LAST:
        raise ValueError("Last value in `threshold` should be 1.")


def _binary_binned_auprc_param_check(num_tasks: int, threshold: torch.Tensor) -> None:
    if num_tasks < 1:
        raise ValueError("`num_tasks` has to be at least 1.")
    _auprc_threshold_check(threshold)


def _binary_binned_auprc_update_input_check(
    input: torch.Tensor, target: torch.Tensor, num_tasks: int, threshold: torch.Tensor
) -> None:
    if input.shape != target.shape:
        raise ValueError(
            "The `input` and `target` should have the same shape, "
            f"got shapes {input.shape} and {target.shape}."
        )
    elif num_tasks == 1:
        if input.ndim not in (1, 2):
            raise ValueError(
                f"`num_tasks = 1`, `input` is expected to be 1D or 2D tensor, but got shape {input.shape}."
            )
    else:
        if input.ndim != 2:
            raise ValueError(
                f"`num_tasks = {num_tasks}`, `input` is expected to be 2D tensor, but got shape {input.shape}."
            )
        elif input.shape[0] != num_tasks:
            raise ValueError(
                f"`num_tasks = {num_tasks}`, `input`'s shape is expected to be ({num_tasks}, num_samples), but got shape {input.shape}."
            )


def _multiclass_binned_auprc_param_check(num_classes: int, threshold: torch.Tensor, average: Optional[str]) -> None:
    average_options = ("macro", "none", None)
    if average not in average_options:
        raise ValueError(f"`average` was not in the allowed value of {average_options}, got {average}.")
    if num_classes < 2:
        raise ValueError("`num_classes` has to be at least 2.")
    _auprc_threshold_check(threshold)


def _multiclass_binned_auprc_update_input_check(input: torch.Tensor, target: torch.Tensor, num_classes: int) -> None:
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


def _multilabel_binned_auprc_param_check(num_labels: int, threshold: torch.Tensor, average: Optional[str]) -> None:
    average_options = ("macro", "none", None)
    if average not in average_options:
        raise ValueError(f"`average` was not in the allowed value of {average_options}, got {average}.")
    if num_labels < 2:
        raise ValueError("`num_labels` has to be at least 2.")
    _auprc_threshold_check(threshold)


def _multilabel_binned_auprc_update_input_check(input: torch.Tensor, target: torch.Tensor, num_labels: int) -> None:
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
