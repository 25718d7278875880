"""BinaryBinnedAUROC / MulticlassBinnedAUROC (reference: classification/binned_auroc.py:31-258).

State layout decision (SURVEY.md §8b open decision i, recorded in DESIGN.md): the reference's
``BinaryBinnedAUROC`` keeps every raw sample (``inputs`` / ``targets`` lists) and builds a
(T, tasks, N) boolean in ``compute``.  Here the state is what the metric actually needs — the
per-threshold counters ``num_tp`` / ``num_fp`` (tasks, T), int64 — so ``update`` does the work in
one streaming kernel pass, memory is O(T), ``merge_state`` is ``+=`` and a multi-GPU sync is one
all-reduce.  Values returned by ``compute`` are unchanged."""
from __future__ import annotations

from typing import Iterable, List, Optional, Tuple, TypeVar, Union

import torch

from ... import _binned
from ..functional.classification.binned_auroc import (
    DEFAULT_NUM_THRESHOLD,
    _binary_binned_auroc_param_check,
    _binary_binned_auroc_update_input_check,
    _multiclass_binned_auroc_compute,
    _multiclass_binned_auroc_param_check,
    _multiclass_binned_auroc_update_input_check,
)
from ..functional.tensor_utils import _create_threshold_tensor
from ..metric import Metric
from ._sample_store import SampleStoreMetric

T = TypeVar("T")


class BinaryBinnedAUROC(Metric[Tuple[torch.Tensor, torch.Tensor]]):
    def __init__(
        self,
        *,
        num_tasks: int = 1,
        threshold: Union[int, List[float], torch.Tensor] = DEFAULT_NUM_THRESHOLD,
        device: Optional[torch.device] = None,
    ) -> None:
        super().__init__(device=device)
        threshold = _create_threshold_tensor(threshold, self.device)
        _binary_binned_auroc_param_check(num_tasks, threshold)
        self.num_tasks = num_tasks
        self.threshold = threshold
        n_thr = len(threshold)
        self._add_state("num_tp", torch.zeros(num_tasks, n_thr, dtype=torch.int64, device=self.device))
        self._add_state("num_fp", torch.zeros(num_tasks, n_thr, dtype=torch.int64, device=self.device))

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        _binary_binned_auroc_update_input_check(input, target, self.num_tasks, self.threshold)
        tp, fp, _ = _binned.binary_counts(input, target, self.threshold)
        self.num_tp += tp.to(self.device)
        self.num_fp += fp.to(self.device)
        return self

    @torch.inference_mode()
    def compute(self) -> Tuple[torch.Tensor, torch.Tensor]:
        """(float64 (num_tasks,), thresholds) — as binary_binned_auroc."""
        return _binned.auroc_from_counts(self.num_tp, self.num_fp).to(self.device), self.threshold

    @torch.inference_mode()
    def merge_state(self: T, metrics: Iterable[T]) -> T:
        for metric in metrics:
            self.num_tp += metric.num_tp.to(self.device)
            self.num_fp += metric.num_fp.to(self.device)
        return self


class MulticlassBinnedAUROC(SampleStoreMetric):
    """Keeps raw samples like the reference (classification/binned_auroc.py:153-258); compute is
    the bug-compatible torch formulation (see functional/classification/binned_auroc.py)."""

    def __init__(
        self,
        *,
        num_classes: int,
        threshold: Union[int, List[float], torch.Tensor] = DEFAULT_NUM_THRESHOLD,
        average: Optional[str] = "macro",
        device: Optional[torch.device] = None,
    ) -> None:
        super().__init__(device=device)
        threshold = _create_threshold_tensor(threshold, self.device)
        _multiclass_binned_auroc_param_check(num_classes, threshold, average)
        self.num_classes = num_classes
        self.threshold = threshold
        self.average = average
        self._init_sample_states()

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        _multiclass_binned_auroc_update_input_check(input, target, self.num_classes)
        self._store(input, target)
        return self

    @torch.inference_mode()
    def compute(self) -> Tuple[torch.Tensor, torch.Tensor]:
        return _multiclass_binned_auroc_compute(
            self._all_inputs(), self._all_targets(), self.num_classes, self.threshold, self.average
        )
