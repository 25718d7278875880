"""BinaryBinnedAUPRC / MulticlassBinnedAUPRC / MultilabelBinnedAUPRC with the reference's
constructors and state names (classification/binned_auprc.py:40-475): streaming counters
``num_tp / num_fp / num_fn`` of shape (tasks, T), (T, C), (T, L)."""
from __future__ import annotations

from typing import List, Optional, Tuple, TypeVar, Union

import torch

from ... import _binned
from ..functional.classification.binned_auprc import (
    DEFAULT_NUM_THRESHOLD,
    _binary_binned_auprc_param_check,
    _binary_binned_auprc_update_input_check,
    _compute_riemann_integrals_from_counts,
    _multiclass_binned_auprc_param_check,
    _multilabel_binned_auprc_param_check,
)
from ..functional.classification.binned_precision_recall_curve import (
    _multiclass_binned_precision_recall_curve_update,
    _multilabel_binned_precision_recall_curve_update,
    _optimization_param_check,
)
from ..functional.tensor_utils import _create_threshold_tensor
from ._counters import CounterMetric

T = TypeVar("T")


class BinaryBinnedAUPRC(CounterMetric):
    def __init__(
        self,
        *,
        num_tasks: int = 1,
        threshold: Union[int, List[float], torch.Tensor] = DEFAULT_NUM_THRESHOLD,
        device: Optional[torch.device] = None,
    ) -> None:
        super().__init__(device=device)
        threshold = _create_threshold_tensor(threshold, self.device)
        _binary_binned_auprc_param_check(num_tasks, threshold)
        self.num_tasks = num_tasks
        self.threshold = threshold
        for name in ("num_tp", "num_fp", "num_fn"):
            self._add_counter(name, num_tasks, len(threshold))

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        _binary_binned_auprc_update_input_check(input, target, self.num_tasks, self.threshold)
        tp, fp, fn = _binned.binary_counts(input, target, self.threshold)  # all tasks in one launch
        self._accumulate(num_tp=tp, num_fp=fp, num_fn=fn)
        return self

    @torch.inference_mode()
    def compute(self) -> Tuple[torch.Tensor, torch.Tensor]:
        """(float32 (num_tasks,), thresholds); NaN -> 0 (reference :143-167)."""
        res = _binned.auprc_from_counts(self.num_tp, self.num_fp, self.num_fn)
        return res.to(torch.float32).to(self.device), self.threshold


class _BinnedAUPRC2D(CounterMetric):
    _multilabel = False

    def _setup(self, width: int, threshold, average: Optional[str], optimization: str) -> None:
        self.threshold = threshold
        self.average = average
        self.optimization = optimization
        for name in ("num_tp", "num_fp", "num_fn"):
            self._add_counter(name, len(threshold), width)

    @torch.inference_mode()
    def compute(self) -> torch.Tensor:
        return _compute_riemann_integrals_from_counts(self.num_tp, self.num_fp, self.num_fn, self.average, self.device)


class MulticlassBinnedAUPRC(_BinnedAUPRC2D):
    def __init__(
        self,
        *,
        num_classes: int,
        threshold: Union[int, List[float], torch.Tensor] = DEFAULT_NUM_THRESHOLD,
        average: Optional[str] = "macro",
        optimization: str = "vectorized",
        device: Optional[torch.device] = None,
    ) -> None:
        super().__init__(device=device)
        _optimization_param_check(optimization)
        threshold = _create_threshold_tensor(threshold, self.device)
        _multiclass_binned_auprc_param_check(num_classes, threshold, average)
        self.num_classes = num_classes
        self._setup(num_classes, threshold, average, optimization)

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        tp, fp, fn = _multiclass_binned_precision_recall_curve_update(
            input, target, num_classes=self.num_classes, threshold=self.threshold, optimization=self.optimization
        )
        self._accumulate(num_tp=tp, num_fp=fp, num_fn=fn)
        return self


class MultilabelBinnedAUPRC(_BinnedAUPRC2D):
    def __init__(
        self,
        *,
        num_labels: int,
        threshold: Union[int, List[float], torch.Tensor] = DEFAULT_NUM_THRESHOLD,
        average: Optional[str] = "macro",
        optimization: str = "vectorized",
        device: Optional[torch.device] = None,
    ) -> None:
        super().__init__(device=device)
        _optimization_param_check(optimization)
        threshold = _create_threshold_tensor(threshold, self.device)
        _multilabel_binned_auprc_param_check(num_labels, threshold, average)
        self.num_labels = num_labels
        self._setup(num_labels, threshold, average, optimization)

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        tp, fp, fn = _multilabel_binned_precision_recall_curve_update(
            input, target, num_labels=self.num_labels, threshold=self.threshold, optimization=self.optimization
        )
        self._accumulate(num_tp=tp, num_fp=fp, num_fn=fn)
        return self
