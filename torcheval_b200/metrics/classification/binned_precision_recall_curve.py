"""Binary / Multiclass / Multilabel BinnedPrecisionRecallCurve with the reference's constructors
and state names (classification/binned_precision_recall_curve.py:31-417)."""
from __future__ import annotations

from typing import List, Optional, Tuple, TypeVar, Union

import torch

from ..functional.classification.binned_precision_recall_curve import (
    _binary_binned_precision_recall_curve_compute,
    _binary_binned_precision_recall_curve_update,
    _binned_precision_recall_curve_param_check,
    _multiclass_binned_precision_recall_curve_compute,
    _multiclass_binned_precision_recall_curve_update,
    _multilabel_binned_precision_recall_curve_compute,
    _multilabel_binned_precision_recall_curve_update,
    _optimization_param_check,
)
from ._counters import CounterMetric

T = TypeVar("T")


def _threshold_tensor(threshold: Union[int, List[float], torch.Tensor], device: torch.device) -> torch.Tensor:
    # reference :70-74: linspace for an int, torch.tensor(...) for anything else
    if isinstance(threshold, int):
        return torch.linspace(0, 1.0, threshold, device=device)
    if isinstance(threshold, torch.Tensor):
        return threshold.clone().detach().to(device)
    return torch.tensor(threshold, device=device)


class BinaryBinnedPrecisionRecallCurve(CounterMetric):
    def __init__(
        self,
        *,
        threshold: Union[int, List[float], torch.Tensor] = 100,
        device: Optional[torch.device] = None,
    ) -> None:
        super().__init__(device=device)
        threshold = _threshold_tensor(threshold, self.device)
        _binned_precision_recall_curve_param_check(threshold)
        self.threshold = threshold
        for name in ("num_tp", "num_fp", "num_fn"):
            self._add_counter(name, len(threshold))

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        tp, fp, fn = _binary_binned_precision_recall_curve_update(input, target, self.threshold)
        self._accumulate(num_tp=tp, num_fp=fp, num_fn=fn)
        return self

    @torch.inference_mode()
    def compute(self) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        return _binary_binned_precision_recall_curve_compute(self.num_tp, self.num_fp, self.num_fn, self.threshold)


class MulticlassBinnedPrecisionRecallCurve(CounterMetric):
    def __init__(
        self,
        *,
        num_classes: int,
        threshold: Union[int, List[float], torch.Tensor] = 100,
        optimization: str = "vectorized",
        device: Optional[torch.device] = None,
    ) -> None:
        super().__init__(device=device)
        _optimization_param_check(optimization)
        threshold = _threshold_tensor(threshold, self.device)
        _binned_precision_recall_curve_param_check(threshold)
        self.num_classes = num_classes
        self.threshold = threshold
        self.optimization = optimization
        for name in ("num_tp", "num_fp", "num_fn"):
            self._add_counter(name, len(threshold), num_classes)

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        tp, fp, fn = _multiclass_binned_precision_recall_curve_update(
            input, target, num_classes=self.num_classes, threshold=self.threshold, optimization=self.optimization
        )
        self._accumulate(num_tp=tp, num_fp=fp, num_fn=fn)
        return self

    @torch.inference_mode()
    def compute(self) -> Tuple[List[torch.Tensor], List[torch.Tensor], torch.Tensor]:
        return _multiclass_binned_precision_recall_curve_compute(
            self.num_tp, self.num_fp, self.num_fn, num_classes=self.num_classes, threshold=self.threshold
        )


class MultilabelBinnedPrecisionRecallCurve(CounterMetric):
    def __init__(
        self,
        *,
        num_labels: int,
        threshold: Union[int, List[float], torch.Tensor] = 100,
        optimization: str = "vectorized",
        device: Optional[torch.device] = None,
    ) -> None:
        super().__init__(device=device)
        _optimization_param_check(optimization)
        threshold = _threshold_tensor(threshold, self.device)
        _binned_precision_recall_curve_param_check(threshold)
        self.num_labels = num_labels
        self.threshold = threshold
        self.optimization = optimization
        for name in ("num_tp", "num_fp", "num_fn"):
            self._add_counter(name, len(threshold), num_labels)

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        tp, fp, fn = _multilabel_binned_precision_recall_curve_update(
            input, target, num_labels=self.num_labels, threshold=self.threshold, optimization=self.optimization
        )
        self._accumulate(num_tp=tp, num_fp=fp, num_fn=fn)
        return self

    @torch.inference_mode()
    def compute(self) -> Tuple[List[torch.Tensor], List[torch.Tensor], torch.Tensor]:
        return _multilabel_binned_precision_recall_curve_compute(
            self.num_tp, self.num_fp, self.num_fn, num_labels=self.num_labels, threshold=self.threshold
        )
