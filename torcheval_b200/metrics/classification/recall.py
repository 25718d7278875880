"""BinaryRecall / MulticlassRecall with the reference's constructors and state names
(classification/recall.py:28-253): additive int64 counters."""
from __future__ import annotations

from typing import Optional, TypeVar

import torch

from ..functional.classification.recall import (
    _binary_recall_compute,
    _binary_recall_update,
    _recall_compute,
    _recall_param_check,
    _recall_update,
)
from ._counters import CounterMetric

T = TypeVar("T")


class BinaryRecall(CounterMetric):
    def __init__(self, *, threshold: float = 0.5, device: Optional[torch.device] = None) -> None:
        super().__init__(device=device)
        self.threshold = threshold
        self._add_counter("num_tp")
        self._add_counter("num_true_labels")

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        num_tp, num_true_labels = _binary_recall_update(input, target, self.threshold)
        self._accumulate(num_tp=num_tp, num_true_labels=num_true_labels)
        return self

    @torch.inference_mode()
    def compute(self) -> torch.Tensor:
        return _binary_recall_compute(self.num_tp, self.num_true_labels)


class MulticlassRecall(CounterMetric):
    def __init__(
        self,
        *,
        num_classes: Optional[int] = None,
        average: Optional[str] = "micro",
        device: Optional[torch.device] = None,
    ) -> None:
        super().__init__(device=device)
        _recall_param_check(num_classes, average)
        self.num_classes = num_classes
        self.average = average
        shape = () if average == "micro" else (num_classes,)
        for name in ("num_tp", "num_labels", "num_predictions"):
            self._add_counter(name, *shape)

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        num_tp, num_labels, num_predictions = _recall_update(input, target, self.num_classes, self.average)
        self._accumulate(num_tp=num_tp, num_labels=num_labels, num_predictions=num_predictions)
        return self

    @torch.inference_mode()
    def compute(self) -> torch.Tensor:
        return _recall_compute(self.num_tp, self.num_labels, self.num_predictions, self.average)
