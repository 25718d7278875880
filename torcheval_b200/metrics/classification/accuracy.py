"""MulticlassAccuracy / BinaryAccuracy with the reference's constructors and state names
(classification/accuracy.py:32-212): ``num_correct`` / ``num_total``, scalars for micro
averaging, (num_classes,) otherwise."""
from __future__ import annotations

from typing import Optional, TypeVar

import torch

from ..functional.classification.accuracy import (
    _accuracy_compute,
    _accuracy_param_check,
    _binary_accuracy_update,
    _multiclass_accuracy_update,
)
from ._counters import CounterMetric

T = TypeVar("T")


class MulticlassAccuracy(CounterMetric):
    def __init__(
        self,
        *,
        average: Optional[str] = "micro",
        num_classes: Optional[int] = None,
        k: int = 1,
        device: Optional[torch.device] = None,
    ) -> None:
        super().__init__(device=device)
        _accuracy_param_check(average, num_classes, k)
        self.average = average
        self.num_classes = num_classes
        self.k = k
        shape = () if average == "micro" else (num_classes or 0,)
        # top-k correctness is a float mask in the reference; counts are integers either way
        self._add_counter("num_correct", *shape)
        self._add_counter("num_total", *shape)

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        num_correct, num_total = _multiclass_accuracy_update(input, target, self.average, self.num_classes, self.k)
        self._accumulate(num_correct=num_correct, num_total=num_total)
        return self

    @torch.inference_mode()
    def compute(self) -> torch.Tensor:
        """NaN before any update (0 / 0), as the reference."""
        return _accuracy_compute(self.num_correct, self.num_total, self.average)


class BinaryAccuracy(MulticlassAccuracy):
    def __init__(self, *, threshold: float = 0.5, device: Optional[torch.device] = None) -> None:
        super().__init__(device=device)
        self.threshold = threshold

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        num_correct, num_total = _binary_accuracy_update(input, target, self.threshold)
        self._accumulate(num_correct=num_correct, num_total=num_total)
        return self
