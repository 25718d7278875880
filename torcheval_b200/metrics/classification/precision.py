"""MulticlassPrecision / BinaryPrecision with the reference's constructors and state names
(classification/precision.py:27-222): ``num_tp`` / ``num_fp`` / ``num_label`` as additive int64
counters (one all-reduce to sync)."""
from __future__ import annotations

from typing import Optional, TypeVar

import torch

from ..functional.classification.precision import (
    _binary_precision_update,
    _precision_compute,
    _precision_param_check,
    _precision_update,
)
from ._counters import CounterMetric

T = TypeVar("T")


class MulticlassPrecision(CounterMetric):
    def __init__(
        self,
        *,
        num_classes: Optional[int] = None,
        average: Optional[str] = "micro",
        device: Optional[torch.device] = None,
    ) -> None:
        super().__init__(device=device)
        _precision_param_check(num_classes, average)
        self.num_classes = num_classes
        self.average = average
        shape = () if average == "micro" else (num_classes,)
        for name in ("num_tp", "num_fp", "num_label"):
            self._add_counter(name, *shape)

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        num_tp, num_fp, num_label = _precision_update(input, target, self.num_classes, self.average)
        self._accumulate(num_tp=num_tp, num_fp=num_fp, num_label=num_label)
        return self

    @torch.inference_mode()
    def compute(self) -> torch.Tensor:
        return _precision_compute(self.num_tp, self.num_fp, self.num_label, self.average)


class BinaryPrecision(MulticlassPrecision):
    def __init__(self, *, threshold: float = 0.5, device: Optional[torch.device] = None) -> None:
        super().__init__(num_classes=2, device=device)
        self.threshold = threshold

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        num_tp, num_fp, num_label = _binary_precision_update(input, target, self.threshold)
        self._accumulate(num_tp=num_tp, num_fp=num_fp, num_label=num_label)
        return self
