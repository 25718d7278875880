"""MulticlassF1Score / BinaryF1Score with the reference's constructors and state names
(classification/f1_score.py:26-227): ``num_tp`` / ``num_label`` / ``num_prediction``."""
from __future__ import annotations

from typing import Optional, TypeVar

import torch

from ..functional.classification.f1_score import (
    _binary_f1_score_update,
    _f1_score_compute,
    _f1_score_param_check,
    _f1_score_update,
)
from ._counters import CounterMetric

T = TypeVar("T")


class MulticlassF1Score(CounterMetric):
    def __init__(
        self,
        *,
        num_classes: Optional[int] = None,
        average: Optional[str] = "micro",
        device: Optional[torch.device] = None,
    ) -> None:
        super().__init__(device=device)
        _f1_score_param_check(num_classes, average)
        self.num_classes = num_classes
        self.average = average
        shape = () if average == "micro" else (num_classes,)
        for name in ("num_tp", "num_label", "num_prediction"):
            self._add_counter(name, *shape)

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        num_tp, num_label, num_prediction = _f1_score_update(input, target, self.num_classes, self.average)
        self._accumulate(num_tp=num_tp, num_label=num_label, num_prediction=num_prediction)
        return self

    @torch.inference_mode()
    def compute(self) -> torch.Tensor:
        """0 before any update (NaN -> 0), as the reference."""
        return _f1_score_compute(self.num_tp, self.num_label, self.num_prediction, self.average)


class BinaryF1Score(MulticlassF1Score):
    def __init__(self, *, threshold: float = 0.5, device: Optional[torch.device] = None) -> None:
        super().__init__(average="micro", device=device)
        self.threshold = threshold

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        num_tp, num_label, num_prediction = _binary_f1_score_update(input, target, self.threshold)
        self._accumulate(num_tp=num_tp, num_label=num_label, num_prediction=num_prediction)
        return self
