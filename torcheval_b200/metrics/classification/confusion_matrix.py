"""MulticlassConfusionMatrix / BinaryConfusionMatrix with the reference's constructors
(classification/confusion_matrix.py:26-314; note the positional ``num_classes``).  The state
``confusion_matrix`` is an exact int64 counter (the reference keeps float32); ``compute()`` and
``normalized()`` return float32 exactly like the reference's class API does."""
from __future__ import annotations

from typing import Optional, TypeVar

import torch

from ..functional.classification.confusion_matrix import (
    _binary_confusion_matrix_update,
    _confusion_matrix_compute,
    _confusion_matrix_param_check,
    _confusion_matrix_update,
)
from ._counters import CounterMetric

T = TypeVar("T")


class MulticlassConfusionMatrix(CounterMetric):
    def __init__(
        self,
        num_classes: int,
        *,
        normalize: Optional[str] = None,
        device: Optional[torch.device] = None,
    ) -> None:
        super().__init__(device=device)
        _confusion_matrix_param_check(num_classes, normalize)
        self.normalize = normalize
        self.num_classes = num_classes
        self._add_counter("confusion_matrix", num_classes, num_classes)

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        # the kernel counts into a fresh matrix: a label out of range raises before the state changes
        self._accumulate(confusion_matrix=_confusion_matrix_update(input, target, self.num_classes))
        return self

    @torch.inference_mode()
    def compute(self) -> torch.Tensor:
        return _confusion_matrix_compute(self.confusion_matrix.to(torch.float32), normalize=self.normalize)

    @torch.inference_mode()
    def normalized(self, normalize: Optional[str] = None) -> torch.Tensor:
        _confusion_matrix_param_check(self.num_classes, normalize)
        return _confusion_matrix_compute(self.confusion_matrix.to(torch.float32), normalize)


class BinaryConfusionMatrix(MulticlassConfusionMatrix):
    def __init__(
        self,
        *,
        threshold: float = 0.5,
        normalize: Optional[str] = None,
        device: Optional[torch.device] = None,
    ) -> None:
        super().__init__(num_classes=2, device=device, normalize=normalize)
        self.threshold = threshold

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        self._accumulate(confusion_matrix=_binary_confusion_matrix_update(input, target, self.threshold))
        return self
