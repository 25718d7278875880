"""BinaryRecallAtFixedPrecision / MultilabelRecallAtFixedPrecision with the reference's
constructors and states (classification/recall_at_fixed_precision.py:29-195)."""
from __future__ import annotations

from typing import List, Optional, Tuple, TypeVar

import torch

from ..functional.classification.recall_at_fixed_precision import (
    _binary_recall_at_fixed_precision_compute,
    _binary_recall_at_fixed_precision_update_input_check,
    _multilabel_recall_at_fixed_precision_compute,
    _multilabel_recall_at_fixed_precision_update_input_check,
)
from ._sample_store import SampleStoreMetric

T = TypeVar("T")


class BinaryRecallAtFixedPrecision(SampleStoreMetric):
    def __init__(self, *, min_precision: float, device: Optional[torch.device] = None) -> None:
        super().__init__(device=device)
        self.min_precision = min_precision
        self._init_sample_states()

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        _binary_recall_at_fixed_precision_update_input_check(input, target, self.min_precision)
        self._store(input, target)
        return self

    @torch.inference_mode()
    def compute(self) -> Tuple[torch.Tensor, torch.Tensor]:
        return _binary_recall_at_fixed_precision_compute(self._all_inputs(), self._all_targets(), self.min_precision)


class MultilabelRecallAtFixedPrecision(SampleStoreMetric):
    def __init__(self, *, num_labels: int, min_precision: float, device: Optional[torch.device] = None) -> None:
        super().__init__(device=device)
        self.num_labels = num_labels
        self.min_precision = min_precision
        self._init_sample_states()

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        _multilabel_recall_at_fixed_precision_update_input_check(input, target, self.num_labels, self.min_precision)
        self._store(input, target)
        return self

    @torch.inference_mode()
    def compute(self) -> Tuple[List[torch.Tensor], List[torch.Tensor]]:
        return _multilabel_recall_at_fixed_precision_compute(
            self._all_inputs(), self._all_targets(), self.num_labels, self.min_precision
        )
