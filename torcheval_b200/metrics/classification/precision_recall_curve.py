"""Binary / Multiclass / MultilabelPrecisionRecallCurve with the reference's constructors and
states (classification/precision_recall_curve.py:32-341): ``inputs`` / ``targets`` sample lists,
``compute`` runs the exact curve on the concatenation."""
from __future__ import annotations

from typing import List, Optional, Tuple, TypeVar

import torch

from ..functional.classification.precision_recall_curve import (
    _binary_precision_recall_curve_compute,
    _binary_precision_recall_curve_update,
    _multiclass_precision_recall_curve_compute,
    _multiclass_precision_recall_curve_update,
    _multilabel_precision_recall_curve_compute,
    _multilabel_precision_recall_curve_update,
)
from ._sample_store import SampleStoreMetric

T = TypeVar("T")


class BinaryPrecisionRecallCurve(SampleStoreMetric):
    def __init__(self, *, device: Optional[torch.device] = None) -> None:
        super().__init__(device=device)
        self._init_sample_states()

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        _binary_precision_recall_curve_update(input, target)
        self._store(input, target)
        return self

    @torch.inference_mode()
    def compute(self) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        return _binary_precision_recall_curve_compute(self._all_inputs(), self._all_targets())


class MulticlassPrecisionRecallCurve(SampleStoreMetric):
    def __init__(self, *, num_classes: Optional[int] = None, device: Optional[torch.device] = None) -> None:
        super().__init__(device=device)
        self.num_classes = num_classes
        self._init_sample_states()

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        _multiclass_precision_recall_curve_update(input, target, self.num_classes)
        self._store(input, target)
        return self

    @torch.inference_mode()
    def compute(self) -> Tuple[List[torch.Tensor], List[torch.Tensor], List[torch.Tensor]]:
        return _multiclass_precision_recall_curve_compute(self._all_inputs(), self._all_targets(), self.num_classes)


class MultilabelPrecisionRecallCurve(SampleStoreMetric):
    def __init__(self, *, num_labels: int, device: Optional[torch.device] = None) -> None:
        super().__init__(device=device)
        self.num_labels = num_labels
        self._init_sample_states()

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        _multilabel_precision_recall_curve_update(input, target, self.num_labels)
        self._store(input, target)
        return self

    @torch.inference_mode()
    def compute(self) -> Tuple[List[torch.Tensor], List[torch.Tensor], List[torch.Tensor]]:
        return _multilabel_precision_recall_curve_compute(self._all_inputs(), self._all_targets(), self.num_labels)
