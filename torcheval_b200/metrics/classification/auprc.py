"""BinaryAUPRC / MulticlassAUPRC / MultilabelAUPRC with the reference's constructors and states
(torcheval/metrics/classification/auprc.py:31-439)."""
from __future__ import annotations

from typing import Optional, TypeVar

import torch

from ..functional.classification.auprc import (
    _binary_auprc_compute,
    _binary_auprc_update_input_check,
    _multiclass_auprc_compute,
    _multiclass_auprc_param_check,
    _multiclass_auprc_update_input_check,
    _multilabel_auprc_compute,
    _multilabel_auprc_param_check,
    _multilabel_auprc_update_input_check,
)
from ._sample_store import SampleStoreMetric

T = TypeVar("T")


class BinaryAUPRC(SampleStoreMetric):
    """Average precision for binary classification; (num_tasks, n) or (n,) inputs."""

    _cat_dim = -1

    def __init__(self, *, num_tasks: int = 1, device: Optional[torch.device] = None) -> None:
        super().__init__(device=device)
        if num_tasks < 1:
            raise ValueError("`num_tasks` must be an integer greater than or equal to 1")
        self.num_tasks = num_tasks
        self._init_sample_states()

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        _binary_auprc_update_input_check(input, target, self.num_tasks)
        self._store(input, target)
        return self

    @torch.inference_mode()
    def compute(self) -> torch.Tensor:
        return _binary_auprc_compute(self._all_inputs(), self._all_targets(), self.num_tasks)


class MulticlassAUPRC(SampleStoreMetric):
    """One-vs-rest average precision; (n, num_classes) scores, (n,) class labels."""

    def __init__(self, *, num_classes: int, average: Optional[str] = "macro", device: Optional[torch.device] = None) -> None:
        super().__init__(device=device)
        _multiclass_auprc_param_check(num_classes, average)
        self.num_classes = num_classes
        self.average = average
        self._init_sample_states()

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        _multiclass_auprc_update_input_check(input, target, self.num_classes)
        self._store(input, target)
        return self

    @torch.inference_mode()
    def compute(self) -> torch.Tensor:
        return _multiclass_auprc_compute(self._all_inputs(), self._all_targets(), self.average)


class MultilabelAUPRC(SampleStoreMetric):
    """Per-label average precision; (n, num_labels) scores and 0/1 targets."""

    def __init__(self, *, num_labels: int, average: Optional[str] = "macro", device: Optional[torch.device] = None) -> None:
        super().__init__(device=device)
        _multilabel_auprc_param_check(num_labels, average)
        self.num_labels = num_labels
        self.average = average
        self._init_sample_states()

    @torch.inference_mode()
    def update(self: T, input: torch.Tensor, target: torch.Tensor) -> T:
        input, target = input.to(self.device), target.to(self.device)
        _multilabel_auprc_update_input_check(input, target, self.num_labels)
        self._store(input, target)
        return self

    @torch.inference_mode()
    def compute(self) -> torch.Tensor:
        return _multilabel_auprc_compute(self._all_inputs(), self._all_targets(), self.num_labels, self.average)
