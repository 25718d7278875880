"""Shared behaviour of the counter-state metrics (binned metrics, confusion matrix, accuracy,
F1): every state is an additive int64 tensor, so ``merge_state`` is ``+=`` over the registered
states (reference pattern, e.g. classification/binned_auprc.py:170-177) and a multi-GPU sync is a
single all-reduce (torcheval_b200.metrics.toolkit).

The reference registers these counters as float32 zeros (classification/confusion_matrix.py:
148-151, binned_auprc.py:94-105), which stop counting exactly at 2^24; int64 is used here
(SURVEY.md §8b open decision ii).  ``load_state_dict`` of a reference checkpoint still works:
float32 counters are accepted and keep accumulating in whatever dtype was loaded."""
from __future__ import annotations

from typing import Iterable, TypeVar

import torch

from ..metric import Metric

TSelf = TypeVar("TSelf")


class CounterMetric(Metric):
    #: tells toolkit.sync_and_compute that the registered tensor states simply add up
    _additive_states: bool = True

    def _add_counter(self, name: str, *shape: int) -> None:
        self._add_state(name, torch.zeros(tuple(shape), dtype=torch.int64, device=self.device))

    def _accumulate(self, **deltas: torch.Tensor) -> None:
        for name, delta in deltas.items():
            state = getattr(self, name)
            state += delta.to(device=state.device, dtype=state.dtype)

    @torch.inference_mode()
    def merge_state(self: TSelf, metrics: Iterable[TSelf]) -> TSelf:
        for metric in metrics:
            for name in self._state_name_to_default:
                state = getattr(self, name)
                state += getattr(metric, name).to(device=state.device, dtype=state.dtype)
        return self
