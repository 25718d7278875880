"""Shared behaviour of the exact (sample-keeping) metrics: ``inputs`` / ``targets`` list states,
append on update, concatenate on merge / prepare (reference pattern, e.g.
classification/auprc.py:104-151).  The concat axis is -1 for (tasks, n) layouts and 0 for
(n, classes) layouts."""
from __future__ import annotations

from typing import Iterable, TypeVar

import torch

from ..metric import Metric, cat_all

TSelf = TypeVar("TSelf")


class SampleStoreMetric(Metric[torch.Tensor]):
    _cat_dim: int = 0

    def _init_sample_states(self) -> None:
        self._add_state("inputs", [])
        self._add_state("targets", [])

    def _store(self, input: torch.Tensor, target: torch.Tensor) -> None:
        self.inputs.append(input)
        self.targets.append(target)

    def _all_inputs(self) -> torch.Tensor:
        return cat_all(self.inputs, self._cat_dim)

    def _all_targets(self) -> torch.Tensor:
        return cat_all(self.targets, self._cat_dim)

    @torch.inference_mode()
    def merge_state(self: TSelf, metrics: Iterable[TSelf]) -> TSelf:
        for metric in metrics:
            if metric.inputs:
                self.inputs.append(cat_all(metric.inputs, self._cat_dim).to(self.device))
                self.targets.append(cat_all(metric.targets, self._cat_dim).to(self.device))
        return self

    @torch.inference_mode()
    def _prepare_for_merge_state(self) -> None:
        if self.inputs and self.targets:
            self.inputs = [self._all_inputs()]
            self.targets = [self._all_targets()]
