"""BinaryAUROC / MulticlassAUROC — stateful wrappers with the reference's constructor and state
contract (torcheval/metrics/classification/auroc.py:34-273): list states ``inputs / targets
(/ weights)``, ``update`` appends, ``compute`` concatenates and runs the ranking kernels.

One deliberate difference (SURVEY.md §8b open decision iii): when ``update`` gets no weight the
reference stores ``ones_like(input, dtype=double)`` — 8 bytes per sample of ones.  Here an
unweighted update stores an EMPTY weight tensor as a marker; the state names and types are
unchanged, and ones are materialised only if weighted and unweighted updates are mixed."""
from __future__ import annotations

from typing import Iterable, List, Optional, TypeVar

import torch

from ..functional.classification.auroc import (
    _binary_auroc_compute,
    _binary_auroc_update_input_check,
    _multiclass_auroc_compute,
    _multiclass_auroc_param_check,
    _multiclass_auroc_update_input_check,
)
from ..metric import Metric, cat_all

TAUROC = TypeVar("TAUROC")
TMulticlassAUROC = TypeVar("TMulticlassAUROC")


def _empty_weight_like(input: torch.Tensor) -> torch.Tensor:
    shape = list(input.shape)
    shape[-1] = 0
    return torch.empty(shape, dtype=torch.double, device=input.device)


def _resolve_weights(inputs: List[torch.Tensor], weights: List[torch.Tensor]) -> Optional[torch.Tensor]:
    """None if every update was unweighted, else the concatenated weights with ones filled in
    for the unweighted updates."""
    if all(w.shape[-1] == 0 for w in weights):
        return None
    if len(weights) == len(inputs):
        full = [
            w if w.shape[-1] == i.shape[-1] else torch.ones_like(i, dtype=torch.double)
            for i, w in zip(inputs, weights)
        ]
        return cat_all(full, -1)
    return cat_all(weights, -1)


class BinaryAUROC(Metric[torch.Tensor]):
    """Exact AUROC for binary classification, one value per task."""

    def __init__(
        self: TAUROC,
        *,
        num_tasks: int = 1,
        device: Optional[torch.device] = None,
        use_fbgemm: Optional[bool] = False,
    ) -> None:
        super().__init__(device=device)
        if num_tasks < 1:
            raise ValueError(
                "`num_tasks` value should be greater than or equal to 1, but received {num_tasks}. "
            )
        if use_fbgemm:
            # the reference forwards to fbgemm_gpu's approximate AUC kernel; it is not part of
            # this build (the exact kernel supersedes it)
            raise ValueError(
                "`use_fbgemm` is enabled but `fbgemm_gpu` is not found. Please "
                "install `fbgemm_gpu` to use this option."
            )
        self.num_tasks = num_tasks
        self._add_state("inputs", [])
        self._add_state("targets", [])
        self._add_state("weights", [])
        self.use_fbgemm = use_fbgemm

    @torch.inference_mode()
    def update(
        self: TAUROC,
        input: torch.Tensor,
        target: torch.Tensor,
        weight: Optional[torch.Tensor] = None,
    ) -> TAUROC:
        input = input.to(self.device)
        target = target.to(self.device)
        if weight is not None:
            weight = weight.to(self.device)
        _binary_auroc_update_input_check(input, target, self.num_tasks, weight)
        self.inputs.append(input)
        self.targets.append(target)
        self.weights.append(_empty_weight_like(input) if weight is None else weight)
        return self

    @torch.inference_mode()
    def compute(self: TAUROC) -> torch.Tensor:
        # like the reference, compute() before any update() raises (torch.cat of an empty list)
        return _binary_auroc_compute(
            cat_all(self.inputs, -1),
            cat_all(self.targets, -1),
            _resolve_weights(self.inputs, self.weights),
            self.use_fbgemm,
        )

    @torch.inference_mode()
    def merge_state(self: TAUROC, metrics: Iterable[TAUROC]) -> TAUROC:
        for metric in metrics:
            if metric.inputs:
                w = _resolve_weights(metric.inputs, metric.weights)
                metric_inputs = cat_all(metric.inputs, -1).to(self.device)
                self.inputs.append(metric_inputs)
                self.targets.append(cat_all(metric.targets, -1).to(self.device))
                self.weights.append(_empty_weight_like(metric_inputs) if w is None else w.to(self.device))
        return self

    @torch.inference_mode()
    def _prepare_for_merge_state(self: TAUROC) -> None:
        if self.inputs and self.targets:
            w = _resolve_weights(self.inputs, self.weights)
            self.inputs = [cat_all(self.inputs, -1)]
            self.targets = [cat_all(self.targets, -1)]
            self.weights = [_empty_weight_like(self.inputs[0]) if w is None else w]


class MulticlassAUROC(Metric[torch.Tensor]):
    """One-vs-rest AUROC for multiclass classification."""

    def __init__(
        self: TMulticlassAUROC,
        *,
        num_classes: int,
        average: Optional[str] = "macro",
        device: Optional[torch.device] = None,
    ) -> None:
        super().__init__(device=device)
        _multiclass_auroc_param_check(num_classes, average)
        self.num_classes = num_classes
        self.average = average
        self._add_state("inputs", [])
        self._add_state("targets", [])

    @torch.inference_mode()
    def update(self: TMulticlassAUROC, input: torch.Tensor, target: torch.Tensor) -> TMulticlassAUROC:
        input = input.to(self.device)
        target = target.to(self.device)
        _multiclass_auroc_update_input_check(input, target, self.num_classes)
        self.inputs.append(input)
        self.targets.append(target)
        return self

    @torch.inference_mode()
    def compute(self: TMulticlassAUROC) -> torch.Tensor:
        return _multiclass_auroc_compute(
            cat_all(self.inputs), cat_all(self.targets), self.num_classes, self.average
        )

    @torch.inference_mode()
    def merge_state(self: TMulticlassAUROC, metrics: Iterable[TMulticlassAUROC]) -> TMulticlassAUROC:
        for metric in metrics:
            if metric.inputs:
                self.inputs.append(cat_all(metric.inputs).to(self.device))
                self.targets.append(cat_all(metric.targets).to(self.device))
        return self

    @torch.inference_mode()
    def _prepare_for_merge_state(self: TMulticlassAUROC) -> None:
        if self.inputs and self.targets:
            self.inputs = [cat_all(self.inputs)]
            self.targets = [cat_all(self.targets)]
