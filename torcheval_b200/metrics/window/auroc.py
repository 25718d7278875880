"""WindowedBinaryAUROC with the reference's constructor, states and window semantics
(torcheval/metrics/window/auroc.py:23-236): the last
``max_num_samples`` samples of every task live in float32 ring buffers ``inputs`` / ``targets`` /
``weights`` of shape (num_tasks, max_num_samples); ``compute`` ranks the filled part of the
window — many short rows, the small-N regime of the ranking kernels."""
from __future__ import annotations

from typing import Iterable, Optional, TypeVar

import torch

from ..functional.classification.auroc import _binary_auroc_compute, _binary_auroc_update_input_check
from ..metric import Metric

TAUROC = TypeVar("TAUROC")


class WindowedBinaryAUROC(Metric[torch.Tensor]):
    def __init__(
        self,
        *,
        num_tasks: int = 1,
        max_num_samples: int = 100,
        device: Optional[torch.device] = None,
    ) -> None:
        super().__init__(device=device)
        if num_tasks < 1:
            raise ValueError(
                "`num_tasks` value should be greater than and equal to 1, but received {num_tasks}. "
            )
        if max_num_samples < 1:
            raise ValueError(
                "`max_num_samples` value should be greater than and equal to 1, but received {max_num_samples}. "
            )
        self.num_tasks = num_tasks
        self._add_state("max_num_samples", max_num_samples)
        self.next_inserted = 0
        self._add_state("total_samples", 0)
        for name in ("inputs", "targets", "weights"):
            self._add_state(name, torch.zeros(self.num_tasks, self.max_num_samples, device=self.device))

    def _write(self, lo: int, hi: int, input, target, weight, src: slice) -> None:
        self.inputs[:, lo:hi] = input[:, src].detach()
        self.targets[:, lo:hi] = target[:, src].detach()
        self.weights[:, lo:hi] = weight[:, src].detach()

    @torch.inference_mode()
    def update(
        self: TAUROC, input: torch.Tensor, target: torch.Tensor, weight: Optional[torch.Tensor] = None
    ) -> TAUROC:
        if weight is None:
            weight = torch.ones_like(input, dtype=torch.double)
        _binary_auroc_update_input_check(input, target, self.num_tasks, weight)
        if input.ndim == 1:
            input, target, weight = input.reshape(1, -1), target.reshape(1, -1), weight.reshape(1, -1)
        n, cap = input.shape[1], self.max_num_samples
        if n >= cap:  # the update alone fills the window: keep its tail
            self._write(0, cap, input, target, weight, slice(n - cap, n))
            self.next_inserted = 0
        else:
            room = cap - self.next_inserted
            if n <= room:
                self._write(self.next_inserted, self.next_inserted + n, input, target, weight, slice(0, n))
                self.next_inserted += n
            else:  # wrap around
                self._write(self.next_inserted, cap, input, target, weight, slice(0, room))
                rest = n - room
                self._write(0, rest, input, target, weight, slice(n - rest, n))
                self.next_inserted = rest
        self.next_inserted %= cap
        self.total_samples += n
        return self

    @torch.inference_mode()
    def compute(self) -> torch.Tensor:
        """AUROC of the window; float64, () for one task else (num_tasks,).  A window whose
        not-yet-overwritten part is all zeros is taken as not full, like the reference (:154)."""
        if torch.all(self.inputs[:, self.next_inserted :] == 0):
            hi = self.next_inserted
            return _binary_auroc_compute(
                self.inputs[:, :hi].squeeze(), self.targets[:, :hi].squeeze(), self.weights[:, :hi].squeeze()
            )
        return _binary_auroc_compute(self.inputs.squeeze(), self.targets.squeeze(), self.weights.squeeze())

    @torch.inference_mode()
    def merge_state(self: TAUROC, metrics: Iterable[TAUROC]) -> TAUROC:
        """The merged window is the concatenation of the windows (its capacity grows)."""
        metrics = list(metrics)
        merged_cap = self.max_num_samples + sum(m.max_num_samples for m in metrics)
        old = (self.inputs, self.targets, self.weights)
        for name in ("inputs", "targets", "weights"):
            setattr(self, name, torch.zeros(self.num_tasks, merged_cap, device=self.device))
        size = min(self.total_samples, self.max_num_samples)
        self.inputs[:, :size], self.targets[:, :size], self.weights[:, :size] = (o[:, :size] for o in old)
        idx = size
        for m in metrics:
            size = min(m.total_samples, m.max_num_samples)
            self.inputs[:, idx : idx + size] = m.inputs[:, :size].to(self.device)
            self.targets[:, idx : idx + size] = m.targets[:, :size].to(self.device)
            self.weights[:, idx : idx + size] = m.weights[:, :size].to(self.device)
            self.total_samples += m.total_samples
            idx += size
        self.max_num_samples = merged_cap
        self.next_inserted = idx % merged_cap
        return self
