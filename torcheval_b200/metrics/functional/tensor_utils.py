"""Small helpers shared by the binned metrics (reference: functional/tensor_utils.py:12-33)."""
from __future__ import annotations

from typing import List, Union

import torch


def _riemann_integral(x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    """Left Riemann sum of y over x for a decreasing x (recall): -sum((x[1:] - x[:-1]) * y[:-1])."""
    return -torch.sum((x[1:] - x[:-1]) * y[:-1])


# What can be wrong with a threshold tensor (the reference's param checks,
# binned_precision_recall_curve.py / binned_auprc.py / binned_auroc.py `_*_param_check`), as bits.
THR_UNSORTED, THR_RANGE, THR_FIRST, THR_LAST = 1, 2, 4, 8
_FAULTS_ATTR = "_te_threshold_faults"


def _faults_of(threshold: torch.Tensor) -> int:
    """The four checks as ONE device->host read (the reference does five, each a stream sync)."""
    t = threshold
    bad = torch.stack(
        [(torch.diff(t) < 0.0).any(), ((t < 0.0) | (t > 1.0)).any(), (t[0] != 0).reshape(()), (t[-1] != 1).reshape(())]
    )
    return sum(bit for bit, b in zip((THR_UNSORTED, THR_RANGE, THR_FIRST, THR_LAST), bad.tolist()) if b)


def _threshold_faults(threshold: torch.Tensor) -> int:
    """Bit mask of the failed checks.  Thresholds built by `_create_threshold_tensor` from an int or
    a list were checked on the host when they were made and cost nothing here; a tensor handed in
    by the caller costs one read.  None: not a case this helper takes (empty or not 1-D) — the
    caller runs the reference's statements one by one, with their own errors."""
    known = getattr(threshold, _FAULTS_ATTR, None)
    if known is not None:
        return known
    if threshold.ndim != 1 or threshold.numel() == 0:
        return None
    return _faults_of(threshold)


def _create_threshold_tensor(
    threshold: Union[int, List[float], torch.Tensor], device: torch.device
) -> torch.Tensor:
    """int n -> n evenly spaced values in [0, 1]; list -> tensor; tensor -> itself."""
    if isinstance(threshold, int) and not isinstance(threshold, bool):
        n = threshold
        threshold = torch.linspace(0, 1.0, n, device=device)
        if n >= 1:  # 0, ..., 1 by construction; a single point is [0.] and misses the closing 1
            setattr(threshold, _FAULTS_ATTR, 0 if n >= 2 else THR_LAST)
    elif isinstance(threshold, list):
        host = torch.tensor(threshold)
        faults = _threshold_faults(host)  # on the host copy: no device sync
        threshold = host.to(device)
        if faults is not None:
            setattr(threshold, _FAULTS_ATTR, faults)
    return threshold
