"""Small helpers shared by the binned metrics (reference: functional/tensor_utils.py:12-33)."""
from __future__ import annotations

from typing import List, Union

import torch


def _riemann_integral(x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    """Left Riemann sum of y over x for a decreasing x (recall): -sum((x[1:] - x[:-1]) * y[:-1])."""
    return -torch.sum((x[1:] - x[:-1]) * y[:-1])


def _create_threshold_tensor(
    threshold: Union[int, List[float], torch.Tensor], device: torch.device
) -> torch.Tensor:
    """int n -> n evenly spaced values in [0, 1]; list -> tensor; tensor -> itself."""
    if isinstance(threshold, int):
        threshold = torch.linspace(0, 1.0, threshold, device=device)
    elif isinstance(threshold, list):
        threshold = torch.tensor(threshold, device=device)
    return threshold
