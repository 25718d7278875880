"""Host-side dispatch for the binned (thresholded) metrics: staging, the (bucket, label)
histogram kernels, suffix counts and the small O(T) epilogues."""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _native as N
from . import ops
from ._ranking import _floor_f32_scores, compute_device, stage


def _labels_i64(t: torch.Tensor) -> torch.Tensor:
    return t if t.dtype == torch.int64 else t.to(torch.int64)


def _thr_f32(threshold: torch.Tensor, dev: torch.device) -> torch.Tensor:
    return stage(threshold, dev, torch.float32).contiguous()


def binary_counts(input: torch.Tensor, target: torch.Tensor, threshold: torch.Tensor):
    """(num_tp, num_fp, num_fn): int64 (tasks, T) on the compute device; input is (n,) or (tasks, n)."""
    dev = compute_device(input)
    x = _floor_f32_scores(stage(input, dev))
    t = _labels_i64(stage(target, dev))
    thr = _thr_f32(threshold, dev)
    T = thr.numel()
    tasks = 1 if x.dim() == 1 else x.shape[0]
    hist = torch.zeros((tasks, T + 1, 2), dtype=torch.int64, device=dev)
    ops.binned_hist_update_(hist, x, t, thr)
    return ops.binned_counts(hist, T)


def counts_2d(input: torch.Tensor, target: torch.Tensor, threshold: torch.Tensor, multilabel: bool):
    """(num_tp, num_fp, num_fn): int64 (C, T) for (n, C) scores, one-vs-rest per class / per label."""
    dev = compute_device(input)
    x = _floor_f32_scores(stage(input, dev))
    t = _labels_i64(stage(target, dev))
    thr = _thr_f32(threshold, dev)
    T, C = thr.numel(), x.shape[1]
    hist = torch.zeros((C, T + 1, 2), dtype=torch.int64, device=dev)
    ops.binned_hist2d_update_(hist, x, t, thr, multilabel)
    return ops.binned_counts(hist, T)


def auroc_from_counts(num_tp: torch.Tensor, num_fp: torch.Tensor) -> torch.Tensor:
    """float64 (tasks,) — trapezoid over the thresholds, 0.5 when degenerate."""
    dev = compute_device(num_tp)
    tp = stage(num_tp, dev, torch.int64).contiguous()
    fp = stage(num_fp, dev, torch.int64).contiguous()
    return ops.binned_integrate(tp, fp, None, N.TE_BINNED_AUROC)


def auprc_from_counts(num_tp: torch.Tensor, num_fp: torch.Tensor, num_fn: torch.Tensor) -> torch.Tensor:
    """float64 (tasks,) — step integration, NaN -> 0."""
    dev = compute_device(num_tp)
    tp = stage(num_tp, dev, torch.int64).contiguous()
    fp = stage(num_fp, dev, torch.int64).contiguous()
    fn = stage(num_fn, dev, torch.int64).contiguous()
    return ops.binned_integrate(tp, fp, fn, N.TE_BINNED_AUPRC)


def pr_curve_from_counts(num_tp: torch.Tensor, num_fp: torch.Tensor, num_fn: torch.Tensor, dim: int = 0):
    """precision / recall with the trailing (1, 0) point (binned_precision_recall_curve.py:113-129),
    float32 like the reference (int64 / int64)."""
    precision = torch.nan_to_num(num_tp / (num_tp + num_fp), 1.0)
    recall = num_tp / (num_tp + num_fn)
    shape = list(precision.shape)
    shape[dim] = 1
    precision = torch.cat([precision, precision.new_ones(shape)], dim=dim)
    recall = torch.cat([recall, recall.new_zeros(shape)], dim=dim)
    return precision, recall
