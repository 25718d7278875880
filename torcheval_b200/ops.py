"""Tensor-level entry points: torch owns device memory and streams, libte_b200.so does the work.

Every function takes CUDA tensors, enqueues on ``torch.cuda.current_stream()`` of the tensors'
device and returns CUDA tensors.  Nothing here computes a metric in PyTorch; if the native
library is missing these functions raise (see ``_native.load``).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _native as N

_DTYPE_CODE = {
    torch.int64: N.TE_DTYPE_I64,
    torch.int32: N.TE_DTYPE_I32,
    torch.uint8: N.TE_DTYPE_U8,
    torch.bool: N.TE_DTYPE_U8,
    torch.float32: N.TE_DTYPE_F32,
    torch.float64: N.TE_DTYPE_F64,
}


def _require_cuda(*tensors: torch.Tensor) -> torch.device:
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError(
                "torcheval_b200 kernels run on CUDA tensors only (no CPU fallback); "
                f"got a tensor on {t.device}"
            )
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError(f"tensors on different devices: {dev} and {t.device}")
    assert dev is not None
    return dev


def _stream_ptr(dev: torch.device) -> int:
    return torch.cuda.current_stream(dev).cuda_stream


def _scratch(nbytes: int, dev: torch.device) -> torch.Tensor:
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=dev)


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


# ----------------------------------------------------------------------------- K4
def cm_update_(
    cm: torch.Tensor,
    pred: torch.Tensor,
    target: torch.Tensor,
    num_classes: int,
    variant: int = N.TE_CM_AUTO,
) -> torch.Tensor:
    """cm (int64 [C, C], updated in place) += counts of (target, pred).  Returns the int64[4]
    device tensor {min_pred, max_pred, min_target, max_target} over out-of-range entries."""
    lib = N.load()
    dev = _require_cuda(cm, pred, target)
    assert cm.dtype == torch.int64 and cm.is_contiguous() and cm.numel() == num_classes * num_classes
    assert pred.dtype == torch.int64 and target.dtype == torch.int64
    pred, target = pred.contiguous(), target.contiguous()
    rng = torch.empty(4, dtype=torch.int64, device=dev)
    sb = lib.te_cm_scratch_bytes_n(num_classes, variant, pred.numel())
    if variant == N.TE_CM_BUCKETED and sb == 0:
        raise ValueError(f"TE_CM_BUCKETED does not apply to num_classes={num_classes}")
    scratch = _scratch(sb, dev)
    with torch.cuda.device(dev):
        N.check(
            "te_cm_update",
            lib.te_cm_update(
                pred.data_ptr(), target.data_ptr(), pred.numel(), num_classes, cm.data_ptr(),
                rng.data_ptr(), scratch.data_ptr(), scratch.numel(), variant, _stream_ptr(dev),
            ),
        )
    return rng


def cm_update_logits_(
    cm: torch.Tensor, logits: torch.Tensor, target: torch.Tensor, num_classes: int
) -> torch.Tensor:
    lib = N.load()
    dev = _require_cuda(cm, logits, target)
    assert cm.dtype == torch.int64 and cm.is_contiguous()
    assert logits.dtype == torch.float32 and logits.dim() == 2 and target.dtype == torch.int64
    if logits.stride(1) != 1:
        logits = logits.contiguous()
    target = target.contiguous()
    rng = torch.empty(4, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        N.check(
            "te_cm_update_logits",
            lib.te_cm_update_logits(
                logits.data_ptr(), logits.stride(0), target.data_ptr(), logits.shape[0], num_classes,
                cm.data_ptr(), rng.data_ptr(), _stream_ptr(dev),
            ),
        )
    return rng


def class_counts_(counts3: torch.Tensor, input: torch.Tensor, target: torch.Tensor, num_classes: int) -> torch.Tensor:
    """counts3 (int64 [3, C], in place) += (num_tp, num_label, num_pred); input is (n,) int64
    classes or (n, C) float32 scores (argmax'ed in the kernel).  Returns the int64[4] range report."""
    lib = N.load()
    dev = _require_cuda(counts3, input, target)
    assert counts3.dtype == torch.int64 and counts3.is_contiguous() and counts3.numel() == 3 * num_classes
    assert target.dtype == torch.int64
    target = target.contiguous()
    rng = torch.empty(4, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        if input.dim() == 2:
            assert input.dtype == torch.float32
            if input.stride(1) != 1:
                input = input.contiguous()
            rc = lib.te_class_counts(None, input.data_ptr(), input.stride(0), target.data_ptr(), input.shape[0],
                                     num_classes, counts3.data_ptr(), rng.data_ptr(), _stream_ptr(dev))
        else:
            assert input.dtype == torch.int64
            input = input.contiguous()
            rc = lib.te_class_counts(input.data_ptr(), None, 0, target.data_ptr(), input.numel(), num_classes,
                                     counts3.data_ptr(), rng.data_ptr(), _stream_ptr(dev))
        N.check("te_class_counts", rc)
    return rng


def match_count(pred: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """int64 scalar tensor: number of positions where pred == target."""
    lib = N.load()
    dev = _require_cuda(pred, target)
    assert pred.dtype == torch.int64 and target.dtype == torch.int64 and pred.shape == target.shape
    pred, target = pred.contiguous(), target.contiguous()
    out = torch.zeros((), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        N.check("te_match_count", lib.te_match_count(pred.data_ptr(), target.data_ptr(), pred.numel(), out.data_ptr(), _stream_ptr(dev)))
    return out


# ----------------------------------------------------------------------------- K3
def binned_hist_update_(
    hist: torch.Tensor, x: torch.Tensor, target: torch.Tensor, threshold: torch.Tensor
) -> None:
    """hist (int64 [tasks, T+1, 2], in place) += (bucket, label) counts of x/target
    ((n,) or (tasks, n), float32 / int64)."""
    lib = N.load()
    dev = _require_cuda(hist, x, target, threshold)
    T = threshold.numel()
    assert x.dtype == torch.float32 and target.dtype == torch.int64
    assert threshold.dtype == torch.float32 and threshold.is_contiguous()
    x, target = x.contiguous(), target.contiguous()
    tasks = 1 if x.dim() == 1 else x.shape[0]
    n = x.shape[-1]
    assert hist.dtype == torch.int64 and hist.is_contiguous() and hist.numel() == tasks * (T + 1) * 2
    scratch = _scratch(lib.te_binned_scratch_bytes(T), dev)
    with torch.cuda.device(dev):
        N.check(
            "te_binned_hist_update",
            lib.te_binned_hist_update(
                x.data_ptr(), n, target.data_ptr(), n, n, tasks, threshold.data_ptr(), T,
                hist.data_ptr(), scratch.data_ptr(), scratch.numel(), _stream_ptr(dev),
            ),
        )


def binned_hist2d_update_(hist: torch.Tensor, x: torch.Tensor, target: torch.Tensor, threshold: torch.Tensor, multilabel: bool) -> None:
    """hist (int64 [C, T+1, 2], in place) += counts for (n, C) scores; target (n,) (multiclass,
    label = target == c) or (n, C) (multilabel)."""
    lib = N.load()
    dev = _require_cuda(hist, x, target, threshold)
    T = threshold.numel()
    assert x.dim() == 2 and x.dtype == torch.float32 and target.dtype == torch.int64
    assert threshold.dtype == torch.float32 and threshold.is_contiguous()
    x, target = x.contiguous(), target.contiguous()
    n, C = x.shape
    assert target.shape == ((n, C) if multilabel else (n,))
    assert hist.dtype == torch.int64 and hist.is_contiguous() and hist.numel() == C * (T + 1) * 2
    scratch = _scratch(lib.te_binned_scratch_bytes(T), dev)
    with torch.cuda.device(dev):
        N.check(
            "te_binned_hist2d_update",
            lib.te_binned_hist2d_update(
                x.data_ptr(), target.data_ptr(), n, C, 1 if multilabel else 0, threshold.data_ptr(), T,
                hist.data_ptr(), scratch.data_ptr(), scratch.numel(), _stream_ptr(dev),
            ),
        )


def binned_counts(hist: torch.Tensor, T: int, want_fn: bool = True):
    """(num_tp, num_fp, num_fn) int64 [tasks, T] from the (bucket, label) histogram."""
    lib = N.load()
    dev = _require_cuda(hist)
    tasks = hist.numel() // ((T + 1) * 2)
    tp = torch.empty((tasks, T), dtype=torch.int64, device=dev)
    fp = torch.empty_like(tp)
    fn = torch.empty_like(tp) if want_fn else None
    with torch.cuda.device(dev):
        N.check(
            "te_binned_counts",
            lib.te_binned_counts(hist.data_ptr(), T, tasks, tp.data_ptr(), fp.data_ptr(), _ptr(fn), _stream_ptr(dev)),
        )
    return tp, fp, fn


def binned_integrate(tp: torch.Tensor, fp: torch.Tensor, fn: Optional[torch.Tensor], mode: int) -> torch.Tensor:
    lib = N.load()
    dev = _require_cuda(tp, fp, fn)
    tasks, T = tp.shape
    out = torch.empty(tasks, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        N.check(
            "te_binned_integrate",
            lib.te_binned_integrate(tp.data_ptr(), fp.data_ptr(), _ptr(fn), T, tasks, mode, out.data_ptr(), _stream_ptr(dev)),
        )
    return out


# ----------------------------------------------------------------------------- K1
def _binary_rank(fn_name: str, score: torch.Tensor, target: torch.Tensor, variant: Optional[int] = None) -> torch.Tensor:
    lib = N.load()
    dev = _require_cuda(score, target)
    assert score.dim() == 1 and score.shape == target.shape and score.dtype == torch.float32
    score, target = score.contiguous(), target.contiguous()
    code = _DTYPE_CODE[target.dtype]
    n = score.numel()
    scratch = _scratch(lib.te_binary_rank_scratch_bytes(n), dev)
    out = torch.empty(8, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        N.check(
            fn_name,
            getattr(lib, fn_name)(
                score.data_ptr(), target.data_ptr(), code, n, out.data_ptr(), scratch.data_ptr(),
                scratch.numel(), *(() if variant is None else (variant,)), _stream_ptr(dev),
            ),
        )
    return out


def binary_auroc_packed(score: torch.Tensor, target: torch.Tensor, variant: int = N.TE_RANK_AUTO) -> torch.Tensor:
    """float64[8] device tensor: [auroc, #pos, #neg, 2U, flags, ...] (see te_binary_auroc).
    `variant`: TE_RANK_AUTO / TE_RANK_SORT / TE_RANK_COUNTING (te_binary_auroc_variant)."""
    return _binary_rank("te_binary_auroc_variant", score, target, variant)


def binary_auprc_packed(score: torch.Tensor, target: torch.Tensor, variant: int = N.TE_RANK_AUTO) -> torch.Tensor:
    return _binary_rank("te_binary_auprc_variant", score, target, variant)


def _binary_rank_general(mode: int, score: torch.Tensor, target: torch.Tensor, weight: Optional[torch.Tensor]) -> torch.Tensor:
    lib = N.load()
    dev = _require_cuda(score, target, weight)
    assert score.dim() == 1 and score.shape == target.shape and score.dtype == torch.float32
    score, target = score.contiguous(), target.contiguous()
    if target.dtype not in _DTYPE_CODE:
        target = target.to(torch.float64)
    code = _DTYPE_CODE[target.dtype]
    n = score.numel()
    scratch = _scratch(lib.te_binary_rank_general_scratch_bytes(n), dev)
    out = torch.empty(8, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        if mode == 0:
            if weight is not None:
                assert weight.shape == score.shape
                weight = weight.to(torch.float64).contiguous()
            rc = lib.te_binary_auroc_general(
                score.data_ptr(), target.data_ptr(), code, _ptr(weight), n, out.data_ptr(),
                scratch.data_ptr(), scratch.numel(), _stream_ptr(dev),
            )
            N.check("te_binary_auroc_general", rc)
        else:
            rc = lib.te_binary_auprc_general(
                score.data_ptr(), target.data_ptr(), code, n, out.data_ptr(), scratch.data_ptr(),
                scratch.numel(), _stream_ptr(dev),
            )
            N.check("te_binary_auprc_general", rc)
    return out


def binary_auroc_general(score: torch.Tensor, target: torch.Tensor, weight: Optional[torch.Tensor] = None) -> torch.Tensor:
    """float64[8]: [auroc, sum w*t, sum w*(1-t), cross term, 0, ...] — any scores, soft labels, weights."""
    return _binary_rank_general(0, score, target, weight)


def binary_auroc_general_batched(score: torch.Tensor, target: torch.Tensor, weight: Optional[torch.Tensor] = None) -> torch.Tensor:
    """(tasks, n) scores / targets / weights -> float64 [tasks, 8], row y as binary_auroc_general of
    row y; all tasks share every launch."""
    lib = N.load()
    dev = _require_cuda(score, target, weight)
    assert score.dim() == 2 and score.shape == target.shape and score.dtype == torch.float32
    score, target = score.contiguous(), target.contiguous()
    if target.dtype not in _DTYPE_CODE:
        target = target.to(torch.float64)
    if weight is not None:
        assert weight.shape == score.shape
        weight = weight.to(torch.float64).contiguous()
    tasks, n = score.shape
    out = torch.empty((tasks, 8), dtype=torch.float64, device=dev)
    step = 65535
    with torch.cuda.device(dev):
        for t0 in range(0, tasks, step):
            k = min(step, tasks - t0)
            scratch = _scratch(lib.te_binary_auroc_general_batched_scratch_bytes(k, n), dev)
            N.check(
                "te_binary_auroc_general_batched",
                lib.te_binary_auroc_general_batched(
                    score[t0:].data_ptr(), target[t0:].data_ptr(), _DTYPE_CODE[target.dtype],
                    None if weight is None else weight[t0:].data_ptr(), k, n, n, out[t0:].data_ptr(),
                    scratch.data_ptr(), scratch.numel(), _stream_ptr(dev),
                ),
            )
    return out


def binary_auprc_general(score: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    return _binary_rank_general(1, score, target, None)


def sort_u32(keys: torch.Tensor, vals: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """Ascending stable sort of uint32 keys held in an int32/uint32 tensor (bit pattern)."""
    lib = N.load()
    dev = _require_cuda(keys, vals)
    assert keys.dim() == 1 and keys.element_size() == 4 and keys.is_contiguous()
    n = keys.numel()
    out_k = torch.empty_like(keys)
    out_v = torch.empty_like(vals) if vals is not None else None
    scratch = _scratch(lib.te_sort_u32_scratch_bytes(n, 1 if vals is not None else 0), dev)
    with torch.cuda.device(dev):
        N.check(
            "te_sort_u32",
            lib.te_sort_u32(
                keys.data_ptr(), out_k.data_ptr(), _ptr(vals), _ptr(out_v), n, scratch.data_ptr(),
                scratch.numel(), _stream_ptr(dev),
            ),
        )
    return out_k, out_v


def dense_rank(x: torch.Tensor, with_values: bool = False):
    """Dense ranks of float64 / int64 scores (any shape, ranked as one pool): (int32 tensor of x's
    shape, distinct values ascending or None).  One host read (the number of distinct values) only
    when the values are asked for."""
    lib = N.load()
    dev = _require_cuda(x)
    assert x.dtype in (torch.float64, torch.int64)
    flat = x.contiguous().reshape(-1)
    n = flat.numel()
    rank = torch.empty(n, dtype=torch.int32, device=dev)
    values = torch.empty(n, dtype=x.dtype, device=dev) if with_values else None
    count = torch.zeros(1, dtype=torch.int32, device=dev)
    scratch = _scratch(lib.te_dense_rank_scratch_bytes(n), dev)
    with torch.cuda.device(dev):
        N.check(
            "te_dense_rank",
            lib.te_dense_rank(
                flat.data_ptr(), N.TE_DTYPE_F64 if x.dtype == torch.float64 else N.TE_DTYPE_I64, n,
                rank.data_ptr(), _ptr(values), count.data_ptr(), scratch.data_ptr(), scratch.numel(),
                _stream_ptr(dev),
            ),
        )
    if with_values:
        values = values[: int(count.item())]
    return rank.reshape(x.shape), values


def ovr_rank(input: torch.Tensor, target: torch.Tensor, multilabel: bool, mode: int) -> torch.Tensor:
    """One-vs-rest AUROC (mode 0) / AUPRC (mode 1) of an (n, C) float32 matrix; float64 [C, 8],
    row c = [value, #pos, #neg, raw, flags, ...] as te_binary_auroc."""
    lib = N.load()
    dev = _require_cuda(input, target)
    assert input.dim() == 2 and input.dtype == torch.float32 and target.dtype == torch.int64
    if input.stride(1) != 1:
        input = input.contiguous()
    target = target.contiguous()
    n, C = input.shape
    assert target.shape == ((n, C) if multilabel else (n,))
    out = torch.empty((C, 8), dtype=torch.float64, device=dev)
    scratch = _scratch(lib.te_ovr_rank_scratch_bytes(n), dev)
    with torch.cuda.device(dev):
        N.check(
            "te_ovr_rank",
            lib.te_ovr_rank(input.data_ptr(), input.stride(0), target.data_ptr(), C if multilabel else 0, n, C,
                            1 if multilabel else 0, mode, out.data_ptr(), scratch.data_ptr(), scratch.numel(), _stream_ptr(dev)),
        )
    return out


# ----------------------------------------------------------------------------- distributed pieces
def _host_splitters(splitters):
    import ctypes

    nsp = len(splitters)
    return (ctypes.c_uint32 * max(nsp, 1))(*[int(v) & 0xFFFFFFFF for v in splitters]), nsp


def rank_pack(score: torch.Tensor, target: torch.Tensor, splitters=(), want_hist: bool = False):
    """(packed keys as int32 tensor [n], info int64[2] = {#positives, flags}, bucket sizes
    int64[len(splitters) + 1]).  `splitters`: ascending key values, either a host sequence of ints
    or an int32 DEVICE tensor holding their bit patterns (no host round trip to choose them).
    want_hist: a fourth result, int32 [buckets, 4 * 256] — the radix-digit histograms of every
    bucket's keys (what the receiver of a bucket needs to start sorting it)."""
    import ctypes

    lib = N.load()
    dev = _require_cuda(score, target)
    assert score.dim() == 1 and score.shape == target.shape and score.dtype == torch.float32
    score, target = score.contiguous(), target.contiguous()
    n = score.numel()
    on_device = isinstance(splitters, torch.Tensor)
    if on_device:
        assert splitters.dtype == torch.int32 and splitters.device == dev and splitters.dim() == 1
        splitters = splitters.contiguous()
        nsp = splitters.numel()
    else:
        sp, nsp = _host_splitters(splitters)
    keys = torch.empty(n, dtype=torch.int32, device=dev)
    info = torch.zeros(2, dtype=torch.int64, device=dev)
    counts = torch.zeros(nsp + 1, dtype=torch.int64, device=dev)
    hist = torch.empty((nsp + 1, 1024), dtype=torch.int32, device=dev) if want_hist else None
    scratch = _scratch(lib.te_rank_pack_scratch_bytes(n), dev)
    with torch.cuda.device(dev):
        if on_device:
            N.check(
                "te_rank_pack_dsplit",
                lib.te_rank_pack_dsplit(score.data_ptr(), target.data_ptr(), _DTYPE_CODE[target.dtype], n,
                                        splitters.data_ptr() if nsp else None, nsp, keys.data_ptr(), info.data_ptr(),
                                        counts.data_ptr(), _ptr(hist), scratch.data_ptr(), scratch.numel(),
                                        _stream_ptr(dev)),
            )
        else:
            N.check(
                "te_rank_pack",
                lib.te_rank_pack(score.data_ptr(), target.data_ptr(), _DTYPE_CODE[target.dtype], n,
                                 ctypes.cast(sp, ctypes.c_void_p), nsp, keys.data_ptr(), info.data_ptr(),
                                 counts.data_ptr(), _ptr(hist), scratch.data_ptr(), scratch.numel(), _stream_ptr(dev)),
            )
    return (keys, info, counts, hist) if want_hist else (keys, info, counts)


def partition_u32(keys: torch.Tensor, splitters, counts_in: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Stable partition of uint32 keys (int32 tensor) by ascending host-side splitters; returns
    (keys grouped by bucket, int64 bucket counts [len(splitters) + 1])."""
    import ctypes

    lib = N.load()
    dev = _require_cuda(keys)
    keys = keys.contiguous()
    n = keys.numel()
    sp, nsp = _host_splitters(splitters)
    out = torch.empty_like(keys)
    counts = torch.zeros(nsp + 1, dtype=torch.int64, device=dev)
    scratch = _scratch(lib.te_partition_u32_scratch_bytes(n), dev)
    with torch.cuda.device(dev):
        N.check(
            "te_partition_u32",
            lib.te_partition_u32(keys.data_ptr(), n, ctypes.cast(sp, ctypes.c_void_p), nsp, _ptr(counts_in),
                                 out.data_ptr(), counts.data_ptr(), scratch.data_ptr(), scratch.numel(), _stream_ptr(dev)),
        )
    return out, counts


def partition_scatter_u32(keys: torch.Tensor, splitters, counts_in: torch.Tensor, dest_ptrs) -> None:
    """Partition `keys` by the splitters and store bucket b at device address dest_ptrs[b]
    (+ the pass's own output index), e.g. into peer-mapped inboxes of other GPUs."""
    import ctypes

    lib = N.load()
    dev = _require_cuda(keys, counts_in)
    keys = keys.contiguous()
    n = keys.numel()
    sp, nsp = _host_splitters(splitters)
    assert len(dest_ptrs) == nsp + 1
    ptrs = (ctypes.c_uint64 * (nsp + 1))(*[int(p) & 0xFFFFFFFFFFFFFFFF for p in dest_ptrs])
    counts = torch.zeros(nsp + 1, dtype=torch.int64, device=dev)
    scratch = _scratch(lib.te_partition_u32_scratch_bytes(n), dev)
    with torch.cuda.device(dev):
        N.check(
            "te_partition_scatter_u32",
            lib.te_partition_scatter_u32(keys.data_ptr(), n, ctypes.cast(sp, ctypes.c_void_p), nsp, _ptr(counts_in),
                                         ctypes.cast(ptrs, ctypes.c_void_p), counts.data_ptr(), scratch.data_ptr(),
                                         scratch.numel(), _stream_ptr(dev)),
        )


def _hist_ptr(hist: Optional[torch.Tensor], dev) -> Optional[int]:
    if hist is None:
        return None
    assert hist.dtype == torch.int32 and hist.numel() == 1024 and hist.device == dev and hist.is_contiguous()
    return hist.data_ptr()


def binary_auroc_from_keys(keys: torch.Tensor, hist: Optional[torch.Tensor] = None, variant: int = N.TE_RANK_AUTO) -> torch.Tensor:
    """float64[8] as te_binary_auroc_from_keys: [auroc, P, N, 2U, flags, 2U_hi32, 2U_lo32, n].
    hist: int32[4 * 256] radix-digit histograms of `keys` if they are already known (used by
    TE_RANK_SORT).  variant: TE_RANK_* (te_binary_auroc_from_keys_variant)."""
    lib = N.load()
    dev = _require_cuda(keys)
    keys = keys.contiguous()
    n = keys.numel()
    out = torch.empty(8, dtype=torch.float64, device=dev)
    scratch = _scratch(lib.te_binary_rank_scratch_bytes(n), dev)
    with torch.cuda.device(dev):
        N.check(
            "te_binary_auroc_from_keys_variant",
            lib.te_binary_auroc_from_keys_variant(keys.data_ptr(), n, _hist_ptr(hist, dev), out.data_ptr(), scratch.data_ptr(),
                                                  scratch.numel(), variant, _stream_ptr(dev)),
        )
    return out


def binary_auprc_from_keys_begin(keys: torch.Tensor, hist: Optional[torch.Tensor] = None):
    """Phase 1 of the rank-sharded average precision: sort + count.  Returns (out, state) where
    out is float64[8] like ``binary_auroc_from_keys`` and ``state`` must be handed, untouched, to
    ``binary_auprc_from_keys_finish``."""
    lib = N.load()
    dev = _require_cuda(keys)
    keys = keys.contiguous()
    n = keys.numel()
    out = torch.empty(8, dtype=torch.float64, device=dev)
    scratch = _scratch(lib.te_binary_rank_scratch_bytes(n), dev)
    with torch.cuda.device(dev):
        N.check(
            "te_binary_auprc_from_keys_begin",
            lib.te_binary_auprc_from_keys_begin(
                keys.data_ptr(), n, _hist_ptr(hist, dev), out.data_ptr(), scratch.data_ptr(), scratch.numel(),
                _stream_ptr(dev)
            ),
        )
    return out, (n, scratch)


def binary_auprc_from_keys_finish(state, positives_above, samples_above=None) -> torch.Tensor:
    """Phase 2: float64[8]; [3] = sum over the local positives of the precision at their
    threshold, given what the ranks holding higher scores contain: two Python ints, or ONE
    float64[2] device tensor {positives_above, samples_above} (no host read needed)."""
    lib = N.load()
    n, scratch = state
    dev = scratch.device
    out = torch.zeros(8, dtype=torch.float64, device=dev)
    if isinstance(positives_above, torch.Tensor):
        above = positives_above.contiguous()
        assert above.dtype == torch.float64 and above.numel() == 2 and above.device == dev
        args = (0.0, 0.0, above.data_ptr())
    else:
        args = (float(positives_above), float(samples_above), None)
    with torch.cuda.device(dev):
        N.check(
            "te_binary_auprc_from_keys_finish",
            lib.te_binary_auprc_from_keys_finish(
                n, args[0], args[1], args[2], out.data_ptr(), scratch.data_ptr(), scratch.numel(), _stream_ptr(dev)
            ),
        )
    return out


def binary_pr_curve(score: torch.Tensor, target: torch.Tensor):
    """Exact precision-recall curve of (score, target == 1): ((precision[K+1], recall[K+1],
    thresholds[K]) float32 on the device, thresholds ascending), flags.  flags != 0: the packed
    representation does not apply (negative scores / labels outside {0, 1}) and the curve is None.
    One host read (K sizes the outputs)."""
    lib = N.load()
    dev = _require_cuda(score, target)
    assert score.dim() == 1 and score.shape == target.shape and score.dtype == torch.float32
    score, target = score.contiguous(), target.contiguous()
    n = score.numel()
    out = torch.zeros(8, dtype=torch.float64, device=dev)
    scratch = _scratch(lib.te_binary_rank_scratch_bytes(n), dev)
    with torch.cuda.device(dev):
        N.check(
            "te_binary_pr_curve_begin",
            lib.te_binary_pr_curve_begin(score.data_ptr(), target.data_ptr(), _DTYPE_CODE[target.dtype], n,
                                         out.data_ptr(), scratch.data_ptr(), scratch.numel(), _stream_ptr(dev)),
        )
        info = out.tolist()
        flags = int(info[4])
        if flags:
            return None, flags
        k = int(info[0])
        precision = torch.empty(k + 1, dtype=torch.float32, device=dev)
        recall = torch.empty(k + 1, dtype=torch.float32, device=dev)
        thresholds = torch.empty(k, dtype=torch.float32, device=dev)
        N.check(
            "te_binary_pr_curve_finish",
            lib.te_binary_pr_curve_finish(n, precision.data_ptr(), recall.data_ptr(), thresholds.data_ptr(),
                                          scratch.data_ptr(), scratch.numel(), _stream_ptr(dev)),
        )
    return (precision, recall, thresholds), 0


def rank_pack_scatter(score: torch.Tensor, target: torch.Tensor, splitters_dev: Optional[torch.Tensor], world: int,
                      rank: int, inbox_ptrs, cap: int, work: torch.Tensor) -> None:
    """te_rank_pack_scatter: pack (score, target) and store every key straight into the slab this
    rank owns inside the destination rank's inbox (inbox_ptrs[b] = rank b's inbox as mapped here).
    `work`: int32[32] device tensor (cleared by the call).  Nothing is returned: fill counts and
    flags travel to the receivers' inbox headers."""
    import ctypes

    lib = N.load()
    dev = _require_cuda(score, target)
    assert score.dim() == 1 and score.shape == target.shape and score.dtype == torch.float32
    score, target = score.contiguous(), target.contiguous()
    assert work.dtype == torch.int32 and work.numel() >= 32 and work.device == dev
    if world > 1:
        assert splitters_dev is not None and splitters_dev.dtype == torch.int32 and splitters_dev.numel() == world - 1
    ptrs = (ctypes.c_uint64 * world)(*[int(p) for p in inbox_ptrs])
    with torch.cuda.device(dev):
        N.check(
            "te_rank_pack_scatter",
            lib.te_rank_pack_scatter(score.data_ptr(), target.data_ptr(), _DTYPE_CODE[target.dtype], score.numel(),
                                     _ptr(splitters_dev) if world > 1 else None, world, rank,
                                     ctypes.cast(ptrs, ctypes.c_void_p), cap, work.data_ptr(), _stream_ptr(dev)),
        )


def binary_auroc_from_slabs(inbox: torch.Tensor, world: int, cap: int) -> torch.Tensor:
    """te_binary_auroc_from_slabs over this rank's inbox (int32 tensor: header + world slabs of
    `cap` keys).  float64[8] as binary_auroc_from_keys; out[4] = OR of all senders' flags."""
    lib = N.load()
    dev = _require_cuda(inbox)
    out = torch.empty(8, dtype=torch.float64, device=dev)
    scratch = _scratch(lib.te_binary_auroc_from_slabs_scratch_bytes(world, cap), dev)
    with torch.cuda.device(dev):
        N.check(
            "te_binary_auroc_from_slabs",
            lib.te_binary_auroc_from_slabs(inbox.data_ptr(), world, cap, out.data_ptr(), scratch.data_ptr(),
                                           scratch.numel(), _stream_ptr(dev)),
        )
    return out


def select_splitters(samples: torch.Tensor, world: int):
    """samples: int32 [world * s] float bit patterns of sampled scores (negative = none), at most
    32768 of them.  Returns (int32[world - 1] splitter bit patterns in key space, int64 copy)."""
    lib = N.load()
    dev = _require_cuda(samples)
    samples = samples.contiguous().view(-1)
    assert samples.dtype == torch.int32 and samples.numel() <= 32768
    sp32 = torch.empty(max(world - 1, 0), dtype=torch.int32, device=dev)
    sp64 = torch.empty(max(world - 1, 0), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        N.check(
            "te_select_splitters",
            lib.te_select_splitters(samples.data_ptr(), samples.numel(), world, _ptr(sp32) if world > 1 else None,
                                    _ptr(sp64) if world > 1 else None, _stream_ptr(dev)),
        )
    return sp32, sp64


def combine_auroc_stats(rows: torch.Tensor, n_received: int) -> torch.Tensor:
    """rows: float64 [world, 8] gathered te_binary_auroc_from_keys results (rank order = ascending
    score ranges) -> float64[8] [auroc, P, N, 2U, 0, n_received, 0, 0]."""
    lib = N.load()
    dev = _require_cuda(rows)
    rows = rows.contiguous()
    assert rows.dtype == torch.float64 and rows.dim() == 2 and rows.shape[1] == 8
    out = torch.empty(8, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        N.check(
            "te_combine_auroc_stats",
            lib.te_combine_auroc_stats(rows.data_ptr(), rows.shape[0], float(n_received), out.data_ptr(), _stream_ptr(dev)),
        )
    return out
