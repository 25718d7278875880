"""ctypes binding of libte_b200.so — the C ABI declared in include/te_b200.h.

There is no fallback: if the library has not been built (``python -m torcheval_b200._build`` or
``__graft_entry__.build()``) loading raises, and every metric that needs a kernel raises with it.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_LIB_PATH = Path(__file__).resolve().parent / "lib" / "libte_b200.so"

TE_OK = 0
TE_CM_AUTO, TE_CM_GLOBAL64, TE_CM_GLOBAL32, TE_CM_SMEM_REPL, TE_CM_SMEM, TE_CM_CLUSTER, TE_CM_BUCKETED = range(7)
TE_BINNED_AUROC, TE_BINNED_AUPRC = 0, 1
TE_DTYPE_I64, TE_DTYPE_I32, TE_DTYPE_U8, TE_DTYPE_F32, TE_DTYPE_F64 = range(5)
TE_RANK_FLAG_BAD_LABEL, TE_RANK_FLAG_NEGATIVE_SCORE, TE_RANK_FLAG_WIDE_SCORE = 1, 2, 4
TE_RANK_AUTO, TE_RANK_SORT, TE_RANK_COUNTING = 0, 1, 2

_vp, _i32, _i64, _sz = C.c_void_p, C.c_int32, C.c_int64, C.c_size_t

# name -> (restype, argtypes); mirrors include/te_b200.h one to one.
SIGNATURES = {
    "te_abi_version": (C.c_int, []),
    "te_status_string": (C.c_char_p, [C.c_int]),
    "te_profile_enable": (C.c_int, [C.c_int]),
    "te_profile_read": (C.c_int, [C.c_char_p, _sz, _vp, _vp, _i32]),
    "te_debug_phase_stats": (C.c_int, [_vp, C.c_int]),
    "te_cm_scratch_bytes": (_sz, [_i32, _i32]),
    "te_cm_scratch_bytes_n": (_sz, [_i32, _i32, _i64]),
    "te_cm_update": (C.c_int, [_vp, _vp, _i64, _i32, _vp, _vp, _vp, _sz, _i32, _vp]),
    "te_cm_update_noalloc": (C.c_int, [_vp, _vp, _i64, _i32, _vp, _vp, _vp]),
    "te_cm_update_logits": (C.c_int, [_vp, _i64, _vp, _i64, _i32, _vp, _vp, _vp]),
    "te_class_counts": (C.c_int, [_vp, _vp, _i64, _vp, _i64, _i32, _vp, _vp, _vp]),
    "te_match_count": (C.c_int, [_vp, _vp, _i64, _vp, _vp]),
    "te_binned_scratch_bytes": (_sz, [_i32]),
    "te_binned_hist_update": (C.c_int, [_vp, _i64, _vp, _i64, _i64, _i32, _vp, _i32, _vp, _vp, _sz, _vp]),
    "te_binned_hist2d_update": (C.c_int, [_vp, _vp, _i64, _i32, _i32, _vp, _i32, _vp, _vp, _sz, _vp]),
    "te_binned_counts": (C.c_int, [_vp, _i32, _i32, _vp, _vp, _vp, _vp]),
    "te_binned_integrate": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp]),
    "te_binary_rank_scratch_bytes": (_sz, [_i64]),
    "te_binary_auroc": (C.c_int, [_vp, _vp, _i32, _i64, _vp, _vp, _sz, _vp]),
    "te_binary_auroc_variant": (C.c_int, [_vp, _vp, _i32, _i64, _vp, _vp, _sz, _i32, _vp]),
    "te_binary_auprc": (C.c_int, [_vp, _vp, _i32, _i64, _vp, _vp, _sz, _vp]),
    "te_binary_auprc_variant": (C.c_int, [_vp, _vp, _i32, _i64, _vp, _vp, _sz, _i32, _vp]),
    "te_binary_rank_general_scratch_bytes": (_sz, [_i64]),
    "te_binary_auroc_general": (C.c_int, [_vp, _vp, _i32, _vp, _i64, _vp, _vp, _sz, _vp]),
    "te_binary_auroc_general_batched_scratch_bytes": (_sz, [_i64, _i64]),
    "te_binary_auroc_general_batched": (C.c_int, [_vp, _vp, _i32, _vp, _i64, _i64, _i64, _vp, _vp, _sz, _vp]),
    "te_binary_auprc_general": (C.c_int, [_vp, _vp, _i32, _i64, _vp, _vp, _sz, _vp]),
    "te_ovr_rank_scratch_bytes": (_sz, [_i64]),
    "te_ovr_rank": (C.c_int, [_vp, _i64, _vp, _i64, _i64, _i32, _i32, _i32, _vp, _vp, _sz, _vp]),
    "te_rank_pack_scratch_bytes": (_sz, [_i64]),
    "te_rank_pack": (C.c_int, [_vp, _vp, _i32, _i64, _vp, _i32, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "te_rank_pack_dsplit": (C.c_int, [_vp, _vp, _i32, _i64, _vp, _i32, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "te_partition_u32_scratch_bytes": (_sz, [_i64]),
    "te_partition_u32": (C.c_int, [_vp, _i64, _vp, _i32, _vp, _vp, _vp, _vp, _sz, _vp]),
    "te_partition_scatter_u32": (C.c_int, [_vp, _i64, _vp, _i32, _vp, _vp, _vp, _vp, _sz, _vp]),
    "te_binary_auroc_from_keys": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _sz, _vp]),
    "te_binary_auroc_from_keys_variant": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _sz, _i32, _vp]),
    "te_binary_auprc_from_keys_begin": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _sz, _vp]),
    "te_binary_auprc_from_keys_finish": (C.c_int, [_i64, C.c_double, C.c_double, _vp, _vp, _vp, _sz, _vp]),
    "te_slab_inbox_bytes": (_sz, [_i32, _i64]),
    "te_slab_capacity": (_i64, [_i64, _i32]),
    "te_rank_pack_scatter": (C.c_int, [_vp, _vp, _i32, _i64, _vp, _i32, _i32, _vp, _i64, _vp, _vp]),
    "te_binary_auroc_from_slabs_scratch_bytes": (_sz, [_i32, _i64]),
    "te_binary_auroc_from_slabs": (C.c_int, [_vp, _i32, _i64, _vp, _vp, _sz, _vp]),
    "te_select_splitters": (C.c_int, [_vp, _i64, _i32, _vp, _vp, _vp]),
    "te_combine_auroc_stats": (C.c_int, [_vp, _i32, C.c_double, _vp, _vp]),
    "te_binary_pr_curve_begin": (C.c_int, [_vp, _vp, _i32, _i64, _vp, _vp, _sz, _vp]),
    "te_binary_pr_curve_finish": (C.c_int, [_i64, _vp, _vp, _vp, _vp, _sz, _vp]),
    "te_sort_u32_scratch_bytes": (_sz, [_i64, _i32]),
    "te_sort_u32": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _vp, _sz, _vp]),
    "te_debug_p2p_store_bench": (C.c_int, [_vp, _i32, _i64, _i32, _i32, _vp]),
    "te_dense_rank_scratch_bytes": (_sz, [_i64]),
    "te_dense_rank": (C.c_int, [_vp, _i32, _i64, _vp, _vp, _vp, _vp, _sz, _vp]),
}


class NativeLibraryMissing(RuntimeError):
    pass


_lib = None


def lib_path() -> Path:
    return Path(os.environ.get("TE_B200_LIB", _LIB_PATH))


def load() -> C.CDLL:
    """Load libte_b200.so once and attach the prototypes. Raises if it is not there."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not path.exists():
        raise NativeLibraryMissing(
            f"{path} not found: torcheval_b200 has no CPU or PyTorch fallback. "
            "Build the sm_100a kernels first: python -m torcheval_b200._build"
        )
    lib = C.CDLL(str(path))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the header and the library drifted apart
        fn.restype = res
        fn.argtypes = args
    if lib.te_abi_version() != 1:
        raise NativeLibraryMissing(f"{path}: ABI version {lib.te_abi_version()} != 1")
    _lib = lib
    return lib


class NativeError(RuntimeError):
    def __init__(self, fn: str, status: int):
        msg = load().te_status_string(status).decode()
        super().__init__(f"{fn} failed with status {status}: {msg}")
        self.status = status


def check(fn: str, status: int) -> None:
    if status != TE_OK:
        raise NativeError(fn, status)
