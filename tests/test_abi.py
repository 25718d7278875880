"""CPU-side checks of the drop-in boundary: libte_b200.so builds, loads without a GPU and exports
exactly what include/te_b200.h declares; the ctypes table mirrors the header."""
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def lib():
    from torcheval_b200 import _build, _native

    _build.build()
    return _native.load()


def declared():
    text = (ROOT / "include" / "te_b200.h").read_text()
    return set(re.findall(r"TE_API\s+[\w\s\*]+?\b(te_\w+)\s*\(", text))


def test_header_symbols_are_exported(lib):
    names = declared()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), f"{n} declared in te_b200.h but not exported"


def test_ctypes_table_matches_header():
    from torcheval_b200 import _native

    assert set(_native.SIGNATURES) == declared()


def test_no_torch_types_in_abi():
    text = (ROOT / "include" / "te_b200.h").read_text()
    code = re.sub(r"/\*.*?\*/", "", text, flags=re.S)  # declarations only, comments cite torch ops
    assert "torch" not in code.lower() and "at::" not in code and "Tensor" not in code


def test_version_and_status_strings(lib):
    assert lib.te_abi_version() == 1
    assert lib.te_status_string(0) == b"ok"
    assert b"scratch" in lib.te_status_string(-2)


def test_scratch_queries_need_no_gpu(lib):
    assert lib.te_cm_scratch_bytes(1000, 2) >= 4_000_000
    assert lib.te_cm_scratch_bytes(1000, 1) == 0
    assert lib.te_binned_scratch_bytes(10000) > 0
    n = 1 << 20
    assert lib.te_binary_rank_scratch_bytes(n) >= 2 * 4 * n
    assert lib.te_sort_u32_scratch_bytes(n, 1) > lib.te_sort_u32_scratch_bytes(n, 0)
    assert lib.te_binary_rank_scratch_bytes(1 << 31) == 0  # beyond the 2^31-1 limit


def test_product_never_imports_the_oracle():
    for p in (ROOT / "torcheval_b200").rglob("*.py"):
        src = p.read_text()
        assert not re.search(r"^\s*(import|from)\s+oracle\b", src, flags=re.M), p
