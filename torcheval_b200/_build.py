"""Build recipe for libte_b200.so (the sm_100a kernels + the C ABI of include/te_b200.h).

Plain nvcc, in-tree output (torcheval_b200/lib/), so that the built library travels with the
source snapshot.  Run as ``python -m torcheval_b200._build`` or through ``__graft_entry__.build()``.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
LIBDIR = PKG / "lib"
LIB = LIBDIR / "libte_b200.so"
INCLUDE = PKG.parent / "include"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC,-O3,-fvisibility=hidden",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
    "-cudart", "static",
] + os.environ.get("TE_EXTRA_NVCC", "").split()


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.sep not in cand or os.path.exists(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _digest() -> str:
    h = hashlib.sha256()
    for p in sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(INCLUDE.glob("*.h"))):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def is_stale() -> bool:
    stamp = LIBDIR / "libte_b200.digest"
    return not (LIB.exists() and stamp.exists() and stamp.read_text().strip() == _digest())


def build_stats_variant() -> Path:
    """libte_b200_stats.so: same sources with -DTE_PHASE_STATS (per-phase cycle counters in the
    radix pass; see te_debug_phase_stats).  Select it with TE_B200_LIB=<path>."""
    LIBDIR.mkdir(exist_ok=True)
    out = LIBDIR / "libte_b200_stats.so"
    cmd = [_nvcc(), *[f for f in NVCC_FLAGS if f not in ("-Xptxas", "-v")], "-DTE_PHASE_STATS", "-I", str(INCLUDE),
           "-shared", "-o", str(out), *map(str, sources())]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed:\n{r.stdout}\n{r.stderr}")
    return out


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every .cu under csrc/ for sm_100a and link libte_b200.so."""
    if not force and not is_stale():
        return LIB
    LIBDIR.mkdir(exist_ok=True)
    objdir = LIBDIR / "obj"
    objdir.mkdir(exist_ok=True)
    nvcc = _nvcc()
    logs = {}

    def compile_one(src: Path) -> Path:
        obj = objdir / (src.stem + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-I", str(INCLUDE), "-c", str(src), "-o", str(obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        logs[src.name] = r.stderr
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, sources()))
    cmd = [nvcc, "-shared", "-cudart", "static", "-o", str(LIB), *map(str, objs)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    (LIBDIR / "ptxas.log").write_text("\n".join(f"== {k}\n{v}" for k, v in sorted(logs.items())))
    (LIBDIR / "libte_b200.digest").write_text(_digest())
    if verbose:
        print((LIBDIR / "ptxas.log").read_text())
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    if "--stats" in sys.argv:
        print(build_stats_variant())
