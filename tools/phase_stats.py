"""Per-phase cycle breakdown of the radix pass (needs the -DTE_PHASE_STATS build):
    python -m torcheval_b200._build --stats
    TE_B200_LIB=torcheval_b200/lib/libte_b200_stats.so python tools/phase_stats.py [log2 n]
"""
import ctypes
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from torcheval_b200 import _native, ops  # noqa: E402

NAMES = ["setup", "load", "rank", "wait-others", "digit-phase", "scatter", "look-back", "output"]


def main():
    logn = int(sys.argv[1]) if len(sys.argv) > 1 else 27
    n = 1 << logn
    lib = _native.load()
    x = torch.rand(n, device="cuda")
    y = torch.randint(0, 2, (n,), device="cuda")
    ops.binary_auroc_packed(x, y)
    torch.cuda.synchronize()
    buf = (ctypes.c_uint64 * 8)()
    lib.te_debug_phase_stats(ctypes.cast(buf, ctypes.c_void_p), 1)
    reps = 3
    for _ in range(reps):
        ops.binary_auroc_packed(x, y)
    torch.cuda.synchronize()
    lib.te_debug_phase_stats(ctypes.cast(buf, ctypes.c_void_p), 0)
    tiles = reps * 4 * ((n + 8191) // 8192)
    tot = sum(buf)
    print(f"n=2^{logn}: {tiles} tiles, mean cycles per tile (thread 0 of each CTA):")
    for name, c in zip(NAMES, buf):
        print(f"  {name:12s} {c / tiles:9.0f}  {c / max(tot, 1):6.1%}")
    print(f"  {'total':12s} {tot / tiles:9.0f}")


if __name__ == "__main__":
    main()
