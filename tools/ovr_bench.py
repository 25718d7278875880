"""One-vs-rest multiclass AUROC / multilabel AUPRC throughput (K2) on one GPU."""
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from torcheval_b200 import ops  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(5)
    for log2n, C in [(20, 1000), (22, 1000), (24, 64), (18, 1000)]:
        n = 1 << log2n
        x = torch.rand((n, C), device=dev, generator=g)
        y = torch.randint(0, C, (n,), device=dev, generator=g)
        for mode, name in [(0, "multiclass_auroc"), (1, "multiclass_auprc")]:
            ops.ovr_rank(x, y, False, mode)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            reps = 2
            for _ in range(reps):
                out = ops.ovr_rank(x, y, False, mode)
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) / reps
            print(f"N=2^{log2n} C={C:5d} {name:18s} {dt * 1e3:9.2f} ms  {n * C / dt / 1e9:7.2f} G scores/s  mean={float(out[:, 0].mean()):.5f}", flush=True)
        del x, y


if __name__ == "__main__":
    main()
