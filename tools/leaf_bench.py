"""Times te_binary_auroc_variant (TE_RANK_SORT vs TE_RANK_COUNTING) and prints the per-kernel
breakdown from the library's own CUDA-event profile.  Usage: python tools/leaf_bench.py [log2n ...]"""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from bench import KernelProfile
from torcheval_b200 import ops


def run(log2n, kind="uniform", reps=5):
    n = 1 << log2n
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.rand(n, device="cuda", generator=g)
    if kind == "band":
        x = 0.875 + x / 8
    elif kind == "grid":
        x = torch.floor(x * 4096) / 4096
    elif kind == "const":
        x = torch.full_like(x, 0.375)
    elif kind == "two":
        x = torch.where(x < 0.3, torch.full_like(x, 0.25), torch.full_like(x, 0.75))
    y = torch.randint(0, 2, (n,), device="cuda", generator=g)
    res = {}
    for variant, name in ((1, "sort"), (2, "counting")):
        for _ in range(2):
            out = ops.binary_auroc_packed(x, y, variant)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with KernelProfile() as prof:
            e0.record()
            for _ in range(reps):
                out = ops.binary_auroc_packed(x, y, variant)
            e1.record()
            torch.cuda.synchronize()
            k = prof.read()
        ms = e0.elapsed_time(e1) / reps
        res[name] = out.cpu().numpy()
        print(f"n=2^{log2n} {kind:8s} {name:9s} {ms:8.3f} ms/step  {n / ms / 1e6:7.1f} G scores/s   auroc={res[name][0]:.12f} flags={res[name][4]}")
        for kn, v in sorted(k.items(), key=lambda kv: -kv[1]["total_ms"]):
            print(f"      {kn:28s} {v['total_ms'] / reps:8.3f} ms/step  ({v['launches'] // reps} launches)")
    assert (res["sort"] == res["counting"]).all(), (res["sort"], res["counting"])
    if os.environ.get("AUPRC") == "1":
        for variant, name in ((1, "sort"), (2, "counting")):
            for _ in range(2):
                out = ops.binary_auprc_packed(x, y, variant)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                out = ops.binary_auprc_packed(x, y, variant)
            e1.record()
            torch.cuda.synchronize()
            print(f"n=2^{log2n} {kind:8s} AUPRC {name:9s} {e0.elapsed_time(e1) / reps:8.3f} ms/step  ap={out[0].item():.12f}")


if __name__ == "__main__":
    sizes = [int(a) for a in sys.argv[1:] if a.isdigit()] or [24, 27, 30]
    for s in sizes:
        for kind in os.environ.get("KINDS", "uniform,band,grid").split(","):
            run(s, kind)
