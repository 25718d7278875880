"""One-vs-rest AUROC (te_ovr_rank) at C classes x 2^k samples, counting-stage path against the
four-pass sort (TE_B200_OVR_LEAF_MIN_N decides), per-kernel times, and equality of the two."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent))
from bench import KernelProfile  # noqa: E402
from torcheval_b200 import ops  # noqa: E402

C = int(sys.argv[1]) if len(sys.argv) > 1 else 64
ks = [int(a) for a in sys.argv[2:]] or [20, 22, 24]
for k in ks:
    n = 1 << k
    g = torch.Generator(device="cuda").manual_seed(k)
    x = torch.rand((n, C), device="cuda", generator=g)
    y = torch.randint(0, C, (n,), device="cuda", generator=g)
    out = ops.ovr_rank(x, y, False, 0)
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with KernelProfile() as prof:
        ev0.record()
        for _ in range(3):
            out = ops.ovr_rank(x, y, False, 0)
        ev1.record()
        torch.cuda.synchronize()
        kern = prof.read()
    ms = ev0.elapsed_time(ev1) / 3
    o = out.cpu().numpy()
    print(f"C={C} n=2^{k} LEAF_MIN_N={os.environ.get('TE_B200_OVR_LEAF_MIN_N', 'default')}: {ms:.3f} ms  {n * C / ms / 1e6:.1f} G scores/s  "
          f"mean={o[:, 0].mean():.9f} sum2U={o[:, 3].sum():.0f} flags={o[:, 4].max()}")
    print("   ", {kk: round(v["total_ms"] / 3, 3) for kk, v in kern.items()})
