"""Confusion-matrix update at C = 1000, N = 2^28: time per update for a given number of route/count
batches (TE_B200_CM_BATCHES is read once per process)."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from bench import KernelProfile
from torcheval_b200 import ops

n, C = 1 << int(os.environ.get("LOG2N", "28")), 1000
g = torch.Generator(device="cuda").manual_seed(3)
pred = torch.randint(0, C, (n,), device="cuda", generator=g)
target = torch.randint(0, C, (n,), device="cuda", generator=g)
cm = torch.zeros((C, C), dtype=torch.int64, device="cuda")
for _ in range(3):
    ops.cm_update_(cm, pred, target, C)
torch.cuda.synchronize()
cm.zero_()
steps = 10
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with KernelProfile() as prof:
    e0.record()
    for _ in range(steps):
        ops.cm_update_(cm, pred, target, C)
    e1.record()
    torch.cuda.synchronize()
    k = prof.read()
ms = e0.elapsed_time(e1) / steps
want = torch.bincount(target * C + pred, minlength=C * C).reshape(C, C) * steps
assert torch.equal(cm, want)
print(f"batches={os.environ.get('TE_B200_CM_BATCHES', 'default')} n=2^{n.bit_length() - 1}: {ms:.4f} ms/update  "
      f"{16 * n / ms / 1e6:.1f} GB/s  frac={16 * n / ms / 1e6 / 6450.9:.3f}  "
      + "  ".join(f"{kn}={v['total_ms'] / steps:.3f}ms/{v['launches'] // steps}" for kn, v in k.items()))
