"""Two calls of te_binary_auroc_variant(TE_RANK_COUNTING) on 2^k uniform scores, for ncu:
the first call is the warm-up, profile the kernels of the second."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from torcheval_b200 import ops

log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 27
variant = int(sys.argv[2]) if len(sys.argv) > 2 else 2
n = 1 << log2n
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.rand(n, device="cuda", generator=g)
y = torch.randint(0, 2, (n,), device="cuda", generator=g)
for _ in range(2):
    out = ops.binary_auroc_packed(x, y, variant)
torch.cuda.synchronize()
print(out.cpu().numpy())
