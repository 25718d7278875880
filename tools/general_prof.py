import sys, torch
sys.path.insert(0, "/root/repo")
sys.path.insert(0, "/root/repo/tools")
from torcheval_b200 import ops
import tune
n = 1 << 27
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn(n, device="cuda", generator=g)
y = torch.randint(0, 2, (n,), device="cuda", generator=g)
w = torch.rand(n, device="cuda", generator=g, dtype=torch.float64)
for name, fn in [("general auroc (logits)", lambda: ops.binary_auroc_general(x, y, None)),
                 ("general auroc (weights)", lambda: ops.binary_auroc_general(x, y, w)),
                 ("general auprc", lambda: ops.binary_auprc_general(x, y))]:
    fn(); torch.cuda.synchronize()
    print(name, f"{tune.timeit(fn):.3f} ms")
    for k, (ms, c) in tune.kernel_times(fn).items():
        print(f"    {k:28s} {ms / max(c, 1):8.3f} ms x{c}")
