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
w32 = torch.rand(n, device="cuda", generator=g, dtype=torch.float32).double()  # float64 tensor, float32-exact values
for name, fn in [("general auroc (logits)", lambda: ops.binary_auroc_general(x, y, None)),
                 ("general auroc (weights, arbitrary float64: gathered by index)", lambda: ops.binary_auroc_general(x, y, w)),
                 ("general auroc (weights, float32-exact: ride as payload)", lambda: ops.binary_auroc_general(x, y, w32)),
                 ("general auprc", lambda: ops.binary_auprc_general(x, y))]:
    fn(); torch.cuda.synchronize()
    print(name, f"{tune.timeit(fn):.3f} ms")
    for k, (ms, c) in tune.kernel_times(fn).items():
        print(f"    {k:28s} {ms / max(c, 1):8.3f} ms x{c}")

# WindowedBinaryAUROC.compute: 64 tasks x window 2^16, weighted
from torcheval_b200.metrics import WindowedBinaryAUROC
K, W = 64, 1 << 16
m = WindowedBinaryAUROC(num_tasks=K, max_num_samples=W, device="cuda")
m.update(torch.rand((K, W), device="cuda", generator=g), torch.randint(0, 2, (K, W), device="cuda", generator=g),
         torch.rand((K, W), device="cuda", generator=g))
m.compute(); torch.cuda.synchronize()
print("WindowedBinaryAUROC.compute (64 tasks x 2^16, weighted)", f"{tune.timeit(lambda: m.compute()):.3f} ms")
