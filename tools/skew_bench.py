"""Exact AUROC step time on score distributions with skewed radix digits (1 GPU).
Run twice to compare TE_B200_HOT_DIGITS=0/1 (the library reads the variable once)."""
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from torcheval_b200 import ops  # noqa: E402


def main():
    log2n = int(os.environ.get("LOG2_N", "28"))
    n = 1 << log2n
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(1)
    y = torch.randint(0, 2, (n,), device=dev, generator=g)
    u = torch.rand(n, device=dev, generator=g)
    cases = {
        "uniform": u,
        "bf16_scores": u.to(torch.bfloat16).float(),
        "fp16_scores": u.to(torch.float16).float(),
        "upper_half": 0.5 + 0.5 * u,
        "1000_values": torch.floor(u * 1000) / 1000,
        "constant": torch.full_like(u, 0.25),
        "sigmoid_logits": torch.sigmoid(4 * torch.randn(n, device=dev, generator=g)),
    }
    print(f"TE_B200_HOT_DIGITS={os.environ.get('TE_B200_HOT_DIGITS', '(default)')}  n=2^{log2n}")
    for name, x in cases.items():
        x = x.contiguous()
        for _ in range(2):
            out = ops.binary_auroc_packed(x, y)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            out = ops.binary_auroc_packed(x, y)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        print(f"  {name:16s} {ms:8.3f} ms/step  {n / ms / 1e6:7.2f} G scores/s  auroc={float(out[0]):.6f}")


if __name__ == "__main__":
    main()
