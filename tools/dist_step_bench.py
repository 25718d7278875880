"""Sharded exact AUROC only (no secondaries): ms per step and the per-kernel breakdown of rank 0,
under torchrun with 2+ GPUs.  LOG2_N = log2 of the TOTAL number of scores (default 30)."""
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))


def main():
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from bench import KernelProfile
    from torcheval_b200 import distributed as tdist

    world, rank = dist.get_world_size(), dist.get_rank()
    n = (1 << int(os.environ.get("LOG2_N", "30"))) // world
    g = torch.Generator(device=dev).manual_seed(20261106 + rank)
    x = torch.rand(n, device=dev, generator=g)
    y = torch.randint(0, 2, (n,), device=dev, generator=g)
    for _ in range(3):
        out = tdist.sharded_binary_auroc_async(x, y)
    torch.cuda.synchronize()
    dist.barrier()
    steps = 20
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with KernelProfile() as prof:
        e0.record()
        for _ in range(steps):
            out = tdist.sharded_binary_auroc_async(x, y)
        e1.record()
        torch.cuda.synchronize()
        k = prof.read()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    rows = [None] * world
    dist.all_gather_object(rows, {kn: round(v["total_ms"] / steps, 3) for kn, v in k.items()})
    if rank == 0:
        print(f"world={world} n_total=2^{(n * world).bit_length() - 1}: {ms.item():.3f} ms/step  flags={out[4].item()} auroc={out[0].item():.9f}")
        for r in (0, world - 1):
            print(f"  rank {r}: {rows[r]}  sum={sum(rows[r].values()):.3f}")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
