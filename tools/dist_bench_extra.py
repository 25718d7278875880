"""Multi-GPU paths other than the headline (run under torchrun): rank-sharded exact BinaryAUPRC
(two-phase integrate) and class-sharded one-vs-rest MulticlassAUROC / MultilabelAUPRC through
toolkit.sync_and_compute.  Device-timed, max over ranks; rank 0 prints one line per workload."""
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))


def timed(fn, world, dev, reps=5):
    for _ in range(2):
        out = fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()), out


def main():
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    from torcheval_b200 import distributed as tdist
    from torcheval_b200.metrics import MulticlassAUROC, MultilabelAUPRC
    from torcheval_b200.metrics.toolkit import sync_and_compute

    g = torch.Generator(device=dev).manual_seed(20261106 + rank)
    n = (1 << 30) // world
    x = torch.rand(n, device=dev, generator=g)
    y = torch.randint(0, 2, (n,), device=dev, generator=g)
    ms, out = timed(lambda: tdist.sharded_binary_auprc(x, y), world, dev)
    if rank == 0:
        print(f"sharded BinaryAUPRC   N=2^30 over {world} GPUs: {ms:8.3f} ms  {(1 << 30) / ms / 1e6:7.1f} G scores/s  ap={float(out[0]):.6f}", flush=True)
    ms, out = timed(lambda: tdist.sharded_binary_auroc(x, y), world, dev)
    if rank == 0:
        print(f"sharded BinaryAUROC   N=2^30 over {world} GPUs: {ms:8.3f} ms  {(1 << 30) / ms / 1e6:7.1f} G scores/s  auroc={float(out[0]):.6f}", flush=True)
    del x, y
    C, nrows = 1000, (1 << 22) // world
    xm = torch.rand((nrows, C), device=dev, generator=g)
    ym = torch.randint(0, C, (nrows,), device=dev, generator=g)
    m = MulticlassAUROC(num_classes=C, average="macro", device=dev).update(xm, ym)
    ms, out = timed(lambda: sync_and_compute(m), world, dev, reps=3)
    if rank == 0:
        print(f"class-sharded MulticlassAUROC  N=2^22 x C=1000 over {world} GPUs: {ms:8.3f} ms  {(1 << 22) * C / ms / 1e6:7.1f} G scores/s  macro={float(out):.6f}", flush=True)
    del m, ym
    nl = nrows // 4
    yl = torch.randint(0, 2, (nl, C), device=dev, generator=g)
    ml = MultilabelAUPRC(num_labels=C, average="macro", device=dev).update(xm[:nl], yl)
    ms, out = timed(lambda: sync_and_compute(ml), world, dev, reps=3)
    if rank == 0:
        print(f"class-sharded MultilabelAUPRC  N=2^20 x L=1000 over {world} GPUs: {ms:8.3f} ms  {(1 << 20) * C / ms / 1e6:7.1f} G scores/s  macro={float(out):.6f}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
