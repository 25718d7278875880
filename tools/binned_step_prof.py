"""Host-side cost of one BinaryBinnedAUPRC step (construct + update + sync_and_compute) at a size
where the kernels are short: what bounds `secondary_binned` at 8 GPUs.  Run alone or under torchrun."""
import cProfile
import io
import os
import pstats
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist

from torcheval_b200.metrics import BinaryBinnedAUPRC
from torcheval_b200.metrics.toolkit import sync_and_compute

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
n = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 20)
x = torch.rand(n, device=dev)
y = torch.randint(0, 2, (n,), device=dev)
T = 10_000


def step():
    m = BinaryBinnedAUPRC(threshold=T, device=dev)
    m.update(x, y)
    return sync_and_compute(m)


def timed(fn, reps):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    host = (time.perf_counter() - t0) * 1e3 / reps
    torch.cuda.synchronize()
    return host, (time.perf_counter() - t0) * 1e3 / reps


for _ in range(20):
    step()
host, total = timed(step, 200)
m = BinaryBinnedAUPRC(threshold=T, device=dev)
parts = {
    "construct": timed(lambda: BinaryBinnedAUPRC(threshold=T, device=dev), 200),
    "update": timed(lambda: m.update(x, y), 200),
    "sync_and_compute": timed(lambda: sync_and_compute(m), 200),
    "compute": timed(lambda: m.compute(), 200),
}
pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    step()
pr.disable()
torch.cuda.synchronize()
if rank == 0:
    print(f"world {world}, n = 2^{n.bit_length() - 1}: step host {host:.3f} ms, host+device {total:.3f} ms")
    for k, (h, t) in parts.items():
        print(f"  {k:18s} host {h:.3f} ms   host+device {t:.3f} ms")
    s = io.StringIO()
    pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(28)
    print("\n".join(l[:150] for l in s.getvalue().splitlines()[:50]))
if world > 1:
    dist.destroy_process_group()
