"""Partition pass on ONE GPU: local output vs the peer-scatter code path pointed at local buffers
(isolates the kernel's own cost from NVLink).  TE_B200_PEER_VEC=0/1 is read once per process."""
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from torcheval_b200 import ops  # noqa: E402


def timeit(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    log2n = int(os.environ.get("LOG2_N", "27"))
    n = 1 << log2n
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(1)
    x = torch.rand(n, device=dev, generator=g)
    y = torch.randint(0, 2, (n,), device=dev, generator=g)
    print(f"n=2^{log2n} PEER_VEC={os.environ.get('TE_B200_PEER_VEC', '(default)')}")
    for world in (2, 4, 8, 16):
        qs = torch.quantile(x[: 1 << 20], torch.arange(1, world, device=dev) / world)
        sp = [((int(v) << 1) & 0xFFFFFFFE) for v in qs.view(torch.int32).tolist()]
        keys, _, counts = ops.rank_pack(x, y, sp)
        cl = counts.tolist()
        bufs = [torch.empty(c + 16, dtype=torch.int32, device=dev) for c in cl]
        dest, excl = [], 0
        for b, c in enumerate(cl):
            dest.append(bufs[b].data_ptr() + 4 * ((b % 4) - excl))
            excl += c
        t_pack = timeit(lambda: ops.rank_pack(x, y, sp))
        t_pack_h = timeit(lambda: ops.rank_pack(x, y, sp, want_hist=True))
        t_loc = timeit(lambda: ops.partition_u32(keys, sp, counts))
        t_sc = timeit(lambda: ops.partition_scatter_u32(keys, sp, counts, dest))
        print(f"  world={world:2d}  pack {t_pack:.3f} ms (+digit hist {t_pack_h:.3f})   partition(local) {t_loc:.3f} ms   scatter(to local buffers) {t_sc:.3f} ms")


if __name__ == "__main__":
    main()
