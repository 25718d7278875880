"""Host->device copy rate of every rank's pinned shard, alone and with all ranks copying at once,
with the pinned pages placed (a) wherever the launcher left the process and (b) on the NUMA node
of the rank's GPU.  Run under torchrun; explains the update() phase of the N-GPU end-to-end step.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29533 tools/h2d_probe.py [log2_records_per_rank]
"""
import os
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist

from bench import gpu_numa_cpus  # noqa: E402


def main():
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 27
    n = 1 << log2n
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    node, cpus = gpu_numa_cpus(local)
    before = sorted(os.sched_getaffinity(0))
    lines = [f"rank {rank}: gpu {local} numa_node {node} node_cpus {len(cpus) if cpus else 0} affinity_at_start {len(before)} cpus"]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def alloc():
        xh = torch.empty(n, dtype=torch.float32, pin_memory=True)
        yh = torch.empty(n, dtype=torch.int64, pin_memory=True)
        xh.fill_(0.5)
        yh.fill_(1)
        return xh, yh

    xd = torch.empty(n, dtype=torch.float32, device=dev)
    yd = torch.empty(n, dtype=torch.int64, device=dev)

    def copy_ms(xh, yh, reps=3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            xd.copy_(xh)
            yd.copy_(yh)
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) * 1e3 / reps

    def measure(tag, xh, yh):
        copy_ms(xh, yh, 1)
        solo = None
        for q in range(world):  # one rank at a time
            barrier()
            if q == rank:
                solo = copy_ms(xh, yh)
        barrier()
        together = copy_ms(xh, yh)  # all ranks at once
        barrier()
        gb = 12 * n / 1e9
        lines.append(f"rank {rank} {tag}: alone {solo:7.2f} ms ({gb / solo * 1e3:5.1f} GB/s)   all {world} ranks at once {together:7.2f} ms ({gb / together * 1e3:5.1f} GB/s)")

    xh, yh = alloc()
    measure("pages as launched ", xh, yh)
    del xh, yh
    if cpus:
        os.sched_setaffinity(0, cpus)
        xh, yh = alloc()
        measure("pages on GPU node ", xh, yh)
        os.sched_setaffinity(0, before)
        measure("same, unbound copy", xh, yh)
    if world > 1:
        out = [None] * world
        dist.all_gather_object(out, lines)
        if rank == 0:
            for ls in out:
                print("\n".join(ls))
        dist.destroy_process_group()
    else:
        print("\n".join(lines))


if __name__ == "__main__":
    main()
