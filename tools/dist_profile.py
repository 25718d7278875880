"""Host-side cost of the sharded AUROC protocol: run under torchrun with 2+ GPUs.
Prints the phase timer lines (TE_B200_DIST_TIMING=1) and a torch.profiler table of one step."""
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
    from torcheval_b200 import distributed as tdist

    n = (1 << int(os.environ.get("LOG2_N", "30"))) // dist.get_world_size()
    g = torch.Generator(device=dev).manual_seed(20261106 + dist.get_rank())
    x = torch.rand(n, device=dev, generator=g)
    y = torch.randint(0, 2, (n,), device=dev, generator=g)
    for _ in range(3):
        tdist.sharded_binary_auroc(x, y)
    torch.cuda.synchronize()
    dist.barrier()
    from torch.profiler import ProfilerActivity, profile

    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        for _ in range(3):
            out = tdist.sharded_binary_auroc(x, y)
        torch.cuda.synchronize()
    if dist.get_rank() == 0:
        print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=40, max_name_column_width=60))
        prof.export_chrome_trace("gpurun_out/dist_trace.json")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
