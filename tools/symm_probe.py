"""Probe: does torch symmetric memory (peer-mapped buffers) work on this box?  torchrun -n 2."""
import os

import torch
import torch.distributed as dist

local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
rank, world = dist.get_rank(), dist.get_world_size()
try:
    import torch.distributed._symmetric_memory as symm_mem

    t = symm_mem.empty(1 << 20, dtype=torch.int32, device=dev)
    hdl = symm_mem.rendezvous(t, dist.group.WORLD)
    t.fill_(rank + 1)
    hdl.barrier()
    peer = (rank + 1) % world
    pt = hdl.get_buffer(peer, (1 << 20,), torch.int32)
    got = int(pt[123].item())
    ptrs = [int(p) for p in hdl.buffer_ptrs]
    hdl.barrier()
    # write into the peer's buffer through the mapped pointer
    pt[:16] = 100 + rank
    hdl.barrier()
    mine = t[:16].tolist()
    print(f"rank {rank}: peer value {got} (want {peer + 1}), buffer_ptrs {[hex(p) for p in ptrs]}, after peer write {mine[:3]}", flush=True)
    import time
    big = symm_mem.empty(1 << 28, dtype=torch.int32, device=dev)
    t0 = time.perf_counter()
    h2 = symm_mem.rendezvous(big, dist.group.WORLD)
    torch.cuda.synchronize()
    print(f"rank {rank}: rendezvous of 1 GiB took {(time.perf_counter() - t0) * 1e3:.1f} ms", flush=True)
    src = torch.ones(1 << 28, dtype=torch.int32, device=dev)
    pb = h2.get_buffer(peer, (1 << 28,), torch.int32)
    h2.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    pb.copy_(src)
    ev1.record()
    torch.cuda.synchronize()
    print(f"rank {rank}: 1 GiB peer copy {ev0.elapsed_time(ev1):.3f} ms = {1.0737 / ev0.elapsed_time(ev1) * 1e3:.0f} GB/s", flush=True)
    h2.barrier()
except Exception as e:  # noqa: BLE001
    print(f"rank {rank}: symmetric memory unavailable: {type(e).__name__}: {e}", flush=True)
dist.destroy_process_group()
