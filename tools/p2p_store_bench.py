"""What one GPU can store into its peers' memory (the ceiling of the slab exchange's sender):
16-byte st.global per lane against bulk async copies, local against peers.  Run under torchrun."""
import ctypes
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

from torcheval_b200 import _native as N

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
lib = N.load()
words = 1 << 26  # 256 MB per destination buffer
buf = symm_mem.empty(words, dtype=torch.int32, device=dev)
hdl = symm_mem.rendezvous(buf, dist.group.WORLD)
ptrs = [int(p) for p in hdl.buffer_ptrs]
sms = torch.cuda.get_device_properties(dev).multi_processor_count
stream = torch.cuda.current_stream(dev).cuda_stream
lines = []


def run(tag, dsts, mode, grid):
    arr = (ctypes.c_uint64 * len(dsts))(*dsts)
    def call():
        N.check("p2p", lib.te_debug_p2p_store_bench(arr, len(dsts), words, mode, grid, stream))
    call()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    gb = 4.0 * words * len(dsts) / 1e9
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    lines.append(f"{tag:34s} mode {mode} grid {grid:4d}: {float(t.item()):7.3f} ms (max over ranks)  {gb / float(t.item()) * 1e3:7.1f} GB/s out of every GPU")
    dist.barrier()


peers = [ptrs[q] for q in range(world) if q != rank]
for mode in (0, 1):
    for grid in (sms, 2 * sms):
        run("local buffer", [ptrs[rank]], mode, grid)
        if peers:
            run(f"{len(peers)} peers (all ranks at once)", peers, mode, grid)
            run("one peer (rank + 1)", [ptrs[(rank + 1) % world]], mode, grid)
if rank == 0:
    print("\n".join(lines))
dist.destroy_process_group()
