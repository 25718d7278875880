"""Single-GPU timing of the two halves of the slab exchange at the 8-GPU shape of the headline
config (2^27 records per rank): the sender kernel with all eight "peer" inboxes on this device
(HBM side only, no NVLink), and the receive side over an inbox holding 2^27 records."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from bench import KernelProfile
from torcheval_b200 import _native, ops

lib = _native.load()
dev = torch.device("cuda")
W, n = 8, 1 << 27
g = torch.Generator(device=dev).manual_seed(1)
x = torch.rand(n, device=dev, generator=g)
y = torch.randint(0, 2, (n,), device=dev, generator=g)
sp = (torch.arange(1, W, device=dev, dtype=torch.float32) / W).view(torch.int32) << 1  # key-space splitters k/8


def timed(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with KernelProfile() as prof:
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        k = prof.read()
    return e0.elapsed_time(e1) / reps, {kn: round(v["total_ms"] / reps, 3) for kn, v in k.items()}


# ---- sender: 2^27 uniform scores, an even 8-way split
cap = int(lib.te_slab_capacity(int(2.0 / 1.3 * n) + 1, W))
nbytes = int(lib.te_slab_inbox_bytes(W, cap))
inboxes = [torch.empty(nbytes // 4, dtype=torch.int32, device=dev) for _ in range(W)]
work = torch.zeros(32, dtype=torch.int32, device=dev)
ms, k = timed(lambda: ops.rank_pack_scatter(x, y, sp, W, 3, [t.data_ptr() for t in inboxes], cap, work))
print(f"sender  2^27 -> 8 local inboxes: {ms:.3f} ms  {k}")
del inboxes
torch.cuda.empty_cache()

# ---- receiver: an inbox with 2^27 records spread over 8 slabs, key range of rank r of 8
for r in (0, 7):
    xr = (x + r) / W
    cap = int(lib.te_slab_capacity(int(2.0 / 1.3 * n) + 1, W))
    inbox = torch.empty(int(lib.te_slab_inbox_bytes(W, cap)) // 4, dtype=torch.int32, device=dev)
    spr = torch.full((W - 1,), 0x7FFFFFFE, dtype=torch.int32, device=dev)  # everything goes to bucket 0 = this inbox
    shard = n // W
    for q in range(W):  # eight senders, one slab each
        ptrs = [inbox.data_ptr()] * W
        ops.rank_pack_scatter(xr[q * shard:(q + 1) * shard], y[q * shard:(q + 1) * shard], spr, W, q, ptrs, cap, work)
    ms, k = timed(lambda: ops.binary_auroc_from_slabs(inbox, W, cap))
    out = ops.binary_auroc_from_slabs(inbox, W, cap).cpu().numpy()
    ref = ops.binary_auroc_packed(xr, y, 1).cpu().numpy()
    assert out[4] == 0 and out[7] == n and (out[:4] == ref[:4]).all(), (out, ref)
    print(f"receiver rank-{r}-like inbox, 2^27 records in 8 slabs: {ms:.3f} ms  {k}")
    del inbox
