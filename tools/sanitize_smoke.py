"""Small-n pass through every kernel family, for compute-sanitizer (memcheck / racecheck):
    compute-sanitizer --tool memcheck python tools/sanitize_smoke.py
Sizes straddle the tile sizes (8192, 24576) so partial tiles, multi-piece segments and the slab
paths are all exercised; results are checked against the oracle so a run that "passes" computed
something."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tests"))
import numpy as np
import torch

import oracle
from torcheval_b200 import _native, ops

lib = _native.load()

# compute-sanitizer is closed on this pool, so the script carries its own two checks: every
# scratch buffer handed to the library is (a) filled with 0xFF first — a kernel that relied on
# scratch being zero where it does not clear it computes garbage and the oracle comparison
# fails — and (b) wrapped in 4 KiB guard zones of 0xA5 that are verified at the end: a write past
# either end of the scratch block shows up.
_guards = []
_GUARD = 4096


def _guarded_scratch(nbytes, dev):
    nbytes = max(int(nbytes), 256)
    pad = (-nbytes) % 256
    buf = torch.empty(nbytes + pad + 2 * _GUARD, dtype=torch.uint8, device=dev)
    buf.fill_(0xFF)
    buf[:_GUARD] = 0xA5
    buf[_GUARD + nbytes + pad:] = 0xA5
    _guards.append((buf, nbytes + pad))
    return buf[_GUARD:_GUARD + nbytes + pad]


ops._scratch = _guarded_scratch
rs = np.random.default_rng(0)
dev = torch.device("cuda")
n = int(os.environ.get("SAN_N", "60001"))
x = rs.random(n, dtype=np.float32)
x[: n // 3] = np.round(x[: n // 3] * 32) / 32
y = rs.integers(0, 2, n)
X, Y = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev)
want, tp, fp = oracle.binary_auroc(x, y)
for variant in (1, 2):  # full sort, counting stage
    a = ops.binary_auroc_packed(X, Y, variant).cpu().numpy()
    assert a[4] == 0 and abs(a[0] - want) < 1e-12 and (a[1], a[2]) == (tp, fp), (variant, a)
band = (0.7 + 1e-5 * x).astype(np.float32)  # a few heavy segments: pieces + slabs + merge
a = ops.binary_auroc_packed(torch.from_numpy(band).to(dev), Y, 2).cpu().numpy()
assert abs(a[0] - oracle.binary_auroc(band, y)[0]) < 1e-12
p = ops.binary_auprc_packed(X, Y).cpu().numpy()
assert abs(p[0] - oracle.binary_auprc(x, y)) < 1e-9
g = ops.binary_auroc_general(X - 0.5, Y, torch.from_numpy(rs.random(n) + 0.5).to(dev)).cpu().numpy()
(pr, rc, th), flags = ops.binary_pr_curve(X, Y)
assert flags == 0
keys = ((x.view(np.uint32) << np.uint32(1)) | y.astype(np.uint32)).view(np.int32)
for variant in (1, 2):
    a = ops.binary_auroc_from_keys(torch.from_numpy(keys).to(dev), None, variant).cpu().numpy()
    assert abs(a[0] - want) < 1e-12
# slab exchange, three simulated ranks on this device
W = 3
cap = int(lib.te_slab_capacity(n, W))
inboxes = [torch.zeros(int(lib.te_slab_inbox_bytes(W, cap)) // 4, dtype=torch.int32, device=dev) for _ in range(W)]
sp = (torch.tensor([0.3, 0.6], device=dev).view(torch.int32) << 1)
work = torch.zeros(32, dtype=torch.int32, device=dev)
cut = [0, n // 5, n // 2, n]
for q in range(W):
    ops.rank_pack_scatter(X[cut[q]:cut[q + 1]], Y[cut[q]:cut[q + 1]], sp, W, q, [t.data_ptr() for t in inboxes], cap, work)
rows = torch.stack([ops.binary_auroc_from_slabs(inboxes[b], W, cap) for b in range(W)])
out = ops.combine_auroc_stats(rows, 0).cpu().numpy()
assert abs(out[0] - want) < 1e-12 and out[4] == 0
# one-vs-rest, confusion matrix (every variant), binned
xm, ym = rs.random((5001, 11), dtype=np.float32), rs.integers(0, 11, 5001)
o = ops.ovr_rank(torch.from_numpy(xm).to(dev), torch.from_numpy(ym).to(dev), False, 0).cpu().numpy()
assert np.allclose(o[:, 0], oracle.multiclass_auroc(xm, ym), rtol=1e-9)
for C, variants in ((10, (0, 1, 2, 3, 4)), (1000, (0, 1, 2, 5, 6))):
    pc, tc = rs.integers(0, C, n), rs.integers(0, C, n)
    wantcm = oracle.confusion_matrix(pc, tc, C)[0]
    for v in variants:
        cm = torch.zeros((C, C), dtype=torch.int64, device=dev)
        ops.cm_update_(cm, torch.from_numpy(pc).to(dev), torch.from_numpy(tc).to(dev), C, v)
        assert np.array_equal(cm.cpu().numpy(), wantcm), (C, v)
thr = np.linspace(0, 1, 1000, dtype=np.float32)
hist = torch.zeros((1, thr.size + 1, 2), dtype=torch.int64, device=dev)
ops.binned_hist_update_(hist, X, Y, torch.from_numpy(thr).to(dev))
assert np.array_equal(hist.cpu().numpy()[0], oracle.binned_hist(x, y, thr))
torch.cuda.synchronize()
for buf, nb in _guards:
    assert bool((buf[:_GUARD] == 0xA5).all()) and bool((buf[_GUARD + nb:] == 0xA5).all()), "scratch guard zone overwritten"
print(f"sanitize_smoke ok ({len(_guards)} guarded scratch blocks intact)")
