"""The slab exchange of the distributed sample sort (csrc/rank_exchange.cuh) on ONE GPU: W
simulated ranks, every "peer" inbox an ordinary buffer of the same device, so that the sender
kernel, the header protocol, the slab view of the receive side and the fix-up are checked
without NCCL.  tests/test_distributed_gpu.py runs the same kernels across real peers."""
import numpy as np
import pytest
import torch

import oracle
from cpu_backend import _exact_2u

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from torcheval_b200 import ops as _ops

    return _ops


def exchange(ops, shards, cap=None, slack=2.0):
    """shards: list of (x, y) numpy pairs, one per simulated rank.  Returns the combined float64[8]
    and the per-rank rows."""
    from torcheval_b200 import _native

    lib = _native.load()
    W = len(shards)
    dev = torch.device("cuda")
    xs = [torch.from_numpy(x).to(dev) for x, _ in shards]
    ys = [torch.from_numpy(y).to(dev) for _, y in shards]
    n_max = max(x.numel() for x in xs)
    if cap is None:
        cap = int(lib.te_slab_capacity(int(slack / 1.3 * n_max) + 1, W))
    nbytes = int(lib.te_slab_inbox_bytes(W, cap))
    inboxes = [torch.full((nbytes // 4,), -7, dtype=torch.int32, device=dev) for _ in range(W)]  # stale garbage
    ptrs = [t.data_ptr() for t in inboxes]
    # splitters from regular samples of every shard, as distributed.device_splitters does
    s = max(1, 8192 // W)
    pool = []
    for x in xs:
        if x.numel():
            idx = torch.div(torch.arange(s, device=dev) * x.numel(), s, rounding_mode="floor")
            pool.append(x.view(torch.int32)[idx])
        else:
            pool.append(torch.full((s,), -1, dtype=torch.int32, device=dev))
    sp32, _ = ops.select_splitters(torch.stack(pool), W)
    work = torch.zeros(32, dtype=torch.int32, device=dev)
    for q in range(W):
        ops.rank_pack_scatter(xs[q], ys[q], sp32 if W > 1 else None, W, q, ptrs, cap, work)
    rows = torch.stack([ops.binary_auroc_from_slabs(inboxes[b], W, cap) for b in range(W)])
    out = ops.combine_auroc_stats(rows, 0)
    return out.cpu().numpy(), rows.cpu().numpy(), cap


@pytest.mark.parametrize("W", [1, 2, 5, 8, 16])
def test_slab_exchange_matches_oracle(ops, W):
    rs = np.random.default_rng(W)
    n = 700_001
    cases = {
        "uniform": rs.random(n, dtype=np.float32),
        "ties": (rs.integers(0, 999, n) / 999).astype(np.float32),
        "binade": (0.5 + 0.5 * rs.random(n, dtype=np.float32)).astype(np.float32),
    }
    y = rs.integers(0, 2, n)
    for name, x in cases.items():
        cut = [0] + sorted(rs.integers(0, n, W - 1).tolist()) + [n]  # uneven shards, some may be empty
        if W >= 5:
            cut[2] = cut[1]  # an empty shard
        shards = [(x[cut[q]:cut[q + 1]], y[cut[q]:cut[q + 1]]) for q in range(W)]
        out, rows, cap = exchange(ops, shards, slack=float(W) + 1.0)  # shards are uneven on purpose: room for all
        assert out[4] == 0 and (rows[:, 4] == 0).all(), (name, rows[:, 4])
        want, tp, fp = oracle.binary_auroc(x, y)
        assert (out[1], out[2]) == (tp, fp), name
        assert (int(out[6]) << 32) | int(out[7]) == _exact_2u(x, y), name
        np.testing.assert_allclose(out[0], want, rtol=1e-12, err_msg=name)
        assert rows[:, 7].sum() == n  # every record arrived exactly once
        # rank order = score order: the key ranges of the ranks do not overlap
        assert (rows[:, 1] + rows[:, 2] == rows[:, 7]).all()


def test_slab_exchange_flags(ops):
    rs = np.random.default_rng(3)
    n = 200_000
    x = rs.random(n, dtype=np.float32)
    y = rs.integers(0, 2, n)
    halves = lambda a: [a[: n // 2], a[n // 2:]]  # noqa: E731
    # a slab too small for what one sender has for one receiver: overflow is reported by every receiver
    out, rows, _ = exchange(ops, list(zip(halves(x), halves(y))), cap=24576)
    assert (rows[:, 4].astype(np.int64) & 0x100).all()
    # negative score on rank 1, soft label on rank 0, score >= 2 on rank 1: all receivers see all flags
    xb, yb = x.copy(), y.astype(np.float32)
    xb[n - 5] = -1.0
    yb[7] = 0.5
    xb[n - 9] = 2.5
    out, rows, _ = exchange(ops, list(zip(halves(xb), halves(yb))))
    assert (rows[:, 4].astype(np.int64) == 7).all(), rows[:, 4]
