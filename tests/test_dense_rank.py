"""te_dense_rank: the 64-bit road (float64 / int64 scores) against numpy's unique, bit for bit, and
the float64 functional entry points against the oracle on scores float32 cannot hold."""
import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from torcheval_b200 import ops as _ops

    return _ops


def _cases_f64(rs):
    n = 100_003
    yield "uniform", rs.random(n)
    yield "ties", rs.integers(0, 50, n) / 49.0
    yield "close_pairs", np.repeat(rs.random(n // 2 + 1), 2)[:n] + np.tile([0.0, 2.0**-50], n // 2 + 1)[:n]
    x = rs.random(n)
    x[::3] = np.nextafter(x[::3], 2.0)  # neighbours that differ in the last bit only
    yield "ulp_neighbours", np.concatenate([x, x[: n // 2]])
    yield "signed", rs.standard_normal(n) * 1e3
    z = rs.standard_normal(n)
    z[::7] = 0.0
    z[3::7] = -0.0
    z[5::11] = np.inf
    z[6::13] = -np.inf
    yield "zeros_infs", z
    yield "tiny_and_huge", np.concatenate([rs.random(n) * 1e-300, rs.random(n) * 1e300, [5e-324, -5e-324, 0.0]])
    yield "one", np.array([0.25])
    yield "constant", np.full(4097, 0.1)
    yield "chunk_edges", np.repeat(np.arange(5, dtype=np.float64), 2048)[:-1] / 7.0


def _check(ops, x, dtype):
    t = torch.as_tensor(np.ascontiguousarray(x)).to(dtype).cuda()
    rank, values = ops.dense_rank(t, with_values=True)
    torch.cuda.synchronize()
    host = t.cpu().numpy()
    want_vals, want_inv = np.unique(host, return_inverse=True)  # (-0.0 == 0.0 in numpy too)
    np.testing.assert_array_equal(rank.cpu().numpy().astype(np.int64), want_inv.reshape(host.shape))
    got_vals = values.cpu().numpy()
    assert got_vals.shape == want_vals.shape
    np.testing.assert_array_equal(got_vals, want_vals)  # (compares -0.0 equal to 0.0)
    rank2, none = ops.dense_rank(t)
    assert none is None
    np.testing.assert_array_equal(rank2.cpu().numpy(), rank.cpu().numpy())


def test_dense_rank_f64_vs_numpy_unique(ops):
    rs = np.random.default_rng(20261107)
    for name, x in _cases_f64(rs):
        _check(ops, x, torch.float64)


def test_dense_rank_i64_vs_numpy_unique(ops):
    rs = np.random.default_rng(5)
    big = rs.integers(-(2**62), 2**62, 70_001)
    _check(ops, big, torch.int64)
    _check(ops, np.concatenate([big, big[:999], [np.iinfo(np.int64).min, np.iinfo(np.int64).max, 0, -1, 1]]), torch.int64)
    _check(ops, rs.integers(2**24, 2**24 + 300, 50_000), torch.int64)


def test_dense_rank_shape_and_empty(ops):
    x = torch.rand((7, 1001), dtype=torch.float64, device="cuda")
    rank, values = ops.dense_rank(x, with_values=True)
    assert rank.shape == x.shape and rank.dtype == torch.int32
    _, inv = np.unique(x.cpu().numpy(), return_inverse=True)
    np.testing.assert_array_equal(rank.cpu().numpy(), inv.reshape(7, 1001))
    e, v = ops.dense_rank(torch.empty(0, dtype=torch.float64, device="cuda"), with_values=True)
    assert e.numel() == 0 and v.numel() == 0


def test_float64_scores_through_the_functional_api(ops):
    """float64 probabilities that differ below float32 resolution: AUROC / AUPRC / PR curve must see
    the float64 order and ties (reference: torch.sort on the float64 tensor, auroc.py:121)."""
    from torcheval_b200.metrics import functional as F

    rs = np.random.default_rng(11)
    n = 200_001
    base = rs.integers(0, 2000, n) / 2000.0
    x = base + rs.integers(0, 4, n) * 2.0**-40  # four float64 values per float32 value
    y = (rs.random(n) < 0.2 + 0.6 * x).astype(np.int64)
    xt, yt = torch.as_tensor(x).cuda(), torch.as_tensor(y).cuda()
    # the C oracle takes float32 scores: hand it the dense ranks (same order, same ties; < 2^24, exact)
    vals, inv = np.unique(x, return_inverse=True)
    assert vals.size < 2**24
    want = oracle.binary_auroc(inv.astype(np.float32), y)[0]
    got = float(F.binary_auroc(xt, yt))
    np.testing.assert_allclose(got, want, rtol=1e-12)
    from oracle import oracle_torch

    np.testing.assert_allclose(got, float(oracle_torch.binary_auroc(torch.as_tensor(x), torch.as_tensor(y))), rtol=1e-12)
    assert abs(got - oracle.binary_auroc(x.astype(np.float32), y)[0]) > 1e-9  # float32 would have merged ties
    got_ap = float(F.binary_auprc(xt, yt))
    np.testing.assert_allclose(got_ap, oracle.binary_auprc(inv.astype(np.float32), y), rtol=1e-5)
    p, r, thr = F.binary_precision_recall_curve(xt, yt)
    assert thr.dtype == torch.float64
    np.testing.assert_array_equal(thr.cpu().numpy(), np.unique(x))
    # int64 scores beyond 2^24
    xi = rs.integers(2**40, 2**40 + 5000, n)
    got_i = float(F.binary_auroc(torch.as_tensor(xi).cuda(), yt))
    np.testing.assert_allclose(got_i, oracle.binary_auroc((xi - 2**40).astype(np.float32), y)[0], rtol=1e-12)
    # negative float32 scores: thresholds come back in the input's dtype
    xn = rs.standard_normal(5000).astype(np.float32)
    p, r, thr = F.binary_precision_recall_curve(torch.as_tensor(xn).cuda(), yt[:5000])
    assert thr.dtype == torch.float32
    np.testing.assert_array_equal(thr.cpu().numpy(), np.unique(xn))
