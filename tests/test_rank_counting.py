"""Parity of the counting-stage AUROC (TE_RANK_COUNTING: two partition passes + shared-memory
histograms per key segment, csrc/rank_leaf.cuh) against the oracle, the reference fixtures and
the four-pass sort it replaces.  2U is an exact integer and must match digit for digit."""
import numpy as np
import pytest
import torch

import oracle
from cpu_backend import _exact_2u

pytestmark = pytest.mark.gpu

SORT, COUNTING = 1, 2


@pytest.fixture(scope="module")
def ops():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from torcheval_b200 import ops as _ops

    return _ops


def dev(a, dtype=None):
    t = torch.as_tensor(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda()


def check(ops, x, y, what=""):
    a = ops.binary_auroc_packed(dev(x), dev(y), COUNTING).cpu().numpy()
    assert a[4] == 0, what
    want, tp, fp = oracle.binary_auroc(x, y) if x.size else (0.5, 0.0, 0.0)
    assert (a[1], a[2]) == (tp, fp), what
    u2 = _exact_2u(x, y) if x.size else 0
    assert (int(a[5]) << 32) | int(a[6]) == u2, what
    np.testing.assert_allclose(a[0], want, rtol=1e-12, atol=1e-15, err_msg=what)
    assert a[7] == x.size


def test_counting_golden(ops, golden):
    for name in golden.cases("binary"):
        c = golden.case(f"binary/{name}")
        x, y = c["x"], c["y"]
        a = ops.binary_auroc_packed(dev(x, torch.float32), dev(y), COUNTING).cpu().numpy()
        if (np.signbit(x) & (x != 0)).any():
            assert int(a[4]) & 2, name
            continue
        if (x >= 2.0).any():
            assert int(a[4]) == 4, name
            continue
        assert a[4] == 0, name
        np.testing.assert_allclose(a[0], c["auroc_f64"], rtol=1e-6, err_msg=name)


def test_counting_edge_sizes(ops):
    rs = np.random.default_rng(5)
    for n in [0, 1, 2, 3, 4, 5, 7, 31, 33, 511, 513, 4095, 4097, 8191, 8192, 8193, 16385, 24577, 131071, 131073]:
        x = (rs.integers(0, 64, n) / 64).astype(np.float32)
        y = rs.integers(0, 2, n)
        check(ops, x, y, f"ties n={n}")
        x = rs.random(n, dtype=np.float32)
        check(ops, x, y, f"uniform n={n}")


@pytest.mark.parametrize("n", [100003, (1 << 20) + 5, (1 << 22) + 1])
def test_counting_distributions(ops, n):
    """Uniform scores (every binade, mostly light segments), one binade, a narrow band (a few
    segments, each counted in several pieces and merged), two values, a constant, bf16-valued
    scores, all-positive / all-negative labels, 1 % positives, tiny scores next to zero and -0.0."""
    rs = np.random.default_rng(n)
    u = rs.random(n, dtype=np.float32)
    y = rs.integers(0, 2, n)
    cases = {
        "uniform": u,
        "binade": (0.5 + 0.5 * u).astype(np.float32),
        "narrow": (0.7 + 1e-4 * u).astype(np.float32),
        "one_segment": (np.float32(0.75) + (rs.integers(0, 1 << 14, n).astype(np.float32) * np.float32(2.0 ** -24))).astype(np.float32),
        "two_values": np.where(u < 0.3, np.float32(0.25), np.float32(0.75)),
        "constant": np.full(n, 0.375, dtype=np.float32),
        "bf16": (u.view(np.uint32) & np.uint32(0xFFFF0000)).view(np.float32),
        "tiny": (u * np.float32(1e-38)).astype(np.float32),
        "just_below_2": np.minimum((1.0 + u).astype(np.float32), np.nextafter(np.float32(2), np.float32(0))),
    }
    for name, x in cases.items():
        check(ops, x, y, name)
    z = u.copy()
    z[::3] = 0.0
    z[1::7] = -0.0
    check(ops, z, y, "zeros")
    check(ops, u, np.ones(n, dtype=np.int64), "all_pos")
    check(ops, u, np.zeros(n, dtype=np.int64), "all_neg")
    check(ops, u, (rs.random(n) < 0.01).astype(np.int64), "skewed_labels")


def test_counting_label_dtypes_and_flags(ops):
    rs = np.random.default_rng(11)
    n = 300_007
    x = rs.random(n, dtype=np.float32)
    y = rs.integers(0, 2, n)
    want = ops.binary_auroc_packed(dev(x), dev(y), SORT).cpu().numpy()
    for dt in (torch.int64, torch.int32, torch.uint8, torch.float32, torch.float64):
        a = ops.binary_auroc_packed(dev(x), dev(y, dt), COUNTING).cpu().numpy()
        np.testing.assert_array_equal(a, want)
    # scores >= 2.0 need the 32nd key bit: reported, and the functional falls back to the full sort
    xw = x.copy()
    xw[12345] = 3.5
    a = ops.binary_auroc_packed(dev(xw), dev(y), COUNTING).cpu().numpy()
    assert int(a[4]) == 4
    xn = x.copy()
    xn[5] = -0.25
    assert int(ops.binary_auroc_packed(dev(xn), dev(y), COUNTING).cpu().numpy()[4]) & 2
    ys = y.astype(np.float32)
    ys[7] = 0.5
    assert int(ops.binary_auroc_packed(dev(x), dev(ys), COUNTING).cpu().numpy()[4]) & 1


def test_counting_matches_sort_2p26_and_closed_form(ops):
    n, G = 1 << 26, 1 << 16
    g = torch.Generator(device="cuda").manual_seed(7)
    lev = torch.randint(0, G, (n,), device="cuda", generator=g)
    y = (torch.rand(n, device="cuda", generator=g) < 0.3).long()
    x = lev.float() / G
    a = ops.binary_auroc_packed(x, y, COUNTING).cpu().numpy()
    b = ops.binary_auroc_packed(x, y, SORT).cpu().numpy()
    np.testing.assert_array_equal(a, b)
    neg = torch.bincount(lev[y == 0], minlength=G).cpu().numpy()
    pos = torch.bincount(lev[y == 1], minlength=G).cpu().numpy()
    np.testing.assert_allclose(a[0], oracle.auroc_from_level_hist(neg, pos), rtol=1e-9)
    # uniform floats: every exponent, 2^26 keys
    x = torch.rand(n, device="cuda", generator=g)
    a = ops.binary_auroc_packed(x, y, COUNTING).cpu().numpy()
    b = ops.binary_auroc_packed(x, y, SORT).cpu().numpy()
    np.testing.assert_array_equal(a, b)


def test_functional_falls_back_for_wide_scores(ops, monkeypatch):
    """scores in [0, 4): the counting stage reports TE_RANK_FLAG_WIDE_SCORE, the host re-issues
    the call with the full sort — same value as the oracle either way."""
    import torcheval_b200._ranking as R

    rs = np.random.default_rng(3)
    n = 200_001
    x = (rs.random(n, dtype=np.float32) * 4).astype(np.float32)
    y = rs.integers(0, 2, n)
    calls = []
    real = ops.binary_auroc_packed

    def spy(s, t, variant=0):
        calls.append(variant)
        return real(s, t, COUNTING if variant == 0 else variant)

    monkeypatch.setattr(R.ops, "binary_auroc_packed", spy)
    got = R.binary_auroc(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda())
    assert calls == [0, SORT]
    np.testing.assert_allclose(got.item(), oracle.binary_auroc(x, y)[0], rtol=1e-12)


def test_counting_from_keys(ops):
    """te_binary_auroc_from_keys_variant (the receive side of the distributed sample sort): packed
    keys in, both structures agree with each other and with the oracle; keys >= 2^31 are flagged."""
    rs = np.random.default_rng(17)
    for n in (0, 5, 8193, 300_001, (1 << 21) + 3):
        x = rs.random(n, dtype=np.float32)
        x[: n // 3] = np.round(x[: n // 3] * 512) / 512
        y = rs.integers(0, 2, n)
        keys = ((x.view(np.uint32) << np.uint32(1)) | y.astype(np.uint32)).view(np.int32)
        a = ops.binary_auroc_from_keys(dev(keys), None, COUNTING).cpu().numpy()
        b = ops.binary_auroc_from_keys(dev(keys), None, SORT).cpu().numpy()
        np.testing.assert_array_equal(a, b)
        if n:
            want, tp, fp = oracle.binary_auroc(x, y)
            assert (a[1], a[2]) == (tp, fp)
            assert (int(a[5]) << 32) | int(a[6]) == _exact_2u(x, y)
    keys = keys.copy()
    keys[3] = np.int32(-5)  # bit 31 set
    assert int(ops.binary_auroc_from_keys(dev(keys), None, COUNTING).cpu().numpy()[4]) == 4


def check_ap(ops, x, y, what=""):
    p = ops.binary_auprc_packed(dev(x), dev(y), COUNTING).cpu().numpy()
    q = ops.binary_auprc_packed(dev(x), dev(y), SORT).cpu().numpy()
    assert p[4] == 0, what
    assert (p[1], p[2]) == (q[1], q[2]) == (float(y.sum()), float(x.size - y.sum())), what
    want = oracle.binary_auprc(x, y) if x.size else -0.0
    np.testing.assert_allclose(p[0], want, rtol=1e-9, atol=1e-15, err_msg=what)
    np.testing.assert_allclose(p[0], q[0], rtol=1e-12, atol=1e-15, err_msg=what)
    if y.sum() == 0:
        assert p[0] == 0 and np.signbit(p[0]), what  # the reference's -0.0 (precision_recall_curve.py:227-229)


def test_counting_auprc(ops):
    """Average precision on the counting stage (two passes over the grouped keys) against the oracle
    and the sort-based kernel: edge sizes, ties, segments of every kind (tiny, single-piece,
    multi-piece), degenerate label sets."""
    rs = np.random.default_rng(9)
    for n in [1, 2, 5, 33, 129, 4097, 24577, 131073]:
        y = rs.integers(0, 2, n)
        check_ap(ops, (rs.integers(0, 64, n) / 64).astype(np.float32), y, f"ties n={n}")
        check_ap(ops, rs.random(n, dtype=np.float32), y, f"uniform n={n}")
    n = (1 << 21) + 7
    u = rs.random(n, dtype=np.float32)
    y = rs.integers(0, 2, n)
    cases = {
        "uniform": u,
        "narrow": (0.7 + 1e-4 * u).astype(np.float32),
        "one_segment": (np.float32(0.75) + (rs.integers(0, 1 << 14, n).astype(np.float32) * np.float32(2.0 ** -24))).astype(np.float32),
        "two_values": np.where(u < 0.3, np.float32(0.25), np.float32(0.75)),
        "constant": np.full(n, 0.375, dtype=np.float32),
        "tiny": (u * np.float32(1e-38)).astype(np.float32),
    }
    for name, x in cases.items():
        check_ap(ops, x, y, name)
    check_ap(ops, u, np.ones(n, dtype=np.int64), "all_pos")
    check_ap(ops, u, np.zeros(n, dtype=np.int64), "all_neg")
    check_ap(ops, u, (rs.random(n) < 0.01).astype(np.int64), "skewed_labels")
    xw = u.copy()
    xw[5] = 2.5
    assert int(ops.binary_auprc_packed(dev(xw), dev(y), COUNTING).cpu().numpy()[4]) == 4
