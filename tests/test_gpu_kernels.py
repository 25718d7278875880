"""Parity of the CUDA kernels (through the C ABI, via torcheval_b200.ops) against the oracle and
the committed reference fixtures.  Counts are bit-exact; AUROC within 1e-6 relative (fp64),
values the reference itself computes in fp32 within 1e-4."""
import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu

RTOL64 = 1e-6
RTOL32 = 1e-4


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


def close(a, b, rtol):
    np.testing.assert_allclose(np.float64(a), np.float64(b), rtol=rtol, atol=rtol * 1e-3)


# =============================================================================== K4
def run_cm(ops, pred, target, C, variant=0, cm=None):
    if cm is None:
        cm = torch.zeros((C, C), dtype=torch.int64, device="cuda")
    rng = ops.cm_update_(cm, pred, target, C, variant)
    torch.cuda.synchronize()
    return cm.cpu().numpy(), rng.cpu().numpy()


def cm_variants(C):
    v = [0, 1, 2]
    if C * C * 128 <= 200 * 1024:
        v.append(3)
    if C * C * 4 <= 200 * 1024:
        v.append(4)
    if C >= 300:
        v.append(5)  # cluster / distributed-smem bins
    if C <= 1240:
        v.append(6)  # bucketed two-pass (route by matrix slice, count in shared memory)
    return v


def test_cm_golden_all_variants(ops, golden):
    for name in golden.cases("cm"):
        c = golden.case(f"cm/{name}")
        C = c["cm"].shape[0]
        for v in cm_variants(C):
            cm, rng = run_cm(ops, dev(c["pred"]), dev(c["target"]), C, v)
            np.testing.assert_array_equal(cm, c["cm"], err_msg=f"{name} variant {v}")
            assert rng[0] == np.iinfo(np.int64).max and rng[1] == np.iinfo(np.int64).min


@pytest.mark.parametrize("C,n", [(2, 100003), (10, 1 << 20), (16, 999), (100, 300001), (1000, (1 << 21) + 3), (300, 70001), (1500, 1 << 20)])
def test_cm_random_vs_oracle(ops, C, n):
    rs = np.random.default_rng(C * 7 + n)
    p = rs.integers(0, C, n)
    t = rs.integers(0, C, n)
    want, bad, _ = oracle.confusion_matrix(p, t, C)
    assert not bad
    for v in cm_variants(C):
        cm, _ = run_cm(ops, dev(p), dev(t), C, v)
        np.testing.assert_array_equal(cm, want, err_msg=f"variant {v}")


def test_cm_skewed_diagonal(ops):
    rs = np.random.default_rng(3)
    n, C = 1 << 20, 1000
    t = rs.integers(0, C, n)
    p = np.where(rs.random(n) < 0.9, t, rs.integers(0, C, n))
    p[: n // 4] = 7
    t[: n // 4] = 7  # one very hot bin
    want, _, _ = oracle.confusion_matrix(p, t, C)
    for v in (1, 2, 5, 6):
        cm, _ = run_cm(ops, dev(p), dev(t), C, v)
        np.testing.assert_array_equal(cm, want)
    # every sample in ONE slice of the matrix (a single target class): the count pass deals all
    # CTAs onto that slice
    t2 = np.full(n, 993)
    want2, _, _ = oracle.confusion_matrix(p, t2, C)
    cm, _ = run_cm(ops, dev(p), dev(t2), C, 6)
    np.testing.assert_array_equal(cm, want2)


def test_cm_bucketed_unaligned_tail_and_accumulate(ops):
    rs = np.random.default_rng(12)
    C, n = 1000, (1 << 22) + 4099
    p, t = rs.integers(0, C, n), rs.integers(0, C, n)
    p[5], t[9], p[n - 1] = C, -2, 12345  # out of range: reported, not counted
    P, T = dev(p), dev(t)
    cm = torch.zeros((C, C), dtype=torch.int64, device="cuda")
    r1 = ops.cm_update_(cm, P[1:], T[1:], C, 6)   # 8-byte aligned only -> scalar loads
    r2 = ops.cm_update_(cm, P, T, C)              # AUTO: n >= 2^22 and C = 1000 -> bucketed
    torch.cuda.synchronize()
    want, bad, _ = oracle.confusion_matrix(p, t, C)
    want1, _, _ = oracle.confusion_matrix(p[1:], t[1:], C)
    assert bad
    np.testing.assert_array_equal(cm.cpu().numpy(), want + want1)
    assert r1.tolist() == [C, 12345, -2, -2] and r2.tolist() == [C, 12345, -2, -2]


def test_cm_full_size_checksums(ops):
    """BASELINE config 2 at full size (N = 2^28, C = 1000): the matrix against torch.bincount of
    target * C + pred on the same device data, for the bucketed path (what AUTO picks) and RED."""
    n, C = 1 << 28, 1000
    g = torch.Generator(device="cuda").manual_seed(20261106)
    p = torch.randint(0, C, (n,), device="cuda", generator=g)
    t = torch.randint(0, C, (n,), device="cuda", generator=g)
    want = torch.bincount(t * C + p, minlength=C * C).view(C, C)
    for v in (0, 1):
        cm = torch.zeros((C, C), dtype=torch.int64, device="cuda")
        ops.cm_update_(cm, p, t, C, v)
        assert torch.equal(cm, want), v
    assert int(want.sum()) == n


def test_cm_unaligned_views_and_accumulate(ops):
    rs = np.random.default_rng(11)
    C, n = 37, 50001
    p, t = rs.integers(0, C, n), rs.integers(0, C, n)
    P, T = dev(p), dev(t)
    cm = torch.zeros((C, C), dtype=torch.int64, device="cuda")
    ops.cm_update_(cm, P[1:], T[1:], C)       # 8-byte aligned only -> scalar path
    ops.cm_update_(cm, P[:1], T[:1], C)       # accumulates
    ops.cm_update_(cm, P, T, C)
    torch.cuda.synchronize()
    want, _, _ = oracle.confusion_matrix(p, t, C)
    np.testing.assert_array_equal(cm.cpu().numpy(), 2 * want)


def test_cm_out_of_range_report(ops):
    C = 5
    p = np.array([0, 1, 2, 3, 4, 9, 1, -3, 2], dtype=np.int64)
    t = np.array([0, 1, 2, 3, 4, 1, 7, 2, -1], dtype=np.int64)
    for v in cm_variants(C):
        cm, rng = run_cm(ops, dev(p), dev(t), C, v)
        assert list(rng) == [-3, 9, -1, 7]
        np.testing.assert_array_equal(cm, np.eye(C, dtype=np.int64))
    want, bad, orng = oracle.confusion_matrix(p, t, C)
    assert bad and orng[1] == 9 and orng[3] == 7
    np.testing.assert_array_equal(want, np.eye(C, dtype=np.int64))


def test_cm_empty(ops):
    cm, rng = run_cm(ops, torch.zeros(0, dtype=torch.int64, device="cuda"), torch.zeros(0, dtype=torch.int64, device="cuda"), 4)
    assert cm.sum() == 0


def test_cm_logits(ops, golden):
    for name in golden.cases("cm_logits"):
        c = golden.case(f"cm_logits/{name}")
        C = c["cm"].shape[0]
        cm = torch.zeros((C, C), dtype=torch.int64, device="cuda")
        ops.cm_update_logits_(cm, dev(c["logits"]), dev(c["target"]), C)
        np.testing.assert_array_equal(cm.cpu().numpy(), c["cm"])
    rs = np.random.default_rng(5)
    for n, C in [(5000, 1000), (3001, 33), (777, 64), (100, 65)]:
        lg = rs.standard_normal((n, C)).astype(np.float32)
        lg[::7, C // 2] = lg[::7].max(axis=1)  # ties -> first index
        lg[3, 1] = np.nan
        t = rs.integers(0, C, n)
        want, _, _ = oracle.confusion_matrix(oracle.argmax_rows(lg), t, C)
        cm = torch.zeros((C, C), dtype=torch.int64, device="cuda")
        ops.cm_update_logits_(cm, dev(lg), dev(t), C)
        np.testing.assert_array_equal(cm.cpu().numpy(), want)
        np.testing.assert_array_equal(oracle.argmax_rows(lg), torch.argmax(torch.as_tensor(lg), dim=1).numpy())


# =============================================================================== K3
def run_binned(ops, x, y, thr):
    x, y, thr = dev(x, torch.float32), dev(y, torch.int64), dev(thr, torch.float32)
    tasks = 1 if x.dim() == 1 else x.shape[0]
    T = thr.numel()
    hist = torch.zeros((tasks, T + 1, 2), dtype=torch.int64, device="cuda")
    ops.binned_hist_update_(hist, x, y, thr)
    tp, fp, fn = ops.binned_counts(hist, T)
    auroc = ops.binned_integrate(tp, fp, fn, 0)
    auprc = ops.binned_integrate(tp, fp, fn, 1)
    torch.cuda.synchronize()
    return [a.cpu().numpy() for a in (hist, tp, fp, fn, auroc, auprc)]


def test_binned_golden(ops, golden):
    for name in golden.cases("binned"):
        c = golden.case(f"binned/{name}")
        hist, tp, fp, fn, auroc, auprc = run_binned(ops, c["x"], c["y"], c["thr"])
        np.testing.assert_array_equal(hist[0], oracle.binned_hist(c["x"], c["y"], c["thr"]))
        np.testing.assert_array_equal(tp[0], c["tp"], err_msg=name)
        np.testing.assert_array_equal(fp[0], c["fp"])
        np.testing.assert_array_equal(fn[0], c["fn"])
        if "auroc" in c:
            close(auroc[0], c["auroc"][0], RTOL64)
        if "auprc" in c:
            close(auprc[0], c["auprc"], RTOL32)
    c = golden.case("binned_multitask/t2")
    hist, tp, fp, fn, auroc, auprc = run_binned(ops, c["x"], c["y"], c["thr"])
    close(auroc, c["auroc"], RTOL64)
    close(auprc, c["auprc"], RTOL32)


@pytest.mark.parametrize("T", [1, 2, 200, 10000, 14000, 30000])
def test_binned_random_vs_oracle(ops, T):
    rs = np.random.default_rng(T)
    n = (1 << 20) + 1
    x = rs.random(n, dtype=np.float32)
    x[:1000] = rs.integers(0, T, 1000) / max(T - 1, 1)  # exactly on thresholds (in fp32 or not)
    x[1000:1010] = [-0.5, 1.5, np.inf, -np.inf, 0.0, 1.0, -0.0, np.nan, 1e-45, 0.99999994]
    y = rs.integers(0, 2, n)
    thr = np.linspace(0, 1, T, dtype=np.float32) if T > 1 else np.array([0.5], dtype=np.float32)
    hist, tp, fp, fn, auroc, auprc = run_binned(ops, x, y, thr)
    # NaN: bucket 0 here, bucket T for searchsorted — drop it from the comparison
    keep = ~np.isnan(x)
    want = oracle.binned_hist(x[keep], y[keep], thr)
    want[0, int(y[1007] != 0)] += 1
    np.testing.assert_array_equal(hist[0], want)
    wtp, wfp, wfn = oracle.binned_counts(want)
    np.testing.assert_array_equal(tp[0], wtp)
    np.testing.assert_array_equal(fp[0], wfp)
    np.testing.assert_array_equal(fn[0], wfn)
    close(auroc[0], oracle.binned_auroc_from_counts(wtp, wfp), 1e-12)
    if T > 1:
        close(auprc[0], oracle.binned_auprc_from_counts(wtp, wfp, wfn), 1e-12)


def test_binned_irregular_thresholds_and_alignment(ops):
    rs = np.random.default_rng(9)
    n = 200003
    x = rs.random(n, dtype=np.float32)
    y = rs.integers(0, 2, n)
    thr = np.sort(np.concatenate([[0.0], np.full(3000, 0.5), 0.5 + rs.random(500) * 1e-4, rs.random(300), [0.25] * 7, [1.0]])).astype(np.float32)
    x[:5000] = rs.choice(thr, 5000)
    want = oracle.binned_hist(x, y, thr)
    hist = run_binned(ops, x, y, thr)[0]
    np.testing.assert_array_equal(hist[0], want)
    # unaligned view -> scalar path
    X, Y, TH = dev(x), dev(y), dev(thr)
    h2 = torch.zeros((1, thr.size + 1, 2), dtype=torch.int64, device="cuda")
    ops.binned_hist_update_(h2, X[1:], Y[1:], TH)
    ops.binned_hist_update_(h2, X[:1], Y[:1], TH)
    np.testing.assert_array_equal(h2.cpu().numpy()[0], want)


def test_binned_multitask_random(ops):
    rs = np.random.default_rng(10)
    K, n, T = 5, 40004, 100
    x = rs.random((K, n), dtype=np.float32)
    y = rs.integers(0, 2, (K, n))
    thr = np.linspace(0, 1, T, dtype=np.float32)
    hist = run_binned(ops, x, y, thr)[0]
    for k in range(K):
        np.testing.assert_array_equal(hist[k], oracle.binned_hist(x[k], y[k], thr))


# =============================================================================== K1: the sort
@pytest.mark.parametrize("n", [0, 1, 5, 511, 8191, 8192, 8193, 100000, (1 << 20) + 77])
def test_sort_u32_keys(ops, n):
    rs = np.random.default_rng(n + 1)
    k = rs.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    out, _ = ops.sort_u32(dev(k.view(np.int32)))
    np.testing.assert_array_equal(out.cpu().numpy().view(np.uint32), np.sort(k))


@pytest.mark.parametrize("n,bits", [(8193, 32), (300001, 32), (300001, 6), ((1 << 21) + 5, 12)])
def test_sort_u32_pairs_stable(ops, n, bits):
    rs = np.random.default_rng(n + bits)
    k = rs.integers(0, 1 << bits, n, dtype=np.uint64).astype(np.uint32)
    if bits == 6:
        k <<= 13  # ties on a middle digit only
    v = np.arange(n, dtype=np.int32)
    ok, ov = ops.sort_u32(dev(k.view(np.int32)), dev(v))
    order = np.argsort(k, kind="stable")
    np.testing.assert_array_equal(ok.cpu().numpy().view(np.uint32), k[order])
    np.testing.assert_array_equal(ov.cpu().numpy(), order.astype(np.int32))


def test_sort_u32_skewed_digits(ops):
    # float-like keys: half of them share the top byte (the exponent) -> long look-back chains
    rs = np.random.default_rng(2)
    n = (1 << 20) + 9
    k = (rs.random(n, dtype=np.float32).view(np.uint32) << 1) | rs.integers(0, 2, n).astype(np.uint32)
    out, _ = ops.sort_u32(dev(k.view(np.int32)))
    np.testing.assert_array_equal(out.cpu().numpy().view(np.uint32), np.sort(k))
    k[:] = 0xDEADBEEF
    out, _ = ops.sort_u32(dev(k.view(np.int32)))
    assert (out.cpu().numpy().view(np.uint32) == 0xDEADBEEF).all()


# =============================================================================== K1: AUROC / AUPRC
def run_rank(ops, x, y, ydtype=torch.int64):
    X, Y = dev(x, torch.float32), dev(y, ydtype)
    a = ops.binary_auroc_packed(X, Y).cpu().numpy()
    p = ops.binary_auprc_packed(X, Y).cpu().numpy()
    return a, p


def test_rank_golden(ops, golden):
    for name in golden.cases("binary"):
        c = golden.case(f"binary/{name}")
        a, p = run_rank(ops, c["x"], c["y"])
        if (np.signbit(c["x"]) & (c["x"] != 0)).any():
            assert int(a[4]) & 2 and int(p[4]) & 2, name  # negative scores are reported, not mis-sorted
            continue
        assert a[4] == 0 and p[4] == 0, name
        close(a[0], c["auroc_f64"], RTOL64)
        close(a[0], c["auroc_fn"], RTOL32)
        close(p[0], c["auprc"], RTOL32)
        assert a[1] == c["y"].sum() and a[2] == (1 - c["y"]).sum()
        if c["y"].sum() == 0:
            assert np.signbit(p[0])
        # exact integer statistic against the oracle
        v, tp, fp = oracle.binary_auroc(c["x"], c["y"])
        if tp * fp > 0:
            assert a[3] == round(2 * v * tp * fp), name


@pytest.mark.parametrize(
    "n,kind",
    [(4097, "uniform"), (8192 * 3, "uniform"), ((1 << 22) + 5, "uniform"), (1 << 21, "ties4096"),
     (1 << 20, "ties3"), (300000, "constant"), (1 << 20, "skew"), (70001, "allpos"), (70001, "allneg")],
)
def test_rank_random_vs_oracle(ops, n, kind):
    rs = np.random.default_rng(n % 1000 + len(kind))
    y = rs.integers(0, 2, n)
    if kind == "uniform":
        x = rs.random(n, dtype=np.float32)
    elif kind == "ties4096":
        x = (rs.integers(0, 4096, n) / 4096).astype(np.float32)
    elif kind == "ties3":
        x = (rs.integers(0, 3, n) / 2).astype(np.float32)
    elif kind == "constant":
        x = np.full(n, 0.7, dtype=np.float32)
    elif kind == "skew":
        x = rs.random(n, dtype=np.float32)
        y = (rs.random(n) < 0.01).astype(np.int64)
    elif kind == "allpos":
        x, y = rs.random(n, dtype=np.float32), np.ones(n, dtype=np.int64)
    else:
        x, y = rs.random(n, dtype=np.float32), np.zeros(n, dtype=np.int64)
    a, p = run_rank(ops, x, y)
    assert a[4] == 0
    v, tp, fp = oracle.binary_auroc(x, y)
    close(a[0], v, RTOL64)
    assert (a[1], a[2]) == (tp, fp)
    close(p[0], oracle.binary_auprc(x, y), RTOL64)
    a2, p2 = run_rank(ops, x, y)
    assert a2[0] == a[0] and a2[3] == a[3] and p2[0] == p[0]  # run-to-run bit identical


def test_rank_target_dtypes_and_flags(ops):
    rs = np.random.default_rng(0)
    n = 50001
    x = rs.random(n, dtype=np.float32)
    y = rs.integers(0, 2, n)
    ref = oracle.binary_auroc(x, y)[0]
    for dt in (torch.int64, torch.int32, torch.uint8, torch.bool, torch.float32, torch.float64):
        a, _ = run_rank(ops, x, y, dt)
        assert a[4] == 0
        close(a[0], ref, 1e-12)
    y2 = y.copy()
    y2[17] = 2
    assert int(run_rank(ops, x, y2)[0][4]) & 1
    assert int(run_rank(ops, x, y.astype(np.float32) * 0.5, torch.float32)[0][4]) & 1
    x2 = x.copy()
    x2[5] = -1e-30
    assert int(run_rank(ops, x2, y)[0][4]) & 2
    x2[5] = -0.0  # signed zero is fine and ties with +0.0
    x2[6] = 0.0
    a, _ = run_rank(ops, x2, y)
    assert a[4] == 0
    close(a[0], oracle.binary_auroc(x2, y)[0], 1e-12)


@pytest.mark.parametrize("n", [3, 8192, 100003, 2_000_001])
def test_rank_constant_digit_passes(ops, n):
    """Score sets whose keys are constant in one or more radix digits (those passes are skipped
    on the device): bf16-valued scores (low 16 key bits = label only), one binade (top byte
    constant), a single value, and a (key, value) sort whose keys have two constant bytes."""
    rs = np.random.default_rng(n)
    y = rs.integers(0, 2, n)
    u = rs.random(n, dtype=np.float32)
    bf16 = (u.view(np.uint32) & np.uint32(0xFFFF0000)).view(np.float32)
    cases = {"bf16": bf16, "binade": (0.5 + 0.5 * u).astype(np.float32), "constant": np.full(n, 0.375, dtype=np.float32),
             "two_values": np.where(u < 0.3, np.float32(0.25), np.float32(0.75))}
    for name, x in cases.items():
        out = ops.binary_auroc_packed(dev(x), dev(y)).cpu().numpy()
        want, tp, fp = oracle.binary_auroc(x, y)
        assert out[4] == 0
        close(out[0], want, RTOL64)
        assert (out[1], out[2]) == (tp, fp), name
        ap = ops.binary_auprc_packed(dev(x), dev(y)).cpu().numpy()
        close(ap[0], oracle.binary_auprc(x, y), RTOL64)
    keys = (rs.integers(0, 256, n).astype(np.uint32) << np.uint32(8)) | np.uint32(0x5A000000) | rs.integers(0, 4, n).astype(np.uint32) << np.uint32(16)
    keys = keys & np.uint32(0xFF03FF00)  # byte 0 constant (0), byte 2 has 4 values, byte 3 constant
    vals = np.arange(n, dtype=np.int32)
    ok, ov = ops.sort_u32(dev(keys.view(np.int32)), dev(vals))
    order = np.argsort(keys, kind="stable")
    np.testing.assert_array_equal(ok.cpu().numpy().view(np.uint32), keys[order])
    np.testing.assert_array_equal(ov.cpu().numpy(), vals[order])  # stable: payload order is fixed


def test_rank_large_closed_form(ops):
    """Size-independent property at a size the sort-based oracle cannot afford: scores on a grid
    of G levels -> AUROC in closed form from the (level, label) histogram."""
    n, G = 1 << 26, 1 << 16
    g = torch.Generator(device="cuda").manual_seed(7)
    lev = torch.randint(0, G, (n,), device="cuda", generator=g)
    y = (torch.rand(n, device="cuda", generator=g) < 0.3).long()
    x = lev.float() / G
    a = ops.binary_auroc_packed(x, y).cpu().numpy()
    neg = torch.bincount(lev[y == 0], minlength=G).cpu().numpy()
    pos = torch.bincount(lev[y == 1], minlength=G).cpu().numpy()
    assert a[4] == 0 and a[1] == pos.sum() and a[2] == neg.sum()
    close(a[0], oracle.auroc_from_level_hist(neg, pos), 1e-9)


def test_full_size_2p30_closed_forms(ops):
    """BASELINE configs 3 and 4 at their full size, N = 2^30 scores on one GPU, through
    size-independent properties: scores on a grid of G levels, so that
      * exact AUROC / AUPRC follow in closed form from the (level, label) histogram,
      * 2U is an exact integer that must match digit for digit,
      * the 10k-threshold binned counts are prefix sums of the same histogram."""
    n, G, T = 1 << 30, 1 << 12, 10_000
    g = torch.Generator(device="cuda").manual_seed(2030)
    lev = torch.randint(0, G, (n,), device="cuda", generator=g)
    y = (torch.rand(n, device="cuda", generator=g) < 0.25).long()
    x = lev.float() / G  # exact in float32
    neg = torch.bincount(lev[y == 0], minlength=G).cpu().numpy().astype(np.int64)
    pos = torch.bincount(lev[y == 1], minlength=G).cpu().numpy().astype(np.int64)
    del lev
    a = ops.binary_auroc_packed(x, y).cpu().numpy()
    assert a[4] == 0 and a[1] == pos.sum() and a[2] == neg.sum()
    below = np.concatenate([[0], np.cumsum(neg)[:-1]])
    u2 = int(np.sum(pos.astype(object) * (2 * below.astype(object) + neg.astype(object))))
    assert (int(a[5]) << 32) | int(a[6]) == u2  # exact integer Mann-Whitney statistic
    close(a[0], oracle.auroc_from_level_hist(neg, pos), 1e-12)
    ap = ops.binary_auprc_packed(x, y).cpu().numpy()
    tp_ge = np.cumsum(pos[::-1])[::-1].astype(np.float64)
    all_ge = np.cumsum((pos + neg)[::-1])[::-1].astype(np.float64)
    want_ap = float(np.sum(pos * (tp_ge / np.maximum(all_ge, 1))) / pos.sum())
    close(ap[0], want_ap, 1e-9)
    thr = torch.linspace(0, 1, T, device="cuda")
    hist = torch.zeros((1, T + 1, 2), dtype=torch.int64, device="cuda")
    ops.binned_hist_update_(hist, x, y, thr)
    tp, fp, fn = ops.binned_counts(hist, T)
    levels = (np.arange(G, dtype=np.float32) / np.float32(G))
    ge = levels[None, :] >= thr.cpu().numpy()[:, None]          # (T, G)
    np.testing.assert_array_equal(tp.cpu().numpy()[0], ge.astype(np.int64) @ pos)
    np.testing.assert_array_equal(fp.cpu().numpy()[0], ge.astype(np.int64) @ neg)
    assert int(hist.sum()) == n


# =============================================================================== K1: general path
def run_general(ops, x, y, w=None, ydtype=None):
    X = dev(x, torch.float32)
    Y = dev(y, ydtype)
    W = None if w is None else dev(w, torch.float64)
    a = ops.binary_auroc_general(X, Y, W).cpu().numpy()
    return a


def test_general_golden(ops, golden):
    for name in golden.cases("binary"):
        c = golden.case(f"binary/{name}")
        a = run_general(ops, c["x"], c["y"])
        close(a[0], c["auroc_f64"], RTOL64)
        assert a[1] == c["y"].sum() and a[2] == (1 - c["y"]).sum()
        p = ops.binary_auprc_general(dev(c["x"], torch.float32), dev(c["y"])).cpu().numpy()
        close(p[0], c["auprc"], RTOL32)
        if c["y"].sum() == 0:
            assert np.signbit(p[0])
    for name in golden.cases("binary_weighted"):
        c = golden.case(f"binary_weighted/{name}")
        a = run_general(ops, c["x"], c["y"], c["w"])
        close(a[0], c["auroc"], RTOL64)


@pytest.mark.parametrize("n,kind", [(2049, "logits"), ((1 << 20) + 3, "logits"), (1 << 19, "ties"), (100000, "soft")])
def test_general_random_vs_oracle(ops, n, kind):
    rs = np.random.default_rng(n % 97 + len(kind))
    w = rs.random(n) * 2
    if kind == "logits":
        x = (rs.standard_normal(n) * 4).astype(np.float32)
        y = rs.integers(0, 2, n).astype(np.float64)
    elif kind == "ties":
        x = (rs.integers(-50, 50, n) / 7).astype(np.float32)
        y = rs.integers(0, 2, n).astype(np.float64)
    else:
        x = (rs.integers(-500, 500, n) / 77).astype(np.float32)
        y = rs.random(n)
    for weight in (None, w):
        a = run_general(ops, x, y, weight, torch.float64)
        v, tp, fp = oracle.binary_auroc(x, y, weight)
        close(a[0], v, RTOL64)
        close(a[1], tp, 1e-9)
        close(a[2], fp, 1e-9)
    if kind != "soft":
        yi = y.astype(np.int64)
        p = ops.binary_auprc_general(dev(x), dev(yi)).cpu().numpy()
        close(p[0], oracle.binary_auprc(x, yi), RTOL64)
        # packed and general agree where both apply
        xp = np.abs(x)
        a1 = ops.binary_auroc_packed(dev(xp), dev(yi)).cpu().numpy()
        a2 = run_general(ops, xp, yi)
        assert a1[4] == 0
        close(a1[0], a2[0], 1e-12)


def test_general_weights_as_payload(ops):
    """Weights that are non-negative float32 values (the usual case: ones, class weights, float32
    tensors cast up) on 0/1 targets travel through the sort as the payload, label in the sign bit;
    anything else is gathered by index.  Both against the oracle, and against each other."""
    rs = np.random.default_rng(21)
    for n in (1, 5, 4097, 300_001):
        x = (rs.standard_normal(n) * 3).astype(np.float32)
        x[: n // 4] = np.round(x[: n // 4])  # ties, and -0.0 / +0.0
        y = rs.integers(0, 2, n)
        w32 = rs.random(n).astype(np.float32).astype(np.float64)
        w32[: n // 10] = 0.0
        a = run_general(ops, x, y, w32, torch.int64)
        v, tp, fp = oracle.binary_auroc(x, y, w32)
        close(a[0], v, RTOL64)
        close(a[1], tp, 1e-12)
        close(a[2], fp, 1e-12)
        # a weight float32 cannot hold, a negative weight, a soft label: each falls back to the gather
        for kind in ("inexact", "negative", "soft"):
            w2, y2 = w32.copy(), y.astype(np.float64)
            if kind == "inexact":
                w2[n // 2] = 1.0 + 2.0 ** -40
            elif kind == "negative":
                w2[n // 2] = -0.5
            else:
                y2[n // 2] = 0.25
            b = run_general(ops, x, y2, w2, torch.float64)
            close(b[0], oracle.binary_auroc(x, y2, w2)[0], RTOL64)


def test_general_batched_tasks_vs_oracle(ops):
    """(tasks, n) rankings in one launch train (grid.y = task): every row against the oracle and
    against the single-task entry point; rows differ in what their payload can carry (float32
    weights ride along, an inexact weight or a soft label makes that ROW gather by index)."""
    rs = np.random.default_rng(77)
    for tasks, n in ((3, 1), (5, 1000), (64, 1 << 14), (2, 300_001), (7, 8192 * 3 + 5)):
        x = (rs.standard_normal((tasks, n)) * 3).astype(np.float32)
        x[:, : n // 3] = np.round(x[:, : n // 3])
        y = rs.integers(0, 2, (tasks, n)).astype(np.float64)
        w = rs.random((tasks, n)).astype(np.float32).astype(np.float64)
        if n > 10:
            w[1, n // 2] = 1.0 + 2.0**-40  # row 1: weights float32 cannot hold
            y[0, n // 3] = 0.25  # row 0: a soft label
            y[tasks - 1] = 1.0  # last row: no negatives -> 0.5
        for weight in (w, None):
            W = None if weight is None else dev(weight, torch.float64)
            got = ops.binary_auroc_general_batched(dev(x), dev(y, torch.float64), W).cpu().numpy()
            assert got.shape == (tasks, 8)
            for k in range(tasks):
                v, tp, fp = oracle.binary_auroc(x[k], y[k], None if weight is None else weight[k])
                close(got[k, 0], v, RTOL64)
                close(got[k, 1], tp, 1e-9)
                close(got[k, 2], fp, 1e-9)
                one = ops.binary_auroc_general(dev(x[k]), dev(y[k], torch.float64), None if weight is None else dev(weight[k], torch.float64)).cpu().numpy()
                close(got[k, 0], one[0], 1e-13)
    # the functional entry point on 2-D input with weights, and the windowed metric on top of it
    from torcheval_b200.metrics import WindowedBinaryAUROC
    from torcheval_b200.metrics.functional.classification.auroc import binary_auroc

    tasks, n = 6, 5000
    x = rs.standard_normal((tasks, n)).astype(np.float32)
    y = rs.integers(0, 2, (tasks, n))
    w = rs.random((tasks, n)).astype(np.float32)
    got = binary_auroc(dev(x), dev(y), num_tasks=tasks, weight=dev(w)).cpu().numpy()
    want = [oracle.binary_auroc(x[k], y[k], w[k].astype(np.float64))[0] for k in range(tasks)]
    np.testing.assert_allclose(got, want, rtol=RTOL64)
    m = WindowedBinaryAUROC(num_tasks=tasks, max_num_samples=3000, device="cuda")
    m.update(dev(x), dev(y), dev(w))
    got = m.compute().cpu().numpy()
    want = [oracle.binary_auroc(x[k, -3000:], y[k, -3000:], w[k, -3000:].astype(np.float64))[0] for k in range(tasks)]
    np.testing.assert_allclose(got, want, rtol=RTOL64)


def test_functional_binary_auroc_dispatch(ops, golden):
    from torcheval_b200.metrics.functional.classification.auroc import binary_auroc

    for name in ("uniform", "logits", "ties64", "all_pos"):
        c = golden.case(f"binary/{name}")
        got = binary_auroc(torch.as_tensor(c["x"]), torch.as_tensor(c["y"]))  # host tensors in, host out
        assert got.device.type == "cpu" and got.dtype == torch.float64 and got.shape == ()
        close(got.item(), c["auroc_f64"], RTOL64)
    c = golden.case("binary_weighted/soft")
    got = binary_auroc(torch.as_tensor(c["x"]).cuda(), torch.as_tensor(c["y"]).cuda(), weight=torch.as_tensor(c["w"]).cuda())
    assert got.is_cuda
    close(got.item(), c["auroc"], RTOL64)
    c = golden.case("binary_multitask/t3")
    got = binary_auroc(torch.as_tensor(c["x"]), torch.as_tensor(c["y"]), num_tasks=3)
    assert got.shape == (3,)
    close(got.numpy(), c["auroc"], RTOL64)


# =============================================================================== distributed pieces
def test_rank_pack_partition_from_keys(ops):
    import cpu_backend

    rs = np.random.default_rng(21)
    for n in (1, 4097, 300001):
        x = rs.random(n, dtype=np.float32)
        x[::17] = 0.5
        if n > 10:
            x[3], x[4] = -0.0, 0.0
        y = rs.integers(0, 2, n)
        keys, info, cnt0 = ops.rank_pack(dev(x), dev(y))
        wk, winfo, _ = cpu_backend.rank_pack(torch.from_numpy(x), torch.from_numpy(y))
        np.testing.assert_array_equal(keys.cpu().numpy(), wk.numpy())
        assert info.tolist() == winfo.tolist() and cnt0.tolist() == [n]
        for sp in ([], [int(0.5 * 2**31) & ~1], sorted(int(v) & ~1 for v in rs.integers(0, 2**31, 7)), [10, 10, 2**32 - 2]):
            wp, wc = cpu_backend.partition_u32(wk, sp)
            k2, _, c2 = ops.rank_pack(dev(x), dev(y), sp)  # bucket sizes straight from the pack pass
            assert c2.tolist() == wc.tolist(), (n, sp)
            sp_dev = torch.tensor(sp, dtype=torch.int64, device="cuda").to(torch.int32)  # bit patterns
            k3, _, c3 = ops.rank_pack(dev(x), dev(y), sp_dev)  # splitters read from device memory
            assert c3.tolist() == wc.tolist() and torch.equal(k3, k2), (n, sp)
            # per-bucket radix-digit histograms computed while packing
            k4, _, c4, h4 = ops.rank_pack(dev(x), dev(y), sp_dev, want_hist=True)
            _, _, _, wh = cpu_backend.rank_pack(torch.from_numpy(x), torch.from_numpy(y), sp, want_hist=True)
            assert torch.equal(k4, k2) and c4.tolist() == wc.tolist()
            np.testing.assert_array_equal(h4.cpu().numpy(), wh.numpy(), err_msg=f"{n} {sp}")
            for known in (None, c2):
                part, counts = ops.partition_u32(k2, sp, known)
                assert counts.tolist() == wc.tolist(), (n, sp)
                got, want_p, off = part.cpu().numpy().view(np.uint32), wp.numpy().view(np.uint32), 0
                for c in wc.tolist():  # same multiset per bucket; the order inside a bucket is free
                    np.testing.assert_array_equal(np.sort(got[off : off + c]), np.sort(want_p[off : off + c]))
                    off += c
        out = ops.binary_auroc_from_keys(keys).cpu().numpy()
        want = cpu_backend.binary_auroc_from_keys(wk).numpy()
        close(out[0], want[0], 1e-12)
        assert list(out[1:4]) == list(want[1:4]) and list(out[5:8]) == list(want[5:8])
        # the same with the digit histograms handed in instead of recomputed from the keys
        _, _, _, h0 = ops.rank_pack(dev(x), dev(y), [], want_hist=True)
        out_h = ops.binary_auroc_from_keys(keys, h0[0].contiguous()).cpu().numpy()
        np.testing.assert_array_equal(out_h, out)
        a1, st1 = ops.binary_auprc_from_keys_begin(keys, h0[0].contiguous())
        a0, st0 = ops.binary_auprc_from_keys_begin(keys)
        assert torch.equal(a1, a0)
        assert torch.equal(ops.binary_auprc_from_keys_finish(st1, 0, 0), ops.binary_auprc_from_keys_finish(st0, 0, 0))
    _, info, _ = ops.rank_pack(dev(np.array([0.5, -1.0], dtype=np.float32)), dev(np.array([0, 1])))
    assert info.tolist() == [1, 2]


def test_select_splitters_and_combine_stats(ops):
    """The two one-CTA helpers of the sample sort against their Python restatements."""
    from cpu_backend import combine_rank_stats

    rs = np.random.default_rng(77)
    for world, s, empty in [(2, 4096, ()), (8, 4096, ()), (3, 1000, (1,)), (16, 2048, (0, 5)), (4, 7, ()), (2, 1, (0, 1))]:
        bits = rs.random((world, s), dtype=np.float32).view(np.int32).copy()
        bits[:, ::5] = bits[:, :1]  # duplicates
        for r in empty:
            bits[r] = -1
        sp32, sp64 = ops.select_splitters(dev(bits), world)
        pool = np.sort(bits[bits >= 0].astype(np.int64))
        m = pool.size
        want = [int((pool[min(m - 1, (i * m) // world)] << 1) & 0xFFFFFFFE) if m else 0 for i in range(1, world)]
        assert sp64.tolist() == want, (world, s)
        assert [v & 0xFFFFFFFF for v in sp32.tolist()] == want
    for world in (1, 2, 8):
        rows = np.zeros((world, 8))
        stats = []
        for r in range(world):
            p_, n_ = int(rs.integers(0, 2**28)), int(rs.integers(0, 2**28))
            u2 = int(rs.integers(0, 2 * p_ * n_ + 1))
            rows[r, 1], rows[r, 2], rows[r, 5], rows[r, 6] = p_, n_, u2 >> 32, u2 & 0xFFFFFFFF
            stats.append((p_, n_, u2))
        out = ops.combine_auroc_stats(dev(rows), 1234).cpu().numpy()
        auroc, P, N_, u2 = combine_rank_stats(stats)
        close(out[0], auroc, 1e-15)
        assert (int(out[1]), int(out[2]), out[5]) == (P, N_, 1234.0) and out[3] == float(u2)


@pytest.mark.parametrize("n", [1, 37, 8192, 8193, 123457, 2_000_003])
def test_partition_scatter_to_separate_buffers(ops, n):
    """te_partition_scatter_u32 with the buckets going to separate buffers at every 4-byte
    misalignment (what the peer inboxes look like to the kernel), against the stable CPU partition."""
    import cpu_backend

    rs = np.random.default_rng(n)
    x = rs.random(n, dtype=np.float32)
    y = rs.integers(0, 2, n)
    keys, _, _ = ops.rank_pack(dev(x), dev(y))
    wk = keys.cpu()
    for sp in ([], [int(0.5 * 2**31) & ~1], sorted(int(v) & ~1 for v in rs.integers(0, 2**31, 7)), [4, 4, 2**32 - 2],
               sorted(int(v) & ~1 for v in rs.integers(0, 2**31, 15))):
        wp, wc = cpu_backend.partition_u32(wk, sp)
        wc = wc.tolist()
        _, _, counts = ops.rank_pack(dev(x), dev(y), sp)
        for slot in (0, 1, 2, 3, 5):
            bufs = [torch.full((c + 16,), -1, dtype=torch.int32, device="cuda") for c in wc]
            dest, excl = [], 0
            for b, c in enumerate(wc):
                dest.append(bufs[b].data_ptr() + 4 * (slot + (b % 3) - excl))
                excl += c
            ops.partition_scatter_u32(keys, sp, counts, dest)
            torch.cuda.synchronize()
            off = 0
            for b, c in enumerate(wc):
                o = slot + (b % 3)
                got = bufs[b].cpu().numpy()
                np.testing.assert_array_equal(np.sort(got[o : o + c]), np.sort(wp.numpy()[off : off + c]), err_msg=f"{n} {sp} {slot} {b}")
                assert (got[:o] == -1).all() and (got[o + c :] == -1).all()  # nothing outside the run is touched
                off += c


@pytest.mark.parametrize("n", [1, 5, 4097, 300001, 3_000_001])
def test_auprc_from_keys_two_phase(ops, n):
    """te_binary_auprc_from_keys_begin/_finish: whole set == te_binary_auprc; and split at a score
    boundary into a low and a high part, each finished with what lies above it, the two sums add
    up to the whole (the decomposition the ranks use)."""
    import cpu_backend

    rs = np.random.default_rng(n)
    x = (rs.integers(0, 4000, n) / 4000).astype(np.float32) if n % 2 else rs.random(n, dtype=np.float32)
    y = rs.integers(0, 2, n)
    keys, _, _ = ops.rank_pack(dev(x), dev(y))
    whole = ops.binary_auprc_packed(dev(x), dev(y)).cpu().numpy()
    loc, st = ops.binary_auprc_from_keys_begin(keys)
    loc = loc.cpu().numpy()
    wa = cpu_backend.binary_auroc_from_keys(keys.cpu()).numpy()
    close(loc[0], wa[0], 1e-12)  # phase 1 also reports the AUROC of the keys
    assert list(loc[1:3]) == list(whole[1:3])
    fin = ops.binary_auprc_from_keys_finish(st, 0, 0).cpu().numpy()
    P = whole[1]
    close(fin[3] / P if P else -0.0, oracle.binary_auprc(x, y), RTOL64)
    assert fin[3] == whole[3]  # same kernel, same order of summation
    if n < 5:
        return
    cutv = np.float32(np.median(x))
    lo, hi = x < cutv, x >= cutv
    klo, _, _ = ops.rank_pack(dev(x[lo]), dev(y[lo]))
    khi, _, _ = ops.rank_pack(dev(x[hi]), dev(y[hi]))
    _, slo = ops.binary_auprc_from_keys_begin(klo)
    _, shi = ops.binary_auprc_from_keys_begin(khi)
    f_hi = ops.binary_auprc_from_keys_finish(shi, 0, 0).cpu().numpy()
    f_lo = ops.binary_auprc_from_keys_finish(slo, int(y[hi].sum()), int(hi.sum())).cpu().numpy()
    want_lo = cpu_backend.binary_auprc_from_keys_finish(np.sort(klo.cpu().numpy().view(np.uint32)), int(y[hi].sum()), int(hi.sum()))
    close(f_lo[3], float(want_lo[3]), 1e-12)
    above = torch.tensor([float(y[hi].sum()), float(hi.sum())], dtype=torch.float64, device="cuda")
    assert ops.binary_auprc_from_keys_finish(slo, above).cpu().numpy()[3] == f_lo[3]  # same numbers from device memory
    close(f_lo[3] + f_hi[3], whole[3], 1e-12)


@pytest.mark.parametrize("n", [1, 2, 511, 513, 40000, 300001, 3_000_001])
def test_pr_curve_vs_cpu(ops, n):
    """te_binary_pr_curve_begin/_finish against the numpy restatement of
    precision_recall_curve.py:208-231 (tests/cpu_backend.py), thresholds bit-exact."""
    import cpu_backend

    rs = np.random.default_rng(n)
    for kind in ("ties", "distinct", "no_pos", "all_pos", "constant"):
        x = rs.random(n, dtype=np.float32)
        y = rs.integers(0, 2, n)
        if kind == "ties":
            x = (rs.integers(0, 300, n) / 300).astype(np.float32)
        elif kind == "no_pos":
            y[:] = 0
        elif kind == "all_pos":
            y[:] = 1
        elif kind == "constant":
            x[:] = 0.25
        (p, r, t), flags = ops.binary_pr_curve(dev(x), dev(y))
        (wp, wr, wt), _ = cpu_backend.binary_pr_curve(torch.from_numpy(x), torch.from_numpy(y))
        assert flags == 0 and t.shape == wt.shape, (n, kind)
        np.testing.assert_array_equal(t.cpu().numpy(), wt.numpy())
        np.testing.assert_allclose(p.cpu().numpy(), wp.numpy(), rtol=2e-7, atol=0)
        np.testing.assert_allclose(r.cpu().numpy(), wr.numpy(), rtol=2e-7, atol=0)
    curve, flags = ops.binary_pr_curve(dev(np.array([0.5, -1.0], dtype=np.float32)), dev(np.array([0, 1])))
    assert curve is None and flags == 2


# =============================================================================== K2: one-vs-rest
@pytest.mark.parametrize("n,C", [(1, 3), (257, 17), (5000, 8), (40001, 13), (3000, 1000)])
def test_ovr_rank_vs_oracle(ops, n, C):
    rs = np.random.default_rng(n + C)
    x = (rs.integers(0, 200, (n, C)) / 200).astype(np.float32) if C < 100 else rs.random((n, C), dtype=np.float32)
    y = rs.integers(0, C, n)
    if C > 3:
        y[y == 2] = 1  # class 2 never occurs
    a = ops.ovr_rank(dev(x), dev(y), False, 0).cpu().numpy()
    p = ops.ovr_rank(dev(x), dev(y), False, 1).cpu().numpy()
    assert (a[:, 4] == 0).all()
    close(a[:, 0], oracle.multiclass_auroc(x, y), RTOL64)
    close(p[:, 0], oracle.multiclass_auprc(x, y), RTOL64)
    np.testing.assert_array_equal(a[:, 1], np.bincount(y, minlength=C))
    ml = rs.integers(0, 2, (n, C))
    if C > 3:
        ml[:, 1] = 0
    q = ops.ovr_rank(dev(x), dev(ml), True, 1).cpu().numpy()
    close(q[:, 0], oracle.multilabel_auprc(x, ml), RTOL64)
    # a strided view (row stride > C) and a column with negative scores
    wide = np.concatenate([x, -np.ones((n, 3), dtype=np.float32)], axis=1)
    a2 = ops.ovr_rank(dev(wide)[:, :C], dev(y), False, 0).cpu().numpy()
    np.testing.assert_array_equal(a2[:, 0], a[:, 0])
    x2 = x.copy()
    x2[:, 0] -= 2.0
    a3 = ops.ovr_rank(dev(x2), dev(y), False, 0).cpu().numpy()
    assert (int(a3[0, 4]) & 2) and (a3[1:, 4] == 0).all()  # (the counting-stage path also reports |score| >= 2)
    np.testing.assert_array_equal(a3[1:, 0], a[1:, 0])


def test_ovr_counting_path_equals_sort_path(ops):
    """te_ovr_rank has two structures for AUROC (four-pass sort; counting-stage path with the classes
    of a batch in grid.y, taken from TE_B200_OVR_LEAF_MIN_N samples on): same integers from both.
    The switch is read once per process, so the comparison runs in a child process."""
    import os
    import subprocess
    import sys

    code = (
        "import numpy as np, torch, sys; sys.path.insert(0, %r)\n"
        "from torcheval_b200 import ops\n"
        "g = torch.Generator(device='cuda').manual_seed(3)\n"
        "for n, C in ((70001, 19), (1 << 19, 9)):\n"
        "    x = torch.rand((n, C), device='cuda', generator=g); x[:, 1] = torch.round(x[:, 1] * 50) / 50\n"
        "    y = torch.randint(0, C, (n,), device='cuda', generator=g)\n"
        "    o = ops.ovr_rank(x, y, False, 0).cpu().numpy()\n"
        "    print(n, C, ' '.join(f'{int(v)}' for v in np.concatenate([o[:, 1], o[:, 2], o[:, 3], o[:, 4]])))\n"
    ) % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for min_n in ("1", "99999999999"):
        env = dict(os.environ, TE_B200_OVR_LEAF_MIN_N=min_n)
        r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(r.stdout)
    assert outs[0] == outs[1] and len(outs[0].splitlines()) == 2


def test_ovr_config5_classes_closed_form(ops):
    """BASELINE configs[4] shape at C = 1000 with N = 2^20 rows (1.05e9 scores): scores on a
    256-level grid, so per class the exact integer 2U, AUROC and the average precision follow in
    closed form from that class's (level, label) histogram — ties on every level."""
    n, C, G = 1 << 20, 1000, 256
    g = torch.Generator(device="cuda").manual_seed(55)
    lev = torch.randint(0, G, (n, C), device="cuda", generator=g, dtype=torch.int16)
    y = torch.randint(0, C, (n,), device="cuda", generator=g)
    y[y == 7] = 8  # class 7 never occurs: AUROC 0.5, AUPRC -0.0
    x = lev.float() / G
    # (class, level) histograms of all rows and of the positive rows
    all_h = torch.zeros(C * G, dtype=torch.int64, device="cuda")
    cls = torch.arange(C, device="cuda", dtype=torch.int64) * G
    for r0 in range(0, n, 1 << 16):
        idx = (lev[r0 : r0 + (1 << 16)].long() + cls).reshape(-1)
        all_h += torch.bincount(idx, minlength=C * G)
    pos_h = torch.bincount(y * G + lev[torch.arange(n, device="cuda"), y].long(), minlength=C * G)
    allh = all_h.reshape(C, G).cpu().numpy()
    pos = pos_h.reshape(C, G).cpu().numpy()
    neg = allh - pos
    del lev, all_h, pos_h
    below = np.cumsum(neg, axis=1) - neg
    u2 = (pos * (2 * below + neg)).sum(axis=1)  # < 2^53: exact in the float64 the kernel reports
    P, N = pos.sum(axis=1), neg.sum(axis=1)
    a = ops.ovr_rank(x, y, False, 0).cpu().numpy()
    assert (a[:, 4] == 0).all()
    np.testing.assert_array_equal(a[:, 1], P)
    np.testing.assert_array_equal(a[:, 2], N)
    np.testing.assert_array_equal(a[:, 3], u2.astype(np.float64))
    want = np.where(P * N == 0, 0.5, u2 / np.maximum(2.0 * P * N, 1))
    np.testing.assert_allclose(a[:, 0], want, rtol=1e-12)
    assert a[7, 0] == 0.5
    p = ops.ovr_rank(x, y, False, 1).cpu().numpy()
    tp_ge = np.cumsum(pos[:, ::-1], axis=1)[:, ::-1].astype(np.float64)
    all_ge = np.cumsum(allh[:, ::-1], axis=1)[:, ::-1].astype(np.float64)
    ap = (pos * (tp_ge / np.maximum(all_ge, 1))).sum(axis=1) / np.maximum(P, 1)
    keep = P > 0
    np.testing.assert_allclose(p[keep, 0], ap[keep], rtol=1e-9)
    assert p[7, 0] == 0 and np.signbit(p[7, 0])
    # multilabel targets (n, C) int64 at C = 1000 (2^17 rows: 1 GB of targets)
    m = 1 << 17
    ml = (torch.rand((m, C), device="cuda", generator=g) < 0.1).long()
    xm = x[:m]
    q = ops.ovr_rank(xm, ml, True, 1).cpu().numpy()
    levm = (xm * G).long()
    pos_m = torch.zeros(C * G, dtype=torch.int64, device="cuda")
    all_m = torch.zeros(C * G, dtype=torch.int64, device="cuda")
    for r0 in range(0, m, 1 << 15):
        idx = (levm[r0 : r0 + (1 << 15)] + cls).reshape(-1)
        all_m += torch.bincount(idx, minlength=C * G)
        pos_m += torch.bincount(idx, weights=ml[r0 : r0 + (1 << 15)].reshape(-1).double(), minlength=C * G).long()
    posm, allm = pos_m.reshape(C, G).cpu().numpy(), all_m.reshape(C, G).cpu().numpy()
    tp_ge = np.cumsum(posm[:, ::-1], axis=1)[:, ::-1].astype(np.float64)
    all_ge = np.cumsum(allm[:, ::-1], axis=1)[:, ::-1].astype(np.float64)
    apm = (posm * (tp_ge / np.maximum(all_ge, 1))).sum(axis=1) / posm.sum(axis=1)
    np.testing.assert_allclose(q[:, 0], apm, rtol=1e-9)
    np.testing.assert_array_equal(q[:, 1], posm.sum(axis=1))


def test_functional_multiclass_with_negative_columns(ops):
    from torcheval_b200.metrics import functional as F

    rs = np.random.default_rng(3)
    n, C = 3001, 6
    x = rs.standard_normal((n, C)).astype(np.float32)  # logits: every column flagged -> general path
    y = rs.integers(0, C, n)
    got = F.multiclass_auroc(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda(), num_classes=C, average=None)
    close(got.cpu().numpy(), oracle.multiclass_auroc(x, y), RTOL32)
    got = F.multiclass_auprc(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda(), average=None)
    close(got.cpu().numpy(), oracle.multiclass_auprc(x, y), RTOL32)


# =============================================================================== edge-size sweep
def test_edge_size_sweep_all_kernels(ops):
    """Sizes around every tile / step / vector boundary of the kernels (0, 1, 3, 4, 5, 511..513,
    4095..4097, 8191..8193, 16383..16385) with heavy ties, through every kernel family."""
    import cpu_backend

    rs = np.random.default_rng(123)
    sizes = [0, 1, 2, 3, 4, 5, 31, 32, 33, 511, 512, 513, 1023, 1025, 4095, 4096, 4097, 8191, 8192, 8193, 16383, 16385, 24577]
    thr = np.linspace(0, 1, 33, dtype=np.float32)
    for n in sizes:
        x = (rs.integers(0, 64, n) / 64).astype(np.float32)
        y = rs.integers(0, 2, n)
        a = ops.binary_auroc_packed(dev(x), dev(y)).cpu().numpy()
        want, tp, fp = oracle.binary_auroc(x, y) if n else (0.5, 0.0, 0.0)
        assert a[4] == 0 and (a[1], a[2]) == (tp, fp), n
        close(a[0], want, RTOL64)
        if n:
            p = ops.binary_auprc_packed(dev(x), dev(y)).cpu().numpy()
            close(p[0], oracle.binary_auprc(x, y), RTOL64)
            (pr, rc, th), flags = ops.binary_pr_curve(dev(x), dev(y))
            (wp, wr, wt), _ = cpu_backend.binary_pr_curve(torch.from_numpy(x), torch.from_numpy(y))
            np.testing.assert_array_equal(th.cpu().numpy(), wt.numpy())
            np.testing.assert_allclose(pr.cpu().numpy(), wp.numpy(), rtol=2e-7)
            np.testing.assert_allclose(rc.cpu().numpy(), wr.numpy(), rtol=2e-7)
            g = ops.binary_auroc_general(dev(x - 0.5), dev(y), None).cpu().numpy()
            close(g[0], want, RTOL64)
        for C in (3, 1000):
            pc, tc = rs.integers(0, C, n), rs.integers(0, C, n)
            wantcm, _, _ = oracle.confusion_matrix(pc, tc, C)
            for v in (0, 6):
                cm, _ = run_cm(ops, dev(pc), dev(tc), C, v)
                np.testing.assert_array_equal(cm, wantcm, err_msg=f"n={n} C={C} v={v}")
        hist = torch.zeros((1, thr.size + 1, 2), dtype=torch.int64, device="cuda")
        ops.binned_hist_update_(hist, dev(x), dev(y), dev(thr))
        np.testing.assert_array_equal(hist.cpu().numpy()[0], oracle.binned_hist(x, y, thr))
        keys = rs.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
        ok, _ = ops.sort_u32(dev(keys.view(np.int32)))
        np.testing.assert_array_equal(ok.cpu().numpy().view(np.uint32), np.sort(keys))
