"""te_cm_update_noalloc: the confusion-matrix update for callers that may not allocate device
memory where they call it (they have no scratch buffer to hand over), so the library's own
static 4 MiB bin array stands in for the scratch of te_cm_update.

Collected last on purpose (file name): a run that stops at the first failure has by then run
every other test.  The check is the strictest one available — counts must be identical, bit for
bit, to what the ordinary entry point te_cm_update produces — and it is made on two calls in a
row with different sizes, because the static array is shared and keeps the previous call's
counts: the second call only passes if the entry point really clears what it uses."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _reference_counts(lib, pred, target, C, variant):
    """The existing entry point, fed a caller-allocated scratch buffer."""
    cm = torch.zeros((C, C), dtype=torch.int64, device="cuda")
    rng = torch.empty(4, dtype=torch.int64, device="cuda")
    nbytes = max(int(lib.te_cm_scratch_bytes(C, variant)), 256)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.te_cm_update(_ptr(pred), _ptr(target), pred.numel(), C, _ptr(cm), _ptr(rng), _ptr(scratch), scratch.numel(), variant, st)
    assert rc == 0
    torch.cuda.synchronize()
    return cm.cpu().numpy(), rng.cpu().numpy()


def _noalloc_counts(lib, pred, target, C, cm=None):
    if cm is None:
        cm = torch.zeros((C, C), dtype=torch.int64, device="cuda")
    rng = torch.empty(4, dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.te_cm_update_noalloc(_ptr(pred), _ptr(target), pred.numel(), C, _ptr(cm), _ptr(rng), st)
    assert rc == 0
    torch.cuda.synchronize()
    return cm.cpu().numpy(), rng.cpu().numpy()


def test_noalloc_matches_te_cm_update_on_two_calls_in_a_row():
    assert torch.cuda.is_available()
    from torcheval_b200 import _native

    lib = _native.load()
    g = torch.Generator(device="cuda").manual_seed(2026)
    # call 1: the full 1024 x 1024 static array is used and left dirty
    C1, n1 = 1024, (1 << 22) + 3
    p1 = torch.randint(0, C1, (n1,), device="cuda", generator=g)
    t1 = torch.randint(0, C1, (n1,), device="cuda", generator=g)
    got1, rng1 = _noalloc_counts(lib, p1, t1, C1)
    for variant in (_native.TE_CM_GLOBAL32, _native.TE_CM_AUTO):
        want1, wrng1 = _reference_counts(lib, p1, t1, C1, variant)
        np.testing.assert_array_equal(got1, want1)  # integer counts: no tolerance
        np.testing.assert_array_equal(rng1, wrng1)
    assert got1.sum() == n1
    # ... and against the CPU oracle itself, not only against the entry point it mirrors
    import oracle

    want_cpu, bad, _ = oracle.confusion_matrix(p1.cpu().numpy(), t1.cpu().numpy(), C1)
    assert not bad
    np.testing.assert_array_equal(got1, want_cpu)

    # call 2, right after: fewer classes, out-of-range labels, an unaligned view, accumulation
    # into a matrix that already holds counts
    C2, n2 = 37, 300_001
    p2 = torch.randint(0, C2, (n2 + 1,), device="cuda", generator=g)[1:]
    t2 = torch.randint(0, C2, (n2 + 1,), device="cuda", generator=g)[1:]
    p2[5], t2[9] = 99, -4
    cm2 = torch.full((C2, C2), 7, dtype=torch.int64, device="cuda")
    got2, rng2 = _noalloc_counts(lib, p2, t2, C2, cm2)
    want2, wrng2 = _reference_counts(lib, p2.clone(), t2.clone(), C2, _native.TE_CM_GLOBAL32)
    np.testing.assert_array_equal(got2, want2 + 7)
    keep = ((p2 >= 0) & (p2 < C2) & (t2 >= 0) & (t2 < C2)).cpu().numpy()  # out-of-range samples are not counted
    want2_cpu, _, _ = oracle.confusion_matrix(p2.cpu().numpy()[keep], t2.cpu().numpy()[keep], C2)
    np.testing.assert_array_equal(got2, want2_cpu + 7)
    np.testing.assert_array_equal(rng2, wrng2)
    assert list(rng2) == [99, 99, -4, -4] and got2.sum() == 7 * C2 * C2 + n2 - 2

    # call 3: the first inputs again — still identical
    got3, _ = _noalloc_counts(lib, p1, t1, C1)
    np.testing.assert_array_equal(got3, got1)


def test_noalloc_refuses_sizes_beyond_its_static_scratch():
    from torcheval_b200 import _native

    lib = _native.load()
    x = torch.zeros(4, dtype=torch.int64, device="cuda")
    cm = torch.zeros(4, dtype=torch.int64, device="cuda")  # never written: the call is refused
    rng = torch.empty(4, dtype=torch.int64, device="cuda")
    rc = lib.te_cm_update_noalloc(_ptr(x), _ptr(x), 4, 1025, _ptr(cm), _ptr(rng), None)
    assert rc == -4
    assert b"static scratch" in lib.te_status_string(rc)
    torch.cuda.synchronize()
    assert int(cm.sum()) == 0
