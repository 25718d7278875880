"""GPU-side tuning runs (under gpurun): per-variant timings that decide defaults.
    python tools/tune.py cm        confusion-matrix variants at C=1000 (and a few other C)
    python tools/tune.py rank      per-kernel breakdown of the exact AUROC path at several N
    python tools/tune.py binned    binned histogram at T=10000
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from torcheval_b200 import ops  # noqa: E402

PEAK = 6548.2


def timeit(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


def kernel_times(fn):
    """{kernel: (total ms, launches)} of one call, from the library's own CUDA-event profile."""
    import ctypes

    from torcheval_b200 import _native

    lib = _native.load()
    lib.te_profile_enable(1)
    fn()
    torch.cuda.synchronize()
    names = ctypes.create_string_buffer(8192)
    ms = (ctypes.c_double * 64)()
    cnt = (ctypes.c_int64 * 64)()
    k = lib.te_profile_read(names, 8192, ctypes.cast(ms, ctypes.c_void_p), ctypes.cast(cnt, ctypes.c_void_p), 64)
    lib.te_profile_enable(0)
    out = {}
    for i, nm in enumerate(names.value.decode().split("\n")[:k]):
        out[nm] = (ms[i], cnt[i])
    return out


def cm():
    for C, logn in [(1000, 28), (1000, 24), (500, 26), (1500, 26), (100, 26), (10, 26), (2, 26), (32, 26)]:
        n = 1 << logn
        p = torch.randint(0, C, (n,), device="cuda")
        t = torch.randint(0, C, (n,), device="cuda")
        cmx = torch.zeros((C, C), dtype=torch.int64, device="cuda")
        for v, name in [(1, "global64"), (2, "global32"), (3, "smem_repl"), (4, "smem"), (5, "cluster"), (6, "bucketed")]:
            if v == 3 and C * C * 128 > 200 * 1024:
                continue
            if v == 4 and C * C * 4 > 200 * 1024:
                continue
            if v == 5 and C < 300:
                continue
            if v == 6 and (C > 1240 or C < 100):
                continue
            try:
                ms = timeit(lambda: ops.cm_update_(cmx, p, t, C, v))
            except Exception as e:  # noqa: BLE001
                print(f"C={C} n=2^{logn} {name}: {e}")
                continue
            gbs = 16 * n / ms / 1e6
            print(f"C={C:5d} n=2^{logn} {name:10s} {ms:8.3f} ms  {gbs:7.1f} GB/s  {gbs / PEAK:5.1%}  {n / ms / 1e6:6.1f} G/s", flush=True)
            if v == 6:
                for kname, (kms, kn) in kernel_times(lambda: ops.cm_update_(cmx, p, t, C, v)).items():
                    print(f"        {kname:12s} {kms / max(kn, 1):8.3f} ms")
        del p, t


def rank():
    import ctypes

    from torcheval_b200 import _native

    lib = _native.load()
    for logn in (20, 24, 27, 30):
        n = 1 << logn
        x = torch.rand(n, device="cuda")
        y = torch.randint(0, 2, (n,), device="cuda")
        ops.binary_auroc_packed(x, y)
        torch.cuda.synchronize()
        lib.te_profile_enable(1)
        reps = 3
        ms = timeit(lambda: ops.binary_auroc_packed(x, y), reps=reps)
        names = ctypes.create_string_buffer(8192)
        tot = (ctypes.c_double * 64)()
        cnt = (ctypes.c_int64 * 64)()
        k = lib.te_profile_read(names, 8192, ctypes.cast(tot, ctypes.c_void_p), ctypes.cast(cnt, ctypes.c_void_p), 64)
        lib.te_profile_enable(0)
        print(f"n=2^{logn}: {ms:.3f} ms/step  {n / ms / 1e6:.2f} G scores/s  step HBM {52 * n / ms / 1e6:.0f} GB/s ({52 * n / ms / 1e6 / PEAK:.1%})")
        per = {"rank_ingest_packed": 16, "onesweep_pass_keys": 8, "rank_final_auroc": 4}
        for i, nm in enumerate(names.value.decode().split("\n")[:k]):
            avg = tot[i] / cnt[i]
            extra = ""
            if nm in per:
                g = per[nm] * n / avg / 1e6
                extra = f"  {g:7.0f} GB/s ({g / PEAK:.1%})"
            print(f"    {nm:22s} {avg:9.4f} ms x{cnt[i] / (reps + 1):.0f}{extra}")
        del x, y
    # general path and AUPRC for reference
    n = 1 << 27
    x = torch.randn(n, device="cuda")
    y = torch.randint(0, 2, (n,), device="cuda")
    print(f"general auroc 2^27: {timeit(lambda: ops.binary_auroc_general(x, y)):.3f} ms")
    print(f"general auprc 2^27: {timeit(lambda: ops.binary_auprc_general(x, y)):.3f} ms")
    x = x.abs()
    print(f"packed  auprc 2^27: {timeit(lambda: ops.binary_auprc_packed(x, y)):.3f} ms")


def binned():
    for logn, T in [(30, 10000), (27, 10000), (27, 200), (27, 5)]:
        n = 1 << logn
        x = torch.rand(n, device="cuda")
        y = torch.randint(0, 2, (n,), device="cuda")
        thr = torch.linspace(0, 1, T, device="cuda")
        hist = torch.zeros((1, T + 1, 2), dtype=torch.int64, device="cuda")
        ms = timeit(lambda: ops.binned_hist_update_(hist, x, y, thr))
        gbs = 12 * n / ms / 1e6
        print(f"binned n=2^{logn} T={T}: {ms:.3f} ms  {gbs:.0f} GB/s ({gbs / PEAK:.1%})  {n / ms / 1e6:.1f} G scores/s", flush=True)
        del x, y


if __name__ == "__main__":
    for a in sys.argv[1:]:
        {"cm": cm, "rank": rank, "binned": binned}[a]()
