"""N > 1 on CPU: world_size-2 (and 3) gloo groups exercising the sample-sort protocol and the
toolkit sync paths (see tests/_dist_worker.py).  The GPU kernels are stood in by the oracle; the
real-NCCL run of the same protocol is tests/test_distributed_gpu.py (needs >= 2 GPUs)."""
import json
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def run_world(world: int, tmp_path, port: int):
    out = tmp_path / "res"
    env = dict(os.environ, TE_TEST_OUT=str(out), OMP_NUM_THREADS="2")
    cmd = [
        sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
        "--master-addr", "127.0.0.1", "--master-port", str(port), str(ROOT / "tests" / "_dist_worker.py"),
    ]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return [json.loads(Path(f"{out}.{k}").read_text()) for k in range(world)]


@pytest.mark.parametrize("world,port", [(2, 29611), (3, 29613)])
def test_sample_sort_and_toolkit_sync(world, port, tmp_path):
    results = run_world(world, tmp_path, port)
    for k in range(1, world):
        assert results[k] == results[0], "every rank must return the same values"
    res = results[0]
    for key, val in res.items():
        if key == "calls":
            assert val[0] >= 12 and val[1] >= 8, val  # the sharded paths ran, not the gather fallback
        elif key.startswith("ovr/"):
            np.testing.assert_allclose(val[0], val[1], rtol=2e-6, err_msg=key)
        elif key.startswith("sharded/") or key.startswith("sharded_ap/"):
            got, want, P, tp, N, fp = val
            np.testing.assert_allclose(got, want, rtol=1e-12, err_msg=key)
            assert (P, N) == (tp, fp), key
        elif key in ("coll2/paths", "sync/over_cap_paths"):
            assert val[0] == val[1], (key, val)
        elif key in ("sync/cm_equal", "sync/cm_input_unchanged", "sync/empty_rank_ok", "coll2/cm_equal"):
            assert val is True, (key, val)
        elif key == "generic/toy_meta":
            assert val[0] == val[1], val
        elif key == "sd/cm_total":
            assert val[0] == val[1]
        else:
            np.testing.assert_allclose(val[0], val[1], rtol=1e-6, err_msg=key)


def test_combine_rank_stats_and_splitters():
    from cpu_backend import choose_splitters, combine_rank_stats

    # two ranks: low range has 2 neg + 1 pos (pos beats 1 neg, ties none), high range 1 neg + 2 pos
    auroc, P, N, u2 = combine_rank_stats([(1, 2, 2), (2, 1, 2)])
    assert (P, N) == (3, 3)
    assert u2 == 2 + 2 + 2 * 2 * 2  # each high positive also beats both low negatives
    assert auroc == u2 / (2 * 9)
    assert combine_rank_stats([(0, 5, 0), (0, 3, 0)])[0] == 0.5
    import torch
    from cpu_backend import combine_rank_stats_tensor

    rs = np.random.default_rng(3)
    for _ in range(20):
        w = int(rs.integers(1, 9))
        p_, n_ = rs.integers(0, 2**28, w), rs.integers(0, 2**28, w)
        u2 = [int(rs.integers(0, 2 * int(a) * int(b) + 1)) for a, b in zip(p_, n_)]
        want = combine_rank_stats([(int(a), int(b), c) for a, b, c in zip(p_, n_, u2)])
        st = torch.tensor([[int(a), int(b), c >> 32, c & 0xFFFFFFFF] for a, b, c in zip(p_, n_, u2)], dtype=torch.int64)
        got = combine_rank_stats_tensor(st)
        assert (int(got[1]), int(got[2]), int(got[3])) == want[1:]
        np.testing.assert_allclose(float(got[0]), want[0], rtol=1e-15)
    assert float(combine_rank_stats_tensor(torch.tensor([[0, 5, 0, 0], [0, 3, 0, 0]]))[0]) == 0.5
    sp = choose_splitters(sorted([5, 9, 9, 9, 9, 13, 21, 40]), 4)
    assert sp == sorted(sp) and all(v % 2 == 0 for v in sp) and len(sp) == 3
    assert choose_splitters([], 2) == [0]
