"""The distributed paths on real GPUs over NCCL (skipped with fewer than 2 devices)."""
import json
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


@pytest.mark.parametrize(
    "no_peer,leaf_min_n,no_slabs,slack",
    [("0", "0", "0", ""), ("0", "0", "0", "0.5"), ("0", "0", "1", ""), ("1", "0", "1", ""), ("0", str(1 << 40), "1", "")],
    ids=["slabs", "slabs_overflow", "count_first_p2p", "count_first_nccl", "count_first_p2p_fullsort"],
)
def test_nccl_sample_sort_and_sync(tmp_path, no_peer, leaf_min_n, no_slabs, slack):
    """slabs: the slab exchange (pack + scatter into per-sender slabs, no host read) with the
    counting stage over the slabs; slabs_overflow: slabs too small on purpose — every rank sees
    the overflow flag, the slabs grow and the count-first protocol answers; count_first_p2p:
    te_rank_pack -> host read of the bucket sizes -> partition fused with P2P stores;
    count_first_nccl: the same with a local partition + NCCL all_to_all_single; leaf_min_n 2^40:
    the received keys go through the four-pass sort instead of the counting stage."""
    world = torch.cuda.device_count()
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    world = min(world, 8)
    out = tmp_path / "res"
    env = dict(os.environ, TE_TEST_OUT=str(out), TE_B200_NO_PEER=no_peer, TE_B200_LEAF_MIN_N=leaf_min_n, TE_B200_NO_SLABS=no_slabs)
    if slack:
        env["TE_B200_SLAB_SLACK"] = slack
    cmd = [
        sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
        "--master-addr", "127.0.0.1", "--master-port", str(29655 + 2 * int(no_peer) + (4 if leaf_min_n != "0" else 0) + 8 * int(no_slabs) + (16 if slack else 0)), str(ROOT / "tests" / "_dist_gpu_worker.py"),
    ]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    results = [json.loads(Path(f"{out}.{k}").read_text()) for k in range(world)]
    for k in range(1, world):
        assert results[k] == results[0]
    for key, val in results[0].items():
        if key.startswith("ovr/"):
            np.testing.assert_allclose(val[0], val[1], rtol=2e-6, err_msg=key)
        elif key.startswith("sharded/") or key.startswith("sharded_ap/"):
            got, want, P, tp, N, fp = val
            np.testing.assert_allclose(got, want, rtol=1e-9, err_msg=key)
            assert (P, N) == (tp, fp)
        elif key in ("sync/cm_equal", "coll/cm_equal"):
            assert val is True
        elif key == "coll/paths":
            assert val[0] == val[1], val
        else:
            np.testing.assert_allclose(val[0], val[1], rtol=1e-6, err_msg=key)
