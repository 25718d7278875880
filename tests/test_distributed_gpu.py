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


@pytest.mark.parametrize("no_peer", ["0", "1"])
def test_nccl_sample_sort_and_sync(tmp_path, no_peer):
    """no_peer=0: partition fused with the exchange (P2P stores into peer inboxes);
    no_peer=1: the fallback, local partition + NCCL all_to_all_single."""
    world = torch.cuda.device_count()
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    world = min(world, 8)
    out = tmp_path / "res"
    env = dict(os.environ, TE_TEST_OUT=str(out), TE_B200_NO_PEER=no_peer)
    cmd = [
        sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
        "--master-addr", "127.0.0.1", "--master-port", "29655" if no_peer == "0" else "29657", str(ROOT / "tests" / "_dist_gpu_worker.py"),
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
        elif key == "sync/cm_equal":
            assert val is True
        else:
            np.testing.assert_allclose(val[0], val[1], rtol=1e-6, err_msg=key)
