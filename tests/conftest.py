import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(Path(__file__).resolve().parent))
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


class Golden:
    """tests/golden/reference_vectors.npz: inputs + outputs of the reference (oracle/gen_golden.py)."""

    def __init__(self):
        self._z = np.load(ROOT / "tests" / "golden" / "reference_vectors.npz")

    def case(self, name: str) -> dict:
        pre = name + "/"
        d = {k[len(pre):]: self._z[k] for k in self._z.files if k.startswith(pre)}
        assert d, f"no golden case {name}"
        return d

    def cases(self, group: str) -> list:
        pre = group + "/"
        return sorted({k.split("/")[1] for k in self._z.files if k.startswith(pre)})


@pytest.fixture(scope="session")
def golden():
    return Golden()
