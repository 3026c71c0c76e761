"""pytest configuration: `gpu` marker, repo root on sys.path, golden fixtures."""
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


class Golden:
    """Attribute access to the arrays of one prefix of tests/golden/golden_v1.npz."""

    def __init__(self, data, prefix):
        self._d = {k[len(prefix) + 1:]: data[k] for k in data.files if k.startswith(prefix + ".")}

    def __getattr__(self, name):
        try:
            v = self._d[name]
        except KeyError as e:
            raise AttributeError(name) from e
        return v.item() if v.ndim == 0 else v

    def table_meta(self):
        rows = np.asarray(self.rows, dtype=np.int64)
        base = np.concatenate([[0], np.cumsum(rows)[:-1]]).astype(np.int64)
        col = (np.arange(self.F) * self.D).astype(np.int64)
        return base, rows, col


@pytest.fixture(scope="session")
def golden():
    data = np.load(ROOT / "tests" / "golden" / "golden_v1.npz")
    return lambda prefix: Golden(data, prefix)
