import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def unit_features(rng, n, d, clusters=0):
    """Seeded unit-norm features like the ResUNet output (model/resunet.py:261 normalises)."""
    f = rng.standard_normal((n, d)).astype(np.float32)
    if clusters:
        c = rng.standard_normal((clusters, d)).astype(np.float32)
        f = c[rng.integers(0, clusters, n)] + 0.3 * f
    f /= np.linalg.norm(f, axis=1, keepdims=True)
    return f.astype(np.float32)


@pytest.fixture
def rng():
    return np.random.default_rng(31)


@pytest.fixture(scope="session")
def ctx():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from corsair_b200 import get_context
    return get_context(0)
