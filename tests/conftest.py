import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run on the GPU box with -m gpu)")


@pytest.fixture(scope="session")
def orc():
    """The CPU oracle (test infrastructure; never imported by the product package)."""
    from oracle import oracle

    oracle.lib()
    return oracle


@pytest.fixture(scope="session")
def assets():
    """PIL-decoded pixels of the reference's four shipped JPEGs (tests/golden/make_golden.py)."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "assets.npz"))
    return {k: g[k] for k in g.files}


@pytest.fixture(scope="session")
def ctx():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import stepth_b200

    return stepth_b200.Context(0)


def rgba_of(rgb: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(np.concatenate([rgb, np.full(rgb.shape[:2] + (1,), 255, np.uint8)], axis=-1))
