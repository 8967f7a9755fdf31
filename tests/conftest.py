import json
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
import s2lc_b200  # noqa: E402,F401  (registers the package alias)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

# torch references on the GPU must be true fp32 (cuDNN / cuBLAS default to TF32 for convs)
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run by the driver with -m gpu)")


@pytest.fixture(scope="session")
def golden_meta():
    with open(os.path.join(GOLDEN_DIR, "golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def golden_arrays():
    return dict(np.load(os.path.join(GOLDEN_DIR, "golden.npz")))


@pytest.fixture(scope="session")
def state_shapes():
    with open(os.path.join(GOLDEN_DIR, "state_shapes.json")) as f:
        return json.load(f)


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    """||a-b||_2 / ||b||_2."""
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))
