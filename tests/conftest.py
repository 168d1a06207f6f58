import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); deselected by -m 'not gpu'")


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    path = os.path.join(ROOT, "tests", "golden", "reference_vectors.npz")
    return dict(np.load(path))


@pytest.fixture(scope="session")
def golden_scalars():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "reference_scalars.json")) as fh:
        return json.load(fh)


@pytest.fixture(autouse=True)
def _fp32_oracle_and_clean_backend_flags():
    """The oracle is a float32 (or float64) evaluation: cuDNN / cuBLAS must not round its operands to TF32, whatever
    algorithm they pick, and backend flags a test flips (the vendored reference suites set cudnn.deterministic /
    benchmark in their seeding helpers, which changes the algorithm eager convolutions choose) must not leak into the
    tests collected after it."""
    import torch
    saved = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32,
             torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32,
     torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark) = saved
