import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name):
    """npz with 'case/key' entries -> {case: {key: array}} (+ top-level keys under '')."""
    z = np.load(os.path.join(GOLDEN, name))
    out = {}
    for k in z.files:
        if "/" in k:
            c, kk = k.split("/", 1)
            out.setdefault(c, {})[kk] = z[k]
        else:
            out.setdefault("", {})[k] = z[k]
    return out


def rel_err(a, b):
    a = torch.as_tensor(a, dtype=torch.float64).flatten()
    b = torch.as_tensor(b, dtype=torch.float64).flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.fixture(scope="session")
def lib_built():
    import multimodal_2025_b200 as m
    m.build()
    return m
