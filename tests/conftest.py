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
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


class Golden:
    """One tests/golden/*.npz case: x, dy, y, gx, sd/<name>, g/<name> (made by oracle/make_golden.py)."""

    def __init__(self, name):
        z = np.load(os.path.join(GOLDEN, name + ".npz"))
        self.raw = {k: z[k] for k in z.files}

    def t(self, key, device="cpu", dtype=None):
        v = torch.from_numpy(self.raw[key]).to(device)
        return v if dtype is None or not v.is_floating_point() else v.to(dtype)

    def sd(self, device="cpu", dtype=None):
        return {k[3:]: self.t(k, device, dtype) for k in self.raw if k.startswith("sd/")}

    def grads(self, device="cpu"):
        return {k[2:]: self.t(k, device) for k in self.raw if k.startswith("g/")}


def load_golden(name):
    return Golden(name)


def relerr(a, b):
    """Norm-wise relative error max|a-b| / max|b|."""
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    denom = b.abs().max().item()
    return (a - b).abs().max().item() / (denom if denom > 0 else 1.0)
