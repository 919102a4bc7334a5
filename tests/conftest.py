"""pytest configuration: `gpu` marker, repo root on sys.path, shared helpers."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def lib():
    from mink_octtree_stablediffusion_b200 import _lib
    return _lib.load()


def random_cloud(n, span, batch, seed, ncols=4, dup_frac=0.0):
    """Random int32 coordinates [n, ncols] (batch first) in [-span, span)."""
    rng = np.random.default_rng(seed)
    c = rng.integers(-span, span, size=(n, ncols), dtype=np.int64)
    c[:, 0] = rng.integers(0, batch, size=n)
    if dup_frac > 0 and n > 1:
        k = int(n * dup_frac)
        src = rng.integers(0, n, size=k)
        dst = rng.integers(0, n, size=k)
        c[dst] = c[src]
    return c.astype(np.int32)
