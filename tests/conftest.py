import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    # a GPU test on a box without a GPU would end in the library's GPUassert + exit;
    # skip instead (the driver runs `-m gpu` only where a B200 is present)
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle
    pyoracle.build()
    return pyoracle


@pytest.fixture(scope="session")
def tsp_matrix():
    with open(os.path.join(GOLDEN, "tsp100.txt")) as f:
        tok = f.read().split()
    n = int(tok[0])
    return np.array(tok[1:1 + n * n], dtype=np.float32).reshape(n, n)


@pytest.fixture()
def pga():
    import libpga_b200
    made = []

    def make(seed=42, quiet=True):
        p = libpga_b200.PGA(seed=seed, quiet=quiet)
        made.append(p)
        return p

    yield make
    for p in made:
        p.deinit()
