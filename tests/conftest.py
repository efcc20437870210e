import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "neural-random-walker_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")

C1 = dict(dt=0.01, da=1.0 / 15.0, v=200, SigS=0.15, L=1)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_cuda():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def engine():
    from nrw.engine import get_engine
    return get_engine()


@pytest.fixture(scope="session")
def c1_csr():
    from nrw import mesh
    return mesh.build_fluence_csr(C1["dt"], C1["da"], C1["v"], C1["SigS"], C1["L"])


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return {n[:-4]: np.load(os.path.join(GOLDEN, n)) for n in os.listdir(GOLDEN) if n.endswith(".npz")}
