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


SWITCHES = ("NRW_K1", "NRW_DET", "NRW_TAB", "NRW_CPS", "NRW_RANGES_PER_WARP", "NRW_RANGE_TAIL", "NRW_TAIL_RANGES",
            "NRW_CNT_MATCH", "NRW_CNT_CACHE", "NRW_CNT_PATH", "NRW_STR")


class _EnginePerSwitches:
    """The library reads its NRW_* measurement switches once, in nrw_create (include/nrw.h).  Parity tests pick a
    code path with ``monkeypatch.setenv("NRW_...")`` and then call ``engine.walk(...)``: this proxy hands every
    call to an engine created under the switches in force at that moment (one engine per combination; device
    tables are plain device pointers and serve every engine of the device)."""

    def __init__(self):
        self._engines = {}

    def _current(self):
        from nrw.engine import Engine, get_engine
        key = tuple(os.environ.get(k) for k in SWITCHES)
        if key not in self._engines:
            self._engines[key] = get_engine() if not any(key) else Engine(0)
        return self._engines[key]

    def __getattr__(self, name):
        return getattr(self._current(), name)


@pytest.fixture(scope="session")
def engine():
    return _EnginePerSwitches()


@pytest.fixture(scope="session")
def c1_csr():
    from nrw import mesh
    return mesh.build_fluence_csr(C1["dt"], C1["da"], C1["v"], C1["SigS"], C1["L"])


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return {n[:-4]: np.load(os.path.join(GOLDEN, n)) for n in os.listdir(GOLDEN) if n.endswith(".npz")}
