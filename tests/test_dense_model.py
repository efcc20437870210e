"""CPU check of the dense-batch schedule of csrc/walk_dense.cuh: the lane-level model of the kernel
(oracle/dense_model.py: fresh batches, two survivor rings per warp, drains per segment) must produce
the integers of the numpy port of the device sampler (oracle/philox_port.py), which the GPU parity
tests hold the kernels to."""
import numpy as np
import pytest

from nrw import mesh, tables
from oracle import dense_model, philox_port


@pytest.fixture(scope="module")
def c1():
    csr = mesh.build_fluence_csr(0.01, 1.0 / 15.0, 200, 0.15, 1)
    src = mesh.source_values(csr.meta["mesh"]["meshpoints"], mesh.source_default)
    nc, _ = tables.classify_source(src, csr.n_rows)
    return csr, nc


def _same(model, port):
    assert np.array_equal(model["score"], port["score"])
    for k in ("steps", "exits", "survivors", "walkers"):
        assert model["counters"][k] == port["counters"][k], k
    assert model["counters"]["max_ring_fill"] < dense_model.DQ


@pytest.mark.parametrize("nt", [64, 4, 3, 2])
def test_fluence_mesh_ragged_starts(c1, nt):
    csr, nc1 = c1
    nc3 = (np.arange(csr.n_rows) % 4).astype(np.uint8)
    nc3[900] = 2
    starts = np.array([435, 0, 900, 14, 14, 899, 450, 16, 464])
    counts = np.array([300, 0, 17, 1, 133, 64, 5, 250, 120])
    total = int(counts.sum())
    carry = np.arange(9) * 2 ** 33 + (2 ** 32 - 150)          # stream ids cross a multiple of 2^32 inside a start
    for nc, cv in ((nc1, [0, 1, 0, 0]), (nc3, [5, 0, 255, 1])):
        at = tables.build_alias(csr, node_class=nc)
        for window, idb in ((None, None), ((299, 340), None), (None, carry), ((total - 1, total), None)):
            port = philox_port.walk_port(csr, starts, counts, nt, 5, alias=at, class_value=cv,
                                         walker_range=window, start_id_base=idb)
            for range_len in (32, 96, 4096):
                model = dense_model.walk_dense_model(at, starts, counts, nt, 5, class_value=cv, walker_range=window,
                                                     start_id_base=idb, range_len=range_len)
                _same(model, port)


def test_long_segments_fill_the_rings(c1):
    """Segments of thousands of ids: the rings run full (the capacity argument of walk_dense.cuh)."""
    csr, nc = c1
    at = tables.build_alias(csr, node_class=nc)
    starts = np.array([435, 464, 16])
    port = philox_port.walk_port(csr, starts, 5000, 64, 9, alias=at, class_value=[0, 1, 0, 0])
    model = dense_model.walk_dense_model(at, starts, 5000, 64, 9, class_value=[0, 1, 0, 0], range_len=4096, n_warps=2)
    _same(model, port)
    assert model["counters"]["max_ring_fill"] >= 32


@pytest.mark.parametrize("seed", list(range(12)))
def test_random_graphs(seed):
    """Random sparse graphs with absorbing rows, single-successor rows and walkers that outlive nt."""
    rng = np.random.default_rng(2000 + seed)
    n = int(rng.integers(5, 120))
    q = np.zeros((n, n))
    for i in range(n):
        kind = rng.random()
        if kind < 0.15:
            continue
        if kind < 0.35:
            q[i, rng.integers(0, n)] = 1.0
            continue
        k = int(rng.integers(2, min(n, 12) + 1))
        cols = rng.choice(n, size=k, replace=False)
        q[i, cols] = rng.dirichlet(np.ones(k))
    n_live = int(rng.integers(max(n // 2, 1), n + 1))
    csr = mesh.csr_from_dense(q, n_live=n_live)
    nc = rng.integers(0, 4, size=n).astype(np.uint8)
    at = tables.build_alias(csr, node_class=nc)
    if at.layout != "compact":
        pytest.skip("wide tables stay with the general kernel")
    n_starts = int(rng.integers(1, 8))
    starts = rng.integers(0, n, size=n_starts)
    counts = rng.integers(0, 400, size=n_starts)
    if counts.sum() == 0:
        counts[0] = 50
    cv = [0, 1, 2, 3] if rng.random() < 0.5 else [int(x) for x in rng.integers(0, 256, size=4)]
    nt = int(rng.integers(2, 30))
    idb = None if rng.random() < 0.5 else (np.arange(1, n_starts + 1, dtype=np.int64) << 32) - rng.integers(0, np.maximum(counts, 1) + 1)
    port = philox_port.walk_port(csr, starts, counts, nt, seed, alias=at, class_value=cv, start_id_base=idb)
    model = dense_model.walk_dense_model(at, starts, counts, nt, seed, class_value=cv, start_id_base=idb,
                                         range_len=int(rng.choice([32, 64, 512])))
    _same(model, port)
