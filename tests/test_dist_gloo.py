"""N>1 host logic on CPU: two gloo ranks shard every start's walkers exactly as nrw.fluence does on
NCCL (shard_range + start_id_base), each rank's tallies come from the numpy port of the device
sampler (oracle/philox_port.py -- checker only), one all-reduce, and the sum must equal the
unsharded run bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import C1, PKG, ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, M, nt, seed, q):
    for p in (PKG, ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    from nrw import dist as ndist, mesh, tables
    from oracle import philox_port
    r, w = ndist.init_from_env(backend="gloo")
    assert (r, w) == (rank, world) and ndist.world() == (rank, world)
    csr = mesh.build_fluence_csr(**C1)
    src = mesh.source_values(csr.meta["mesh"]["meshpoints"], mesh.source_default)
    node_class, _ = tables.classify_source(src, csr.n_rows)
    at = tables.build_alias(csr, node_class=node_class)
    starts = np.arange(0, csr.n_live, 37)
    lo, hi = ndist.shard_range(M, rank, world)
    id_base = starts.astype(np.int64) * M + lo
    out = philox_port.walk_port(csr, starts, hi - lo, nt, seed, alias=at, class_value=[0, 1, 0, 0],
                                start_id_base=id_base)
    score = torch.from_numpy(out["score"].copy())
    hist = torch.from_numpy(out["node_hist"].copy())
    cnt = torch.tensor([out["counters"]["steps"], out["counters"]["exits"]], dtype=torch.int64)
    ndist.allreduce_sum_([score, hist, None, cnt])
    if rank == 0:
        q.put((score.numpy(), hist.numpy(), cnt.numpy()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_sharded_tallies_allreduce_to_the_unsharded_result():
    for p in (PKG, ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    from nrw import mesh, tables
    from oracle import philox_port
    M, nt, seed = 41, 64, 11                    # odd M: ragged shards (21 + 20)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, M, nt, seed, q)) for r in range(2)]
    for p in procs:
        p.start()
    score, hist, cnt = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    csr = mesh.build_fluence_csr(**C1)
    src = mesh.source_values(csr.meta["mesh"]["meshpoints"], mesh.source_default)
    node_class, _ = tables.classify_source(src, csr.n_rows)
    at = tables.build_alias(csr, node_class=node_class)
    starts = np.arange(0, csr.n_live, 37)
    ref = philox_port.walk_port(csr, starts, M, nt, seed, alias=at, class_value=[0, 1, 0, 0],
                                start_id_base=starts.astype(np.int64) * M)
    assert np.array_equal(score, ref["score"])
    assert np.array_equal(hist, ref["node_hist"])
    assert cnt[0] == ref["counters"]["steps"] and cnt[1] == ref["counters"]["exits"]


def test_shard_range_covers_every_walker_once_for_any_world():
    from nrw import dist as ndist
    for total in (0, 1, 7, 100, 12501):
        for world in (1, 2, 3, 4, 8):
            for align in (1, 32):
                got = [ndist.shard_range(total, r, world, align) for r in range(world)]
                assert got[0][0] == 0 and got[-1][1] == total
                for (a, b), (c, d) in zip(got, got[1:]):
                    assert b == c and a <= b
