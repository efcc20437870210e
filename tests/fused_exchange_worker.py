"""Worker of test_gpu_multi.py::test_fused_walk_exchange_two_ranks (run under torchrun, one rank per GPU): the walk
kernel's own tail sums scores + counters over ranks (csrc/walk_tail.cuh) -- against dist.all_reduce of the per-rank
tallies and against rank 0's unsharded walk of the same seed, over many back-to-back solves on both buffers; then the
same through the drop-in solver (nrw.fluence.solve_fluence) and the C-ABI host entry."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "neural-random-walker_b200"))

from nrw import dist as ndist, mesh, tables   # noqa: E402
from nrw.engine import get_engine             # noqa: E402
from nrw.fluence import solve_fluence         # noqa: E402


def main():
    rank, world = ndist.init_from_env()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    eng = get_engine(local)
    csr = mesh.build_fluence_csr_cached(0.01, 1.0 / 15.0, 200, 0.15, 1)
    src = mesh.source_values(csr.meta["mesh"]["meshpoints"], mesh.source_default)
    nc, _ = tables.classify_source(src, csr.n_rows)
    tabs = eng.upload(csr, node_class=nc)
    starts = np.arange(900)
    M = 2001                                                 # ragged shards
    lo, hi = ndist.shard_range(M, rank, world)
    id_base = starts.astype(np.int64) * M + lo
    ex = ndist.TallyExchange(len(starts), eng)
    want_p2p = os.environ.get("NRW_P2P_REDUCE", "1") != "0"
    assert ex.pair.p2p == want_p2p, (ex.pair.p2p, want_p2p)
    pws = [eng.prepare_walk(tabs, starts, hi - lo, 64, 0, score=ex.scores(h), class_value=[0, 1, 0, 0],
                            start_id_base=id_base, counters=ex.counters(h))[0] for h in (0, 1)]
    for k in range(14):
        seed = 50 + k
        for pw in pws:
            pw.set_seed(seed)
        out = ex.launch(pws).clone()
        mine = ex.pair.half(k & 1).clone()
        dist.all_reduce(mine)
        assert torch.equal(out, mine), (rank, k)
        assert int(ex.status_view().item()) == 0
        one = eng.walk(tabs, starts, M, 64, seed, score=True, class_value=[0, 1, 0, 0],
                       start_id_base=starts.astype(np.int64) * M)
        assert torch.equal(out[:900], one.score), (rank, k)
        assert torch.equal(out[900:905], one.counters[:5]), (rank, k)
    # the drop-in solver under torchrun: sharded by walker id inside solve_fluence
    res = solve_fluence(csr, M, 64, 0, 0.01, 200, rng="philox", seed=7)
    one = eng.walk(tabs, starts, M, 64, 7, score=True, class_value=[0, 1, 0, 0], start_id_base=starts.astype(np.int64) * M)
    assert np.array_equal(res.score, one.score.cpu().numpy()) and res.counters["steps"] == one.counters_dict()["steps"]
    # the C-ABI host entry with the exchange: reduced scores on every rank's host
    if ex.fused:
        at = tabs.alias_host
        score = np.zeros(900, np.int64)
        cnt = np.zeros(8, np.int64)
        ex2 = ndist.TallyExchange(900, eng)
        solve = eng.prepare_fluence_scores_host(at, csr.n_rows, csr.n_live, starts.astype(np.int32),
                                                np.arange(901, dtype=np.int64) * (hi - lo), 0, 900 * (hi - lo), 64,
                                                [0, 1, 0, 0], score, cnt, start_id_base=id_base, exchange=ex2)
        for seed in (7, 8, 7):
            solve(seed)
        assert np.array_equal(score, one.score.cpu().numpy()) and cnt[0] == one.counters_dict()["steps"]
    if rank == 0:
        print("fused_exchange_worker ok p2p=%s world=%d" % (ex.pair.p2p, world), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
