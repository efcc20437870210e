#!/usr/bin/env python
"""Latency of the path's one collective (the sum of the per-start scores + counters over ranks), under torchrun:
NCCL all_reduce, the stand-alone peer-memory kernel (nrw_p2p_allreduce_i64) and the exchange fused into the walk's
tail, the latter on an EMPTY walk (no ids: launch + tail only) so that what is timed is the exchange itself.
Usage: torchrun --nproc-per-node N tools/allreduce_latency.py [n_words]"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "neural-random-walker_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def timed(fn, reps=300, warm=30):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / reps * 1e3], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    from nrw import dist as ndist, mesh
    from nrw.engine import Engine
    rank, world = ndist.init_from_env()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
    out = {"world": world, "n_words": n + 8}
    x = torch.ones(n + 8, dtype=torch.int64, device="cuda")
    out["nccl_all_reduce_us"] = timed(lambda: dist.all_reduce(x))
    csr = mesh.build_fluence_csr_cached(0.01, 1.0 / 15.0, 200, 0.15, 1)
    for tag, sw in (("push", {"NRW_XCHG": None}), ("pull", {"NRW_XCHG": "pull"})):
        eng = Engine(local, switches=sw)
        tabs = eng.upload(csr)
        red = ndist.ScoreAllReduce(n + 8, eng)
        out["standalone_kernel_%s_us" % tag] = timed(red.reduce)
        red.check()
        starts = np.arange(min(n, 900))
        ex = ndist.TallyExchange(n, eng)
        # an empty id window [0, 0): the launch walks nothing; its tail runs the exchange
        pws = [eng.prepare_walk(tabs, starts, 1, 64, 0, walker_range=(0, 0), score=ex.scores(h)[:len(starts)],
                                class_value=[0, 1, 0, 0], counters=ex.counters(h))[0] for h in (0, 1)]
        out["empty_walk_fused_%s_us" % tag] = timed(lambda: ex.launch(pws))
        assert int(ex.status_view().item()) == 0
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
