"""Worker of test_gpu_multi.py::test_p2p_score_allreduce_equals_nccl (run under torchrun, one rank per GPU)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "neural-random-walker_b200"))

from nrw import dist as ndist           # noqa: E402
from nrw.engine import Engine           # noqa: E402


def main():
    rank, world = ndist.init_from_env()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    eng = Engine(local)
    n = int(os.environ.get("NRW_TEST_N", "10000"))
    red = ndist.ScoreAllReduce(n, eng)
    want_p2p = os.environ.get("NRW_P2P_REDUCE", "1") != "0"
    assert red.p2p == want_p2p, (red.p2p, want_p2p)
    g = torch.Generator(device="cuda").manual_seed(100 + rank)
    for k in range(25):                                     # both buffers, many epochs, back to back
        buf = red.buffer()
        assert not buf.any(), (rank, k)                     # reduce() left the next step's buffer zeroed
        buf.copy_(torch.randint(-2 ** 40, 2 ** 40, (n,), generator=g, device="cuda", dtype=torch.int64))
        ref = buf.clone()
        out = red.reduce()
        dist.all_reduce(ref)
        assert torch.equal(out, ref), (rank, k)
    red.check()
    if rank == 0:
        print("p2p_reduce_worker ok p2p=%s world=%d" % (red.p2p, world), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
