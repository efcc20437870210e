#!/usr/bin/env python
"""Latency of the path's one collective (all-reduce of the per-start scores) by dtype, under torchrun."""
import json
import os
import sys

import torch
import torch.distributed as dist


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
    out = {}
    for name, dt in (("int64", torch.int64), ("float64", torch.float64), ("float32", torch.float32), ("int32", torch.int32)):
        x = torch.ones(n, dtype=dt, device="cuda")
        for _ in range(20):
            dist.all_reduce(x)
        torch.cuda.synchronize()
        dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(200):
            dist.all_reduce(x)
        b.record()
        torch.cuda.synchronize()
        out[name] = a.elapsed_time(b) / 200 * 1e3
    if dist.get_rank() == 0:
        print(json.dumps({"world": dist.get_world_size(), "n": n, "us_per_allreduce": out}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
