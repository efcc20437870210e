"""Multi-GPU: one process per GPU, walkers sharded by global id, ONE allreduce of the tallies.

Walkers never interact and every trajectory is a pure function of (seed, global walker id), so the
ranks simply take disjoint contiguous id ranges of the same start list; the transition tables
(<= a few MB) are replicated.  The only exchange is the sum of the per-start score / histogram
buffers at the end (NCCL over NVLink on the GPUs; ``gloo`` in the CPU tests).  The reference has
no distributed code at all; this replaces its serial loop over start nodes
(fluence_mini_app.py:397).
"""
from __future__ import annotations

import os


def world():
    """(rank, world_size) from torch.distributed if initialised, else from the torchrun env."""
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(), dist.get_world_size()
    except Exception:
        pass
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def init_from_env(backend=None):
    """Initialise torch.distributed from torchrun's environment (no-op for a single process)."""
    import torch
    import torch.distributed as dist
    ws = int(os.environ.get("WORLD_SIZE", "1"))
    if ws <= 1 or (dist.is_available() and dist.is_initialized()):
        return world()
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    if backend == "nccl":
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dist.init_process_group(backend=backend)
    return world()


def shard_range(total: int, rank: int, world_size: int, align: int = 1):
    """Contiguous slice [lo, hi) of ``total`` walker ids for ``rank``; boundaries are multiples of
    ``align`` (except the last) so that a warp's id range never straddles two ranks' work."""
    if world_size <= 1:
        return 0, int(total)
    per = -(-int(total) // world_size)
    if align > 1:
        per = -(-per // align) * align
    lo = min(rank * per, total)
    hi = min(lo + per, total)
    return int(lo), int(hi)


def interleave_starts(n_starts: int, world_size: int):
    """Order of start indices that balances the ~50 % instant-exit start nodes across contiguous
    id shards: not needed when every rank takes an id range of EVERY start (see shard_by_start)."""
    return list(range(n_starts))


def allreduce_sum_(tensors):
    """In-place sum over ranks of each tensor (int64 / float64); the single collective of the path."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return tensors
    for t in tensors:
        if t is not None:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return tensors


def broadcast_int(value: int, src: int = 0) -> int:
    """Rank ``src``'s value on every rank (the Philox seed: with the reference's time-based default
    seed, fluence_mini_app.py:517-521, every process would otherwise draw its own)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return int(value)
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    t = torch.tensor([int(value)], dtype=torch.int64, device=dev)
    dist.broadcast(t, src=src)
    return int(t.item())
