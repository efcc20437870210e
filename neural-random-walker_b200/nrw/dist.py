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


def allreduce_sum_(tensors):
    """In-place sum over ranks of each tensor (int64 / float64); the single collective of the path."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return tensors
    for t in tensors:
        if t is not None:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return tensors


class ScoreAllReduce:
    """The path's one collective -- the sum over ranks of the per-start score vector -- as ONE kernel
    per rank over NVLink peer memory (csrc/p2p_reduce.cuh) instead of an NCCL call: 80 KB per peer is
    latency-bound, and NCCL's launch + synchronisation costs 28-39 us at N = 8 against a 0.63 ms step.

    Two score buffers live in torch's symmetric memory (mapped into every peer); the walk of step k
    writes ``buffer()`` directly, ``reduce()`` publishes / awaits epoch k and sums all peers' vectors
    into ``out``.  Falls back to NCCL (same results: the sum is integer) when symmetric memory is not
    available, when ``NRW_P2P_REDUCE=0``, or on a single rank; all ranks take the same decision."""

    def __init__(self, n: int, engine):
        import torch
        import torch.distributed as dist
        self.n, self.engine, self.torch = int(n), engine, torch
        self.rank, self.world = world()
        self.epoch = 0
        self.p2p = False
        dev = engine.device
        self.out = torch.zeros(self.n, dtype=torch.int64, device=dev)
        self._bufs = torch.zeros(2 * self.n, dtype=torch.int64, device=dev)
        if self.world > 1 and dist.is_initialized() and dist.get_backend() == "nccl":
            ok = 0
            if os.environ.get("NRW_P2P_REDUCE", "1") != "0":
                try:
                    import torch.distributed._symmetric_memory as symm
                    group = dist.group.WORLD.group_name
                    bufs = symm.empty(2 * self.n, dtype=torch.int64, device=dev)
                    flags = symm.empty(max(self.world, 8), dtype=torch.int64, device=dev)
                    bufs.zero_()
                    flags.zero_()
                    self._hb = symm.rendezvous(bufs, group)
                    self._hf = symm.rendezvous(flags, group)
                    if (int(self._hb.buffer_ptrs[self.rank]) != bufs.data_ptr()
                            or int(self._hf.buffer_ptrs[self.rank]) != flags.data_ptr()):
                        raise RuntimeError("symmetric allocation does not start at the mapped base")
                    self._hb.barrier(channel=0)                      # every rank's flags are zero from here on
                    self._sym = (bufs, flags)
                    self._err = torch.zeros(1, dtype=torch.int32, device=dev)
                    ok = 1
                except Exception:
                    ok = 0
            agreed = torch.tensor([ok], dtype=torch.int32, device=dev)
            dist.all_reduce(agreed, op=dist.ReduceOp.MIN)
            self.p2p = bool(agreed.item())
            if self.p2p:
                self._bufs = self._sym[0]

    def half(self, h: int):
        """int64[n] view of score buffer h (0 / 1); reduce number k (from 0) sums buffer k & 1."""
        return self._bufs[h * self.n:(h + 1) * self.n]

    def buffer(self):
        """The view the NEXT reduce() will sum: pass it to the walk as its score buffer (zero it first)."""
        return self.half(self.epoch & 1)

    def reduce(self):
        """Sum of every rank's ``buffer()`` on every rank (int64[n], valid until the next reduce()).
        Also zeroes the OTHER buffer -- the one the next step's walk accumulates into -- so callers
        never zero score buffers themselves."""
        buf = self.buffer()
        half = self.epoch & 1
        self.epoch += 1
        other = self.half(half ^ 1)
        if self.world == 1:
            other.zero_()
            return buf
        if not self.p2p:
            import torch.distributed as dist
            self.out.copy_(buf)
            dist.all_reduce(self.out, op=dist.ReduceOp.SUM)
            other.zero_()
            return self.out
        e = self.engine
        e._check(e.lib.nrw_p2p_allreduce_i64(e.handle, self.out.data_ptr(), self._hb.buffer_ptrs_dev,
                                             self._hf.buffer_ptrs_dev, half * self.n, (half ^ 1) * self.n,
                                             self.n, self.rank, self.world, self.epoch,
                                             self._err.data_ptr(), e.stream))
        return self.out

    def settle(self) -> bool:
        """Collective: after some warm-up reductions, agree on whether the peer-memory path works on
        every rank (no time-out anywhere); if not, all ranks switch to NCCL for good.  Returns p2p."""
        if self.world > 1 and self.p2p:
            import torch.distributed as dist
            bad = self._err.clone()
            dist.all_reduce(bad, op=dist.ReduceOp.MAX)
            if int(bad.item()) != 0:
                self.p2p = False
                self._err.zero_()
                self._bufs.zero_()
                self.epoch = 0
        return self.p2p

    def check(self):
        """Synchronises; raises if a peer did not arrive at some reduce()."""
        if self.p2p and int(self._err.item()) != 0:
            raise RuntimeError("nrw.dist.ScoreAllReduce: a peer did not reach the reduction within 5 s")


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
