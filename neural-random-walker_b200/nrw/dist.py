"""Multi-GPU: one process per GPU, walkers sharded by global id, ONE allreduce of the tallies.

Walkers never interact and every trajectory is a pure function of (seed, global walker id), so the
ranks simply take disjoint contiguous id ranges of the same start list; the transition tables
(<= a few MB) are replicated.  The only exchange is the sum of the per-start score / histogram
buffers at the end (NCCL over NVLink on the GPUs; ``gloo`` in the CPU tests).  The reference has
no distributed code at all; this replaces its serial loop over start nodes
(fluence_mini_app.py:397).
"""
from __future__ import annotations

import logging
import os

rw_logger = logging.getLogger("NRW")


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


class _SymmetricPair:
    """Two alternating int64 tally buffers of ``n`` words (padded to an even count: 16-byte aligned halves), the
    output region the sums over ranks land in, a status word, and a flag array -- in torch's symmetric memory when
    every rank can map every peer (NVLink / NVSwitch box, NCCL backend), else in plain device memory.  Layout (words):
    [0, n_pad) half 0 | [n_pad, 2 n_pad) half 1 | [2 n_pad, 3 n_pad) sums | [3 n_pad] status.  All ranks take the same decision (one MIN all-reduce); whenever the
    peer-memory path is not taken on a multi-rank job the reason is logged once."""

    def __init__(self, n: int, engine):
        import torch
        import torch.distributed as dist
        self.n, self.engine, self.torch = int(n), engine, torch
        self.n_pad = (self.n + 1) & ~1
        self.rank, self.world = world()
        self.p2p = False
        dev = engine.device
        self.words = 3 * self.n_pad + 2
        self.bufs = torch.zeros(self.words, dtype=torch.int64, device=dev)
        self.hb = self.hf = None
        if self.world > 1 and dist.is_initialized():
            ok, why = 0, ""
            if dist.get_backend() != "nccl":
                why = "backend %s has no peer memory" % dist.get_backend()
            elif os.environ.get("NRW_P2P_REDUCE", "1") == "0":
                why = "NRW_P2P_REDUCE=0"
            else:
                try:
                    import torch.distributed._symmetric_memory as symm
                    group = dist.group.WORLD.group_name
                    bufs = symm.empty(self.words, dtype=torch.int64, device=dev)
                    flags = symm.empty(8 * max(self.world, 2), dtype=torch.int64, device=dev)   # 8 CTAs x world words (stand-alone kernel)
                    bufs.zero_()
                    flags.zero_()
                    hb = symm.rendezvous(bufs, group)
                    hf = symm.rendezvous(flags, group)
                    if (int(hb.buffer_ptrs[self.rank]) != bufs.data_ptr()
                            or int(hf.buffer_ptrs[self.rank]) != flags.data_ptr()):
                        raise RuntimeError("symmetric allocation does not start at the mapped base")
                    hb.barrier(channel=0)                            # every rank's flags are zero from here on
                    self._sym = (bufs, flags, hb, hf)
                    ok = 1
                except Exception as exc:                             # no symmetric memory on this box / torch
                    why = "symmetric memory unavailable (%s: %s)" % (type(exc).__name__, exc)
            agreed = torch.tensor([ok], dtype=torch.int32,
                                  device=dev if dist.get_backend() == "nccl" else "cpu")
            dist.all_reduce(agreed, op=dist.ReduceOp.MIN)
            self.p2p = bool(agreed.item())
            if self.p2p:
                self.bufs, _, self.hb, self.hf = self._sym
            else:
                rw_logger.warning("nrw.dist: rank %d sums tallies with the backend's all_reduce instead of the "
                                  "peer-memory kernel: %s", self.rank, why or "a peer could not map symmetric memory")

    def half(self, h: int):
        return self.bufs[h * self.n_pad:h * self.n_pad + self.n]

    @property
    def out_offset(self):
        return 2 * self.n_pad

    def sums(self):
        """int64[n] view of the region the exchange writes the sums over ranks into."""
        return self.bufs[2 * self.n_pad:2 * self.n_pad + self.n]

    def status(self):
        return self.bufs[3 * self.n_pad:3 * self.n_pad + 1]


class ScoreAllReduce:
    """The path's one collective as a STAND-ALONE step: the sum over ranks of an int64 tally vector as ONE kernel
    per rank over NVLink peer memory (csrc/p2p_reduce.cuh) instead of an NCCL call.  The walks themselves run the
    same exchange in their own tail (``TallyExchange`` below); this class serves tallies that no single walk launch
    produced and the tests that hold the kernel to ``dist.all_reduce``.

    Two buffers live in symmetric memory; the producer of step k writes ``buffer()``, ``reduce()`` publishes /
    awaits epoch k and sums all peers' vectors into ``out``.  Falls back to the backend's all_reduce (same result:
    the sum is integer) when symmetric memory is not available -- the reason is logged."""

    def __init__(self, n: int, engine):
        import torch
        self.pair = _SymmetricPair(n, engine)
        self.n, self.engine, self.torch = int(n), engine, torch
        self.rank, self.world = self.pair.rank, self.pair.world
        self.epoch = 0
        self.out = self.pair.sums()                       # the peers push their slices of the sums here
        self._err = torch.zeros(1, dtype=torch.int32, device=engine.device)

    @property
    def p2p(self):
        return self.pair.p2p

    def half(self, h: int):
        """int64[n] view of buffer h (0 / 1); reduce number k (from 0) sums buffer k & 1."""
        return self.pair.half(h)

    def buffer(self):
        """The view the NEXT reduce() will sum."""
        return self.half(self.epoch & 1)

    def reduce(self):
        """Sum of every rank's ``buffer()`` on every rank (int64[n], valid until the next reduce()).
        Also zeroes the OTHER buffer -- the one the next step accumulates into."""
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
        e, pr = self.engine, self.pair
        e._check(e.lib.nrw_p2p_allreduce_i64(e.handle, pr.hb.buffer_ptrs_dev, pr.hf.buffer_ptrs_dev,
                                             half * pr.n_pad, (half ^ 1) * pr.n_pad, pr.out_offset,
                                             self.n, self.rank, self.world, self.epoch,
                                             self._err.data_ptr(), e.stream))
        return self.out

    def check(self):
        """Synchronises; raises if a peer did not arrive at some reduce() (its result is then invalid)."""
        if self.p2p and int(self._err.item()) != 0:
            raise RuntimeError("nrw.dist.ScoreAllReduce: a peer did not reach the reduction within 5 s")


class TallyExchange:
    """Scores and counters of a fused solve in ONE int64 buffer per rank, double-buffered, with everything that
    used to surround the walk launch folded into the walk's own tail (csrc/walk_tail.cuh, nrw_walk.reduce): the
    sum over ranks over NVLink peer memory (N > 1) and the zeroing of the next launch's buffer.  One kernel per
    solve on any number of GPUs; no memset, no separate collective.

        ex = TallyExchange(n_starts, engine)
        pws = [engine.prepare_walk(..., score=ex.scores(h), counters=ex.counters(h))[0] for h in (0, 1)]
        result = ex.launch(pws)          # device int64[n_starts + 8]: sums over ranks; valid until the launch after next

    Without peer memory on a multi-rank job (gloo, no NVLink mapping) the walk is followed by the backend's
    all_reduce; the choice is logged (``_SymmetricPair``)."""

    def __init__(self, n_starts: int, engine):
        import torch
        from .engine import NRW_N_COUNTERS, _Reduce
        self.n_starts = int(n_starts)
        self.n = self.n_starts + NRW_N_COUNTERS
        self.pair = _SymmetricPair(self.n, engine)
        self.engine, self.torch = engine, torch
        self.rank, self.world = self.pair.rank, self.pair.world
        self.epoch = 0
        r = _Reduce()
        r.rank, r.world = (self.rank, self.world) if self.pair.p2p else (0, 1)
        if self.pair.p2p:
            r.peer_bufs_dev, r.peer_flags_dev = self.pair.hb.buffer_ptrs_dev, self.pair.hf.buffer_ptrs_dev
        r.own_buf, r.n, r.out_offset = self.pair.bufs.data_ptr(), self.n, self.pair.out_offset
        self._r = r

    @property
    def fused(self):
        """True when the walk's tail does the whole exchange (single rank, or peer memory on every rank)."""
        return self.world == 1 or self.pair.p2p

    def scores(self, h):
        return self.pair.half(h)[:self.n_starts]

    def counters(self, h):
        return self.pair.half(h)[self.n_starts:]

    def arm(self):
        """The nrw_reduce descriptor of the next launch (tallies in half epoch & 1, the other half zeroed in the
        tail); advances the epoch.  Returns (half index, ctypes pointer)."""
        import ctypes as C
        h = self.epoch & 1
        self.epoch += 1
        r = self._r
        r.offset, r.zero_offset, r.epoch = h * self.pair.n_pad, (h ^ 1) * self.pair.n_pad, self.epoch
        return h, C.pointer(r)

    def launch(self, pws):
        """One solve: walk into the current half, exchange in the tail.  Returns the device view of the summed
        tallies (scores, then counters)."""
        h, rp = self.arm()
        pw = pws[h]
        pw.w.reduce = rp
        pw.launch_kernel()
        if self.world == 1:
            return self.pair.half(h)
        out = self.pair.sums()
        if not self.pair.p2p:
            import torch.distributed as dist
            out.copy_(self.pair.half(h))
            dist.all_reduce(out, op=dist.ReduceOp.SUM)
        return out

    def status_view(self):
        """Device int64[1]: 0 after a good exchange, 1 when a peer did not arrive (copy it back with the result)."""
        return self.pair.status()


def broadcast_array(arr, src: int = 0):
    """Rank ``src``'s float64 array on every rank (same shape everywhere)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return arr
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    t = torch.from_numpy(np.ascontiguousarray(arr, dtype=np.float64)).to(dev)
    dist.broadcast(t, src=src)
    return t.cpu().numpy()


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
