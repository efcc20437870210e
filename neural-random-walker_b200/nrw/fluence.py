"""Fused fluence solve: all start nodes of run() (fluence_mini_app.py:397-438) in one pass.

The reference calls its back end once per start node, materialises tempdata[(meshSize+1), nt] and
then loops over all nt*meshSize cells to form one number.  Here
  * replay mode walks all start nodes in order on the reference's uniform stream (K2) and K3 sums
    the sparse events in the reference's own (t, k) fp64 order -> rwSol bit-identical;
  * philox mode runs K1 once with the per-start score tally; the dense tempdata never exists.
"""
from __future__ import annotations

import logging
from dataclasses import dataclass, field

import numpy as np

from . import dist as _dist
from . import mesh as _mesh
from . import tables as _tables
from .dropin import device_tables, exp_time_weights, resolve_rng
from .engine import COUNTER_NAMES as _COUNTER_NAMES, NRW_N_COUNTERS, get_engine

rw_logger = logging.getLogger("NRW")

MAX_FIXED_BITS = 32            # most fractional bits of the time-weighted / multi-class fixed-point score


@dataclass
class FluenceResult:
    rwSol: np.ndarray                      # [(NS-1), (ND-1)] in the reference's index layout (Q2)
    per_start: np.ndarray                  # [meshSize] value for start node i (before the Q2 mapping)
    walker_steps: int                      # uniform draws == walker-steps (the throughput unit)
    rng: str
    stderr: np.ndarray | None = None       # [meshSize] standard error of per_start (philox mode)
    counters: dict = field(default_factory=dict)
    kernel: str = ""
    score: np.ndarray | None = None        # int64 [meshSize]: the integer tallies behind per_start (philox mode)
    split: dict = field(default_factory=dict)   # seconds: tables (build + upload, 0 when cached), device (launch .. host copy)


def scatter_to_rwsol(per_start: np.ndarray, NS: int, ND: int) -> np.ndarray:
    """Write per-start values into rwSol with the reference's index map (fluence_mini_app.py:415-419,
    :438) in start order: rwSol[0, ND-2] is written twice (i = 0, then i = ND-1 wins) and
    rwSol[NS-2, ND-2] is never written."""
    nd1 = ND - 1
    i = np.arange(len(per_start))
    aind = i % nd1
    aind = np.where(aind == 0, nd1, aind) - 1
    posind = ((i - aind) / nd1).astype(np.int64)          # int() truncates toward zero
    rwSol = np.zeros((NS - 1, nd1))
    rwSol[posind, aind] = per_start                       # later i overwrite earlier: numpy keeps the last
    return rwSol


_fused_cache: dict = {}


def _fused_solver(eng, tabs, starts, M, lo_m, hi_m, nt, sampler, cv):
    """solve(seed) -> pinned int64[n_starts + NRW_N_COUNTERS] host view (scores, then counters) of one fused
    fluence walk of walkers [lo_m, hi_m) of every start, summed over ranks.  Built once per (tables, M, slice,
    nt, weights): the start list, two launch descriptors (the tally buffers alternate, nrw.dist.TallyExchange)
    and the pinned mirror are reused.  One solve = ONE kernel (the walk sums over ranks and zeroes the next
    buffer in its own tail) + one copy into pinned memory."""
    import torch
    key = (id(tabs), M, lo_m, hi_m, int(nt), sampler, tuple(float(x) for x in cv))
    hit = _fused_cache.get(key)
    if hit is not None and hit.tabs is tabs:
        return hit
    if len(_fused_cache) > 16:
        _fused_cache.clear()
    n = len(starts)
    ex = _dist.TallyExchange(n, eng)
    pinned = torch.empty(n + NRW_N_COUNTERS + 1, dtype=torch.int64).pin_memory()
    host = pinned.numpy()
    counts = np.full(n, hi_m - lo_m, dtype=np.int64)
    id_base = starts.astype(np.int64) * M + lo_m             # the unsharded stream ids (start * M + m)
    pws, res = [], None
    for h in (0, 1):
        pw, res = eng.prepare_walk(tabs, starts, counts, nt, 0, sampler=sampler, score=ex.scores(h), class_value=cv,
                                   start_id_base=id_base, counters=ex.counters(h))
        pws.append(pw)
    peers = ex.world > 1 and ex.pair.p2p

    def solve(seed):
        for pw in pws:
            pw.set_seed(seed)
        out = ex.launch(pws)
        pinned[:n + NRW_N_COUNTERS].copy_(out, non_blocking=True)
        if peers:
            pinned[n + NRW_N_COUNTERS:].copy_(ex.status_view(), non_blocking=True)
        torch.cuda.current_stream(eng.device).synchronize()
        if peers and host[n + NRW_N_COUNTERS] != 0:
            raise RuntimeError("nrw: a peer rank did not reach the tally exchange within 5 s; the sums are invalid")
        return host[:n + NRW_N_COUNTERS]
    solve.tabs, solve.kernel, solve.keepalive, solve.exchange = tabs, res.kernel, (pws, res, ex, pinned), ex
    _fused_cache[key] = solve
    return solve


def source_classes(csr, source):
    """(source value per node, 2-bit class per row, value per class) of a fluence mesh, cached on the CSR: the
    source term is a Python call per node (fluence_mini_app.py:30-41)."""
    cache = csr.meta.setdefault("_source_cache", {})
    if source not in cache:
        sv = _mesh.source_values(csr.meta["mesh"]["meshpoints"], source)
        try:
            cache[source] = (sv,) + tuple(_tables.classify_source(sv, csr.n_rows))
        except ValueError:                                    # > 3 distinct values: replay mode can still run
            cache[source] = (sv, None, None)
    return cache[source]


def fluence_tables(csr, source):
    """(engine, device tables, source values, node classes, class values) of the Philox-mode fluence solve;
    uploaded once per (matrix, source)."""
    srcvals, node_class, class_value = source_classes(csr, source)
    if node_class is None:
        _tables.classify_source(srcvals, csr.n_rows)          # raises the explanatory ValueError
    eng, tabs = device_tables(csr, node_class=node_class, key=(id(csr), "fluence", id(node_class)))
    return eng, tabs, srcvals, node_class, class_value


def solve_fluence(csr: _mesh.TransitionCSR, M: int, nt: int, SigA, dt, v, source=_mesh.source_default,
                  rng=None, sampler="alias", seed=None, want_stderr=False) -> FluenceResult:
    """rwSol for every start node of a fluence mesh built by ``mesh.build_fluence_csr``."""
    m = csr.meta["mesh"]
    meshSize, NS, ND = m["meshSize"], m["NS"], m["ND"]
    srcvals, node_class, class_value = source_classes(csr, source)
    mode = resolve_rng(rng, meshSize * int(M))
    rank, world = _dist.world()
    starts = np.arange(meshSize)
    if mode == "replay":
        # one shared stream: the walk is sequential ("replicas only").  Under torchrun rank 0
        # replays and every other rank receives the result, so that all ranks return the same rwSol
        if rank == 0:
            eng, tabs = device_tables(csr)
            rep = eng.replay(tabs, starts, M, nt)
            tw = exp_time_weights(SigA, dt, nt)
            out = eng.assemble_events(rep, tw, srcvals, v, dt, M, nt, meshSize)
            packed = np.concatenate([out.cpu().numpy(), [float(rep.draws_used)]])
        else:
            rw_logger.info("replay mode is sequential (one shared stream): rank %d waits for rank 0", rank)
            packed = np.zeros(meshSize + 1)
        packed = _dist.broadcast_array(packed)
        per_start, draws = packed[:meshSize].copy(), int(packed[meshSize])
        return FluenceResult(scatter_to_rwsol(per_start, NS, ND), per_start, draws, "replay")

    # ---- philox: one K1 launch, per-start score in registers -------------------------------
    import time
    t0 = time.perf_counter()
    eng, tabs, _, _, _ = fluence_tables(csr, source)
    t_tables = time.perf_counter() - t0
    if seed is None:
        from .dropin import philox_seed_from_global_stream
        seed = philox_seed_from_global_stream()
    seed = _dist.broadcast_int(seed)                      # one seed for all ranks
    n_classes = int((class_value != 0).sum())
    uniform_time = (float(SigA) == 0.0) and n_classes <= 1
    if uniform_time:
        cv = (class_value != 0).astype(np.float64)        # count hits; scale by S0 afterwards
        unit = float(class_value[class_value != 0][0]) if n_classes else 0.0
        tw = None
    else:
        # fixed-point score: llrint(w(t) * S * 2^bits) summed into one int64 per start.  w(t) <= 1, so the sum is
        # below M * (nt - 2) * max(S) * 2^bits: take the most fractional bits (<= 32) that keep it under 2^62
        bound = float(M) * max(int(nt) - 2, 1) * float(np.abs(class_value).max())
        bits = MAX_FIXED_BITS
        while bits > 0 and bound * 2.0 ** bits >= 2.0 ** 62:
            bits -= 1
        if bits < 16:
            raise ValueError("M * nt * max(source) = %.3g leaves %d fractional bits for the 64-bit score tally; "
                             "walk fewer steps or walkers per solve" % (bound, bits))
        cv = class_value * 2.0 ** bits
        unit = 2.0 ** -bits
        tw = exp_time_weights(SigA, dt, nt)
    # every rank walks the same slice of EVERY start's walkers (balanced: starts differ in lifetime)
    lo_m, hi_m = _dist.shard_range(int(M), rank, world)
    if tw is None and not want_stderr:
        # the common solve (count hits, no per-walker variance): one prepared launch per (tables, M, nt),
        # scores and counters in ONE device buffer -- one fill, one launch, one collective, one copy back
        t0 = time.perf_counter()
        fs = _fused_solver(eng, tabs, starts, int(M), lo_m, hi_m, nt, sampler, cv)
        t_prep = time.perf_counter() - t0
        t0 = time.perf_counter()
        host = fs(seed)
        t_dev = time.perf_counter() - t0
        score = host[:meshSize].astype(np.float64)
        c = {k: int(x) for k, x in zip(_COUNTER_NAMES, host[meshSize:].view(np.uint64))}
        per_start = score * (unit * v * dt) / M
        return FluenceResult(scatter_to_rwsol(per_start, NS, ND), per_start, c["steps"], "philox",
                             counters=c, kernel=fs.kernel, score=host[:meshSize].copy(),
                             split=dict(tables_s=t_tables, launch_descriptors_s=t_prep,
                                        walk_collective_copy_back_s=t_dev))
    counts = np.full(meshSize, hi_m - lo_m, dtype=np.int64)
    id_base = starts.astype(np.int64) * int(M) + lo_m
    res = eng.walk(tabs, starts, counts, nt, seed, sampler=sampler, score=True,
                   score_sq=want_stderr, class_value=cv, time_weight=tw, start_id_base=id_base)
    _dist.allreduce_sum_([res.score, res.score_sq, res.counters])
    score_i = res.score.cpu().numpy()
    score = score_i.astype(np.float64)
    scale = unit * v * dt
    per_start = score * scale / M
    stderr = None
    if want_stderr:
        sq = res.score_sq.cpu().numpy()
        var = np.maximum(sq / M - (score / M) ** 2, 0.0) * scale * scale
        stderr = np.sqrt(var / M)
    c = res.counters_dict()
    return FluenceResult(scatter_to_rwsol(per_start, NS, ND), per_start, c["steps"], "philox",
                         stderr=stderr, counters=c, kernel=res.kernel, score=score_i)
