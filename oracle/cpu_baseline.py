"""Timed CPU legs for bench.py (cpu_baseline / --impl reference) -- TEST INFRASTRUCTURE.

kind "port": /root/reference is pure Python and cannot travel to the GPU box, so what is timed is
the line-for-line restatement of the reference's hot loop (restated.run_pc_literal ==
fluence_mini_app.py:224-247: np.cumsum of the dense matrix, then the triple loop with one
np.random.uniform() + one np.searchsorted per walker-step) followed by the reference's assembly
loop (:428-434), per start node, exactly as the reference's run() does.  The reference's 0.53 s
CSV re-parse per call (:223) is NOT included -- that favours the reference.  The reference is
single-threaded, so cores = 1.
"""
from __future__ import annotations

import time

import numpy as np

from . import restated as R


def literal_sample(dense_q, mesh, M, nt, start_nodes, budget_s, SigA=0.0, dt=0.01, v=200):
    """Run the literal port for the given start nodes until the budget is spent.
    Returns dict(walker_steps, seconds, starts_done)."""
    meshSize, meshpoints = mesh["meshSize"], mesh["meshpoints"]
    steps, done = 0, 0
    t0 = time.perf_counter()
    for i in start_nodes:
        td = R.run_pc_literal(dense_q, nt, [[int(i), int(M)]], meshSize)
        R.assemble_start_literal(td, meshSize, meshpoints, SigA, dt, v, M)
        steps += int(td[:meshSize, :nt - 1].sum())
        done += 1
        if time.perf_counter() - t0 > budget_s:
            break
    return dict(walker_steps=steps, seconds=time.perf_counter() - t0, starts_done=done)


def vectorised_sample(csr_arrays, meshSize, M, nt, start_nodes, budget_s):
    """The count-based numpy restatement (NOT the reference's algorithmic cost; reported beside it)."""
    indptr, cols, cum = csr_arrays
    steps, done = 0, 0
    t0 = time.perf_counter()
    for i in start_nodes:
        _, used = R.run_pc_counts(indptr, cols, cum, nt, [[int(i), int(M)]], meshSize, dense=False)
        steps += used
        done += 1
        if time.perf_counter() - t0 > budget_s:
            break
    return dict(walker_steps=steps, seconds=time.perf_counter() - t0, starts_done=done)
