#!/usr/bin/env python
"""Side measurements of the BASELINE configs that are not the bench.py headline (run on a B200):
C3 simple_flux (SigA>0, --use_sinks) at M walkers per state, C4 refined fluence mesh through the
fused solve.  Prints one JSON object per config.  Usage: python tools/side_benches.py [c3] [c4]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "neural-random-walker_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def c3(M=100_000_000):
    import torch
    import simple_flux_miniapp as app
    from nrw import dropin, mesh
    from nrw.engine import get_engine
    q = mesh.build_simple_flux_dense(0.005, 8.0, 2.0)
    csr = mesh.csr_from_dense(q, n_live=2)
    eng, tabs = dropin.device_tables(csr)
    nt = 1002
    out = {}
    for reps in (1, 3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for r in range(reps):
            res = eng.walk(tabs, [0, 1], [M, M], nt, 100 + r, count_tn=True, t_cap=nt, count_per_start=True)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / reps
    c = res.counters_dict()
    cnt = res.count_tn.cpu().numpy().astype(np.float64)
    sol = (5.0 * cnt[:, :, 0] + 3.0 * cnt[:, :, 1]) / M            # rwSol[i, t] (simple_flux:327)
    t = np.arange(nt) * 0.005
    e1, e2 = np.exp(-2.0 * t), np.exp(-10.0 * t)
    true = np.stack([2.5 * (e1 + e2) + 1.5 * (e1 - e2), 2.5 * (e1 - e2) + 1.5 * (e1 + e2)])
    err = float(np.abs(sol[:, 1:1000] - true[:, 1:1000]).max())
    out = {"config": "C3 simple_flux -sa 2 -ss 8 -dt 0.005 --use_sinks, M=%d per state, nt=%d" % (M, nt),
           "kernel": res.kernel, "walker_steps": c["steps"], "seconds": dt,
           "walker_steps_per_s": c["steps"] / dt, "max_abs_err_vs_closed_form": err,
           "exits": c["exits"], "survivors": c["survivors"]}
    print(json.dumps(out), flush=True)


def c4(M=100):
    import torch
    from nrw import mesh
    from nrw.fluence import solve_fluence
    t0 = time.perf_counter()
    csr = mesh.build_fluence_csr(0.001, 1.0 / 63.0, 200, 0.15, 1)
    t_build = time.perf_counter() - t0
    t0 = time.perf_counter()
    r = solve_fluence(csr, M, 30000, 0, 0.001, 200, rng="philox", seed=1)      # includes table build + upload
    torch.cuda.synchronize()
    t_first = time.perf_counter() - t0
    reps = 5
    t0 = time.perf_counter()
    for i in range(reps):
        r = solve_fluence(csr, M, 30000, 0, 0.001, 200, rng="philox", seed=2 + i)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    out = {"config": "C4 refined fluence mesh dt=0.001 da=1/63: %d nodes nnz=%d, M=%d" % (csr.n_live, csr.nnz, M),
           "kernel": r.kernel, "walker_steps": r.walker_steps, "csr_build_s": t_build,
           "first_solve_s_incl_alias_tables": t_first, "seconds_per_solve_host_to_host": dt,
           "walker_steps_per_s": r.walker_steps / dt, "steps_per_walker": r.walker_steps / (csr.n_live * M),
           "rwSol_shape": list(r.rwSol.shape), "rwSol_sum": float(r.rwSol.sum())}
    print(json.dumps(out), flush=True)
    for M2 in (1000,):
        t0 = time.perf_counter()
        r = solve_fluence(csr, M2, 30000, 0, 0.001, 200, rng="philox", seed=11)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print(json.dumps({"config": "C4 M=%d" % M2, "walker_steps": r.walker_steps, "seconds": dt,
                          "walker_steps_per_s": r.walker_steps / dt, "kernel": r.kernel}), flush=True)


def c4d(M=20_000_000, nt=40):
    """BASELINE configs[3]: the refined mesh through the densityrw surface (time-resolved occupancy)."""
    import torch
    from densityrw.DensityRW import CSRTransitions, DensityRW
    from nrw import dropin, mesh
    csr = mesh.build_fluence_csr(0.001, 1.0 / 63.0, 200, 0.15, 1)
    starts = [79443, 40000, 120000, 1000]
    walk = DensityRW(nt, CSRTransitions(csr), init_walkers={(s,): M // len(starts) for s in starts})
    np.random.seed(3)
    walk.run(rng="philox")                                            # tables, first launch
    t0 = time.perf_counter()
    counts = walk.run(rng="philox")
    t_run = time.perf_counter() - t0
    eng, tabs = dropin.device_tables(walk.csr)
    res = None
    for rep in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = eng.walk(tabs, starts, [M // len(starts)] * len(starts), nt, 50 + rep, count_tn=True, t_cap=nt,
                       count_path=True)
        torch.cuda.synchronize()
        t_walk = time.perf_counter() - t0
    c = res.counters_dict()
    print(json.dumps({"config": "C4 through densityrw.DensityRW: %d nodes, %d walkers from %d nodes, %d steps"
                                % (csr.n_rows, M, len(starts), nt), "kernel": res.kernel,
                      "walker_steps": c["steps"], "walk_seconds_device_resident": t_walk,
                      "walker_steps_per_s": c["steps"] / t_walk, "DensityRW_run_seconds_host_to_host": t_run,
                      "counts_shape": list(counts.shape), "counts_sum": float(counts.sum())}), flush=True)


if __name__ == "__main__":
    which = sys.argv[1:] or ["c3", "c4", "c4d"]
    if "c4d" in which:
        c4d()
    if "c3" in which:
        c3()
    if "c4" in which:
        c4()
