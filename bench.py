#!/usr/bin/env python
"""bench.py -- walker-steps/s of the fused fluence solve on N B200s (one process per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path over one batch: the fused fluence solve (K1s walk_score, per-start score
tally) of every start node of the BASELINE C5 mesh (dt=0.01, da=0.02: 100 x 100 = 10 000 nodes, nnz 505 000)
with `--walkers-per-node-per-gpu` walkers per node on every GPU (default 12 500, so that N=8 is the 1e9-walker
mesh of BASELINE.json configs[4]; weak scaling) -- or `--total-walkers-per-node` split over the ranks (strong
scaling; 100000 = the 1e9-walker mesh on any N).  On N > 1 the step includes the path's one collective, the sum
of the per-start scores over ranks, which the walk kernel runs in its own tail over NVLink peer memory
(csrc/walk_tail.cuh): ONE kernel launch per step on any N.  One walker-step = one uniform draw + one transition
look-up + one tally (fluence_mini_app.py:243-245 of the reference).  DESIGN.md section 5 defines every JSON key.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "neural-random-walker_b200")
for _p in (PKG, ROOT):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "walker-steps/s"
C5 = dict(dt=0.01, da=0.02, v=200, SigS=0.15, L=1)
C4 = dict(dt=0.001, da=1.0 / 63.0, v=200, SigS=0.15, L=1)
C1 = dict(dt=0.01, da=1.0 / 15.0, v=200, SigS=0.15, L=1)
NT = 30000                      # the script default (-nt): walkers run until they exit
NT_EXPECT = 400                 # horizon of the exact chain propagation: P(alive at t = 400) < 1e-100 on C5 / C1


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--walkers-per-node-per-gpu", type=int, default=12500)
    ap.add_argument("--total-walkers-per-node", type=int, default=0,
                    help="strong scaling: this many walkers per node in total, split over the ranks "
                         "(100000 = the 1e9-walker mesh of BASELINE configs[4] on any number of GPUs)")
    ap.add_argument("--tally", default="score", choices=["score", "score+hist"],
                    help="score: per-start fluence score (rwSol); score+hist: also the shared-memory "
                         "occupancy histogram over nodes")
    ap.add_argument("--no-extras", action="store_true", help="skip the side legs (C2 / C3 / C4 / cpu_baseline / run(args))")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="budget of the cpu_baseline sample")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# helpers
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock + throttle reasons during the timed region (NVML, 5 ms period)."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap", 0x80: "hw_power_brake", 0x2: "applications_clocks_setting"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(float(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                mask = int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.005)

    def __enter__(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def algorithmic_bytes_per_step(csr, steps_from_node):
    """B(r) = 4 (state read) + 4 (state write) + 8 (row extent) +
    8*ceil(log2 nnz(r)) (fp64 cumulative probes of the reference's searchsorted; 0 if nnz=1) +
    4 (column id) + 4 (tally); averaged over the walker-steps actually taken from each row."""
    nnz = csr.row_lengths()[:csr.n_live].astype(np.float64)
    probes = np.where(nnz > 1, np.ceil(np.log2(np.maximum(nnz, 1))), 0.0)
    b = 24.0 + 8.0 * probes
    w = steps_from_node[:csr.n_live].astype(np.float64)
    return float((b * w).sum() / max(w.sum(), 1.0))


def time_kernel(torch, launch, flush, reps=10):
    """Mean CUDA-event time (ms) of the prepared walk's kernel alone: L2 flushed and tallies zeroed outside the events."""
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        flush.fill_(1)
        launch.zero()
        a.record()
        launch.launch_kernel()
        b.record()
    torch.cuda.synchronize()
    return float(np.mean([a.elapsed_time(b) for a, b in ev]))


# ------------------------------------------------------------------------------------------------
# CPU legs (the reference itself when it is mounted, else the oracle's literal port of its loop; the only
# places bench.py executes oracle/)
# ------------------------------------------------------------------------------------------------
_cpu_cache = {}
CPU_M, CPU_NT = 100, 64         # the reference's default M; nt = 64: every walker is dead by t ~ 50


def _cpu_setup():
    if not _cpu_cache:
        from nrw import mesh, tables
        csr = mesh.build_fluence_csr(**C5)
        _cpu_cache.update(csr=csr, m=csr.meta["mesh"], q=csr.to_dense(),   # the reference walks the dense matrix
                          cum=tables.row_cum(csr), ref=_load_reference())
    return _cpu_cache


def _load_reference():
    """The reference's own fluence_mini_app module if a copy is reachable: $NRW_REFERENCE_ROOT, baseline/_ref,
    /root/reference (the build container).  None on the GPU box, where only this repository exists."""
    for root in (os.environ.get("NRW_REFERENCE_ROOT"), os.path.join(ROOT, "baseline", "_ref"), "/root/reference"):
        if root and os.path.isfile(os.path.join(root, "fluence_mini_app.py")):
            try:
                from oracle import ref_import
                ref_import.REFERENCE_ROOT = root
                return ref_import.fluence(), root
            except Exception as exc:                      # a broken copy: say so and time the port
                sys.stderr.write("bench: reference at %s not importable (%s); timing the port\n" % (root, exc))
    return None


def _reference_start_node(c, start, seed):
    """One start node through the UNMODIFIED reference: its run_pc (fluence_mini_app.py:218-247) and its own
    assembly arithmetic (:428-434).  run_pc re-parses the matrix CSV on every call (:223; 2.5 GB of text for this
    mesh): np.genfromtxt is answered from memory instead, which favours the reference."""
    ref = c["ref"][0]
    m, M, nt = c["m"], CPU_M, CPU_NT
    real = np.genfromtxt
    np.genfromtxt = lambda *a, **k: c["q"]
    try:
        np.random.seed(seed)
        t0 = time.perf_counter()
        td = ref.run_pc(matrix_filename="curmat.csv", neuraltime=nt, initial_walkers=[[int(start), M]],
                        meshSize=m["meshSize"], runtime_args={})
        tempval = 0
        for j in range(1, td.shape[1] - 1):                                     # :428-434
            for k in range(0, m["meshSize"]):
                tempval = tempval + td[k, j] * np.exp(-0.0 * 0.01 * j) * 200 * ref.sourceTerm(
                    m["meshpoints"][k, 0], m["meshpoints"][k, 1]) * 0.01
        dt = time.perf_counter() - t0
    finally:
        np.genfromtxt = real
    return int(td[:m["meshSize"], :nt - 1].sum()), dt


def cpu_start_node(start, seed):
    """(walker-steps, seconds) of one start node (M = 100, nt = 64) on one core: the reference, else the port."""
    c = _cpu_setup()
    if c["ref"] is not None:
        return _reference_start_node(c, start, seed)
    from oracle import cpu_baseline
    np.random.seed(seed)
    lit = cpu_baseline.literal_sample(c["q"], c["m"], CPU_M, CPU_NT, [start], 1e9)
    return lit["walker_steps"], lit["seconds"]


def cpu_kind():
    c = _cpu_setup()
    return ("reference", "the unmodified reference at %s (run_pc + its assembly arithmetic; CSV parse answered "
            "from memory)" % c["ref"][1]) if c["ref"] is not None else \
           ("port", "oracle/restated.py's literal port of run_pc + assembly (fluence_mini_app.py:224-247, :428-434; "
            "the Python reference cannot travel to the GPU box)")


def cpu_leg(budget_s, seed=1):
    """cpu_baseline of the main line: the reference algorithm as it is -- single-threaded -- on random start nodes
    of the bench mesh at the reference's M = 100, nt = 64, bounded by budget_s."""
    c = _cpu_setup()
    starts = np.random.default_rng(seed).permutation(c["m"]["meshSize"])
    steps, secs, done = 0, 0.0, 0
    for s in starts:
        a, b = cpu_start_node(int(s), seed + done)
        steps, secs, done = steps + a, secs + b, done + 1
        if secs > budget_s:
            break
    kind, what = cpu_kind()
    return dict(value=steps / max(secs, 1e-9), unit=METRIC, cores=1, kind=kind,
                sample="%s on the C5 mesh (10 000 nodes, dense 10001^2 matrix), M=%d, nt=%d, %d random start nodes = "
                       "%d walker-steps in %.1f s on 1 core; the reference's per-call CSV re-parse is excluded"
                       % (what, CPU_M, CPU_NT, done, steps, secs), host_cores=os.cpu_count())


def cpu_vectorised_leg(budget_s, seed=1):
    from oracle import cpu_baseline
    c = _cpu_setup()
    csr = c["csr"]
    starts = np.random.default_rng(seed).permutation(c["m"]["meshSize"])
    np.random.seed(seed)
    vec = cpu_baseline.vectorised_sample((csr.indptr, csr.cols.astype(np.int64), c["cum"]),
                                         c["m"]["meshSize"], CPU_M, CPU_NT, starts, budget_s)
    return vec["walker_steps"] / max(vec["seconds"], 1e-9)


def _ref_worker(task):
    """One start node in a forked worker: the dense matrix is inherited copy-on-write; the per-call np.cumsum of
    it (fluence_mini_app.py:224) is the worker's own."""
    seed, = task
    start = int(np.random.default_rng(seed).integers(0, 10000))
    return cpu_start_node(start, seed)


def run_reference_arm(args):
    """--impl reference: the reference's CPU implementation of the path on the box's host cores.  The reference is
    single-threaded Python; its loop over start nodes (:397-407) is independent per node, so to give the host its
    best case a step here walks one start node (M=100 walkers) of the C5 mesh on EVERY host core at once (forked
    workers; bounded by memory: each worker builds its own 800 MB cumulative matrix).  `config` describes what
    THIS arm ran (the reference's own M and RNG), not the GPU arm's workload."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import multiprocessing as mp
    _cpu_setup()
    workers = max(1, min(os.cpu_count() or 1, 16))
    try:
        import psutil
        workers = max(1, min(workers, int(psutil.virtual_memory().available // (3 << 30))))
    except Exception:
        workers = min(workers, 4)
    steps_total, secs_total = 0, 0.0
    t_all = time.perf_counter()
    pool = mp.get_context("fork").Pool(workers) if workers > 1 else None
    try:
        for i in range(args.warmup + args.steps):
            tasks = [(1 + i * workers + k,) for k in range(workers)]
            t0 = time.perf_counter()
            out = pool.map(_ref_worker, tasks) if pool else [_ref_worker(t) for t in tasks]
            dt = time.perf_counter() - t0
            if i >= args.warmup:
                steps_total += sum(o[0] for o in out)
                secs_total += dt
    finally:
        if pool:
            pool.close()
            pool.join()
    value = steps_total / max(secs_total, 1e-9)
    kind, what = cpu_kind()
    sample = ("%s on the C5 mesh (10 000 nodes, dense 10001^2 matrix), M=%d, nt=%d: %d steps x %d start nodes walked "
              "concurrently on %d host cores = %d walker-steps in %.1f s wall; the reference's per-call CSV re-parse "
              "is excluded" % (what, CPU_M, CPU_NT, args.steps, workers, workers, steps_total, secs_total))
    config = {"workload": "fluence C5 mesh: dt=0.01 da=0.02 v=200 SigS=0.15 SigA=0 L=1 -> 100x100=10000 start nodes, "
                          "nnz=505000 (dense 10001^2 fp64 matrix, as the reference holds it); per step %d randomly "
                          "chosen start nodes x M=%d walkers, nt=%d, rng = numpy legacy MT19937 (one uniform per "
                          "walker-step), sampler = np.searchsorted on np.cumsum rows, default source; %d forked "
                          "single-threaded workers" % (workers, CPU_M, CPU_NT, workers),
              "same_mesh_as_gpu_arm": True, "walkers_per_node": CPU_M, "nt": CPU_NT, "workers": workers,
              "note": "a bounded sample of the GPU arm's mesh at the reference's own M and RNG: the reference needs "
                      "~10 s per start node on this mesh, the GPU arm's 1.25e8-walker step would take it years"}
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": METRIC,
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * secs_total / max(args.steps, 1),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": config,
            "cpu_baseline": dict(value=value, unit=METRIC, cores=workers, kind=kind, sample=sample,
                                 host_cores=os.cpu_count()),
            "e2e": {"value": value, "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": time.perf_counter() - t_all}
    _emit(json.dumps(line))


def workload_config(args, world, Mg, strong):
    return {"workload": "fluence C5 mesh: dt=0.01 da=0.02 v=200 SigS=0.15 SigA=0 L=1 -> 100x100=10000 "
                        "start nodes, nnz=505000, %d walkers/node/GPU x %d GPU(s) = %.3g walkers per step%s, "
                        "nt=30000, default source, rng=philox, sampler=alias"
                        % (Mg, world, 1e4 * Mg * world,
                           " (the 1e9-walker mesh of BASELINE configs[4])" if Mg * world == 100000 else
                           " (x8 GPUs = the 1e9-walker mesh of BASELINE configs[4])" if Mg == 12500 and not strong else ""),
            "walkers_per_node_per_gpu": Mg, "n_nodes": 10000,
            "tally": args.tally, "parallelism": "walker-id sharding; the one all-reduce of the score buffer runs in the "
                                                "walk kernel's tail over NVLink peer memory (one launch per step)",
            "l2": "tables are shared-memory resident by design; a 512 MiB buffer is overwritten "
                  "between timed steps (untimed) to flush L2"}


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as tdist
    from nrw import dist as ndist
    from nrw import mesh, tables
    from nrw.engine import Engine, get_engine

    rank, world = ndist.init_from_env()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    eng = get_engine(local)

    csr = mesh.build_fluence_csr_cached(**C5)
    m = csr.meta["mesh"]
    n_nodes = m["meshSize"]
    src = mesh.source_values(m["meshpoints"], mesh.source_default)
    node_class, class_value = tables.classify_source(src, csr.n_rows)
    tabs = eng.upload(csr, node_class=node_class)
    strong = args.total_walkers_per_node > 0
    Mg = args.total_walkers_per_node // world if strong else args.walkers_per_node_per_gpu
    M_total = Mg * world
    starts = np.arange(n_nodes)
    id_base = starts.astype(np.int64) * M_total + rank * Mg      # rank's slice of every start's walkers
    want_hist = args.tally == "score+hist"
    # scores + counters in one double-buffered tally buffer; the walk's tail sums it over ranks (N > 1) and zeroes
    # the next step's half: one kernel per step
    ex = ndist.TallyExchange(n_nodes, eng)
    pws, hists = [], []
    for h in (0, 1):
        pw, r = eng.prepare_walk(tabs, starts, Mg, NT, seed=1, class_value=[0, 1, 0, 0], node_hist=want_hist,
                                 start_id_base=id_base, score=ex.scores(h), counters=ex.counters(h))
        pws.append(pw)
        hists.append(r.node_hist)
    kernel_name, geometry = r.kernel, list(r.launch_geometry)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)

    def step():
        if want_hist:
            hists[ex.epoch & 1].zero_()
        return ex.launch(pws)                                     # walk + (N > 1) the single collective, one kernel

    def barrier():
        if world > 1:
            tdist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        out = step()
    barrier()
    job_steps = int(out[n_nodes].item())                          # walker-steps per step, all ranks (reduced counter)
    steps_per_launch = int(ex.counters((ex.epoch - 1) & 1)[0].item())     # this rank's share (identical every step)

    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
           for _ in range(args.steps)]
    with ClockSampler(local) as clk:
        barrier()
        t_wall = time.perf_counter()
        for a, b in evs:
            flush.fill_(1)                                        # L2 flush, outside the events
            a.record()
            out = step()
            b.record()
        barrier()
        t_wall = time.perf_counter() - t_wall
    ms = np.array([a.elapsed_time(b) for a, b in evs])
    reduced = out.clone()
    status = int(ex.status_view().item())
    assert status == 0, "a peer rank did not reach the tally exchange"
    collective = None
    if world > 1:
        # the last exchange against the backend's all_reduce on the same per-rank vector (integer sum: identical)
        chk = ex.pair.half((ex.epoch - 1) & 1).clone()
        tdist.all_reduce(chk, op=tdist.ReduceOp.SUM)
        assert torch.equal(chk, reduced), "the walk's fused exchange differs from dist.all_reduce"
        collective = ("fused into the walk kernel's tail over NVLink peer memory (csrc/walk_tail.cuh)" if ex.pair.p2p
                      else "NCCL all_reduce after the walk (no peer memory)")
    total_ms = torch.tensor([float(ms.sum())], dtype=torch.float64, device=dev)
    if world > 1:
        tdist.all_reduce(total_ms, op=tdist.ReduceOp.MAX)
    total_ms = float(total_ms.item())
    value = job_steps * args.steps / (total_ms * 1e-3)

    # ---- dominant kernel alone (own score buffer, no exchange): the roofline's `achieved` ----------------------
    launch, res = eng.prepare_walk(tabs, starts, Mg, NT, seed=1, class_value=[0, 1, 0, 0], node_hist=want_hist,
                                   start_id_base=id_base, score=True)
    for _ in range(3):
        launch()
    kernel_ms = time_kernel(torch, launch, flush)
    assert int(res.counters.cpu()[0].item()) == steps_per_launch
    collective_us = 1e3 * (float(np.median(ms)) - kernel_ms) if world > 1 else None

    # ---- ablation (kernel alone, same launch): draws on deterministic rows too / the general kernel.  The library
    # reads its switches once per engine, so each variant gets its own engine.
    ablation = {}
    if rank == 0 and not args.no_extras:
        for tag, sw in (("always_draw_NRW_DET=0", {"NRW_DET": "0"}), ("general_kernel_NRW_K1=general", {"NRW_K1": "general"})):
            e2 = Engine(local, switches=sw)
            l2, r2 = e2.prepare_walk(tabs, starts, Mg, NT, seed=1, class_value=[0, 1, 0, 0], node_hist=want_hist,
                                     start_id_base=id_base, score=True)
            for _ in range(3):
                l2()
            ms_ab = time_kernel(torch, l2, flush)
            assert int(r2.counters.cpu()[0].item()) == steps_per_launch      # same walk, same count
            assert torch.equal(r2.score, res.score)                           # and the same tallies
            ablation[tag] = {"kernel": r2.kernel, "kernel_ms": ms_ab, "walker_steps_per_s": steps_per_launch / (ms_ab * 1e-3)}
            e2.close()

    line = None
    if rank == 0:
        # occupancy (walker-steps taken FROM each node): deterministic-step share and the algorithmic-bytes side note
        occ_launch, occ = eng.prepare_walk(tabs, starts, Mg, NT, seed=1, node_hist=True, start_id_base=id_base)
        occ_launch()
        arrivals = occ.node_hist.cpu().numpy().astype(np.float64)
        steps_from = arrivals[:n_nodes] + Mg
        steps_from[csr.row_lengths()[:n_nodes] == 0] = 0
        lens = csr.row_lengths()[:n_nodes]
        det_frac = float(steps_from[lens == 1].sum() / max(steps_from.sum(), 1.0))
        clocks = clk.summary()
        roofline = roofline_object(eng, csr, steps_from, res.kernel, kernel_ms, steps_per_launch, det_frac, clocks, Mg)
        always = ablation.get("always_draw_NRW_DET=0")
        line = {"metric": METRIC, "value": value, "unit": METRIC, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps,
                "ms_p50": float(np.percentile(ms, 50)), "ms_p95": float(np.percentile(ms, 95)),
                "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None,
                "dtype": "u32/i64", "data": "synthetic", "config": workload_config(args, world, Mg, strong),
                "value_always_draw": (always["walker_steps_per_s"] if always and world == 1 else None),
                "drawn_steps_per_s": value * (1.0 - det_frac),
                "headline_note": "value counts every walker-step of the reference's loop (fluence_mini_app.py:243-245); "
                                 "%.1f %% of them leave rows with a single absorbing successor and are taken without "
                                 "computing a draw (bit-identical tallies).  value_always_draw = the kernel alone "
                                 "drawing on every step (N = 1); drawn_steps_per_s = value x drawn share"
                                 % (100.0 * det_frac),
                "roofline": roofline,
                "walker_steps_per_step": job_steps, "walkers_per_step": n_nodes * Mg * world,
                "gpu_launches": args.steps, "launches_per_step": 1, "wall_s_timed_region": t_wall,
                "launch_geometry": geometry, "kernel": kernel_name, "ablation_kernel_only": ablation or None,
                "collective": collective, "collective_us": collective_us, "engine_config": eng.config,
                "clocks": clocks}
        line["fluence_err"] = fluence_err_leg(csr, src, reduced[:n_nodes].cpu().numpy(), M_total)

    # ---- N > 1: the drop-in's sharded solve against rank 0's unsharded solve of the same seed -------------------
    if world > 1:
        mrp = multi_rank_parity_leg(eng, tabs, csr, starts, M_total, rank)
        if rank == 0:
            line["multi_rank_parity"] = mrp

    # ---- e2e: the C-ABI host entry with HOST buffers: H2D of tables + start list, walk (+ exchange), D2H of the
    # (reduced) scores inside the timed region
    e2e = e2e_leg(eng, tabs, csr, starts, Mg, id_base, rank, world, dev, args, barrier)
    if rank == 0:
        line["e2e"] = e2e

    # ---- e2e through the reference's own entry point: fluence_mini_app.run(args) ---------------------------------
    if not args.no_extras:
        era = run_args_leg(rank, world, barrier)
        if rank == 0:
            line["e2e_run_args"] = era

    # ---- side legs: BASELINE configs[2] (C3, at N = 1 and 2), and on one GPU the resource rooflines, configs[1]
    # (C2) with its fluence error, configs[3] (C4) and the CPU baseline
    if not args.no_extras and world <= 2:
        c3 = c3_leg(eng, rank, world, flush)
        if rank == 0:
            line["c3"] = c3
    if rank == 0 and world == 1 and not args.no_extras:
        line["resources"] = resource_leg(eng, tabs, starts, Mg, id_base, flush, steps_per_launch, kernel_ms, det_frac, kernel=line["kernel"])
        line["roofline"]["mem"].update(
            smem_load_frac=line["resources"]["score_kernel"]["frac_of_smem_load_peak"],
            smem_atomic_frac_hist_mode=line["resources"]["hist_kernel"]["frac_of_smem_atomic_peak"])
        line["c2"] = c2_leg(eng)
        line["c4"] = c4_leg(eng, flush)
        cb = cpu_leg(args.cpu_seconds)
        cb["vectorised_port_value"] = cpu_vectorised_leg(3.0)
        line["cpu_baseline"] = cb
    elif rank == 0:
        line["cpu_baseline"] = None
    if rank == 0:
        _emit(json.dumps(line))
    if world > 1:
        tdist.barrier()
        tdist.destroy_process_group()


def roofline_object(eng, csr, steps_from, kernel, kernel_ms, steps_per_launch, det_frac, clocks, Mg):
    """The binding resource of the walk is the SM's instruction issue (integer ALU + IMAD pipes), not memory: tables
    live in shared memory, walker state in registers, HBM carries 0.4 MB per launch.  frac = warp-instructions per
    launch / (SMs x 4 schedulers x SM clock x kernel time): one warp-instruction per scheduler and cycle is the
    ceiling.  The instruction count is deterministic per (workload, seed) and comes from the committed ncu capture
    (profiles/walk_traffic.json, regenerated by tools/ncu_summary.py --traffic-json); a capture of another kernel
    or another workload is refused."""
    facts, stale = {}, None
    tp = os.path.join(ROOT, "profiles", "walk_traffic.json")
    if os.path.isfile(tp):
        try:
            facts = json.load(open(tp))
        except Exception as exc:
            stale = "unreadable: %s" % exc
    else:
        stale = "profiles/walk_traffic.json missing"
    # same kernel and the same workload: the walker-steps of the captured launch, or -- on a rank of an N > 1 job, which
    # walks another slice of the stream ids of the same workload -- within 0.1 % of it (the instruction count is then
    # scaled by the ratio of the walker-steps and the line says so)
    cap_steps = int(facts.get("walker_steps_per_launch", -1)) if facts else -1
    if facts and (facts.get("kernel") != kernel or abs(cap_steps - steps_per_launch) > 1e-3 * steps_per_launch):
        stale = ("stale capture: %s with %s walker-steps per launch, this run: %s with %d"
                 % (facts.get("kernel"), facts.get("walker_steps_per_launch"), kernel, steps_per_launch))
    if facts and not stale:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        from ncu_summary import csrc_sha256
        if facts.get("csrc_sha256") != csrc_sha256():
            stale = "stale capture: the CUDA sources changed since it was taken (csrc_sha256 differs)"
    sm_mhz = clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965.0
    peak = eng.n_sms * 4 * sm_mhz * 1e6                       # warp-instructions / s the schedulers can issue
    peak_hbm, peak_src = measured_peak()
    b_step = algorithmic_bytes_per_step(csr, steps_from)
    r = {"bound": "issue", "unit": "warp-inst/s", "peak": peak,
         "peak_source": "%d SMs x 4 warp schedulers x %.0f MHz (median SM clock sampled during the timed region)"
                        % (eng.n_sms, sm_mhz),
         "kernel": kernel, "kernel_ms": kernel_ms, "walker_steps_per_launch": steps_per_launch,
         "deterministic_step_fraction": det_frac}
    if stale:
        r.update(achieved=None, frac=None, traffic=None, stale=stale)
        dram = None
    else:
        wi = float(facts["warp_inst_per_launch"]) * steps_per_launch / cap_steps
        if cap_steps != steps_per_launch:
            r["instruction_count_scaled_from_capture_by"] = steps_per_launch / cap_steps
        dram = facts.get("dram_bytes_per_launch")
        drawn = steps_per_launch * (1.0 - det_frac)
        r.update(achieved=wi / (kernel_ms * 1e-3), frac=wi / (kernel_ms * 1e-3) / peak, traffic=dram,
                 warp_inst_per_launch=wi, lane_efficiency=facts.get("threads_per_warp_inst", 0.0) / 32.0,
                 warp_inst_per_32_walker_steps=32.0 * wi / steps_per_launch,
                 warp_inst_per_32_drawn_steps=32.0 * wi / drawn,
                 issue_active_pct_ncu=facts.get("issue_active_pct"), alu_pipe_pct_ncu=facts.get("alu_pipe_pct"),
                 fmaheavy_pipe_pct_ncu=facts.get("fmaheavy_pipe_pct"),
                 smem_bank_conflict_share_ncu=facts.get("smem_bank_conflict_share"), source=facts.get("source"))
    r["mem"] = {"dram_bytes": dram, "hbm_frac": (dram / (kernel_ms * 1e-3) / 1e9 / peak_hbm) if dram else None,
                "hbm_peak_gbs": peak_hbm, "hbm_peak_source": peak_src,
                "algorithmic_bytes_per_walker_step": b_step,
                "hbm_equivalent_of_algorithmic_bytes": steps_per_launch * b_step / (kernel_ms * 1e-3) / 1e9 / peak_hbm,
                "note": "algorithmic bytes are the reference algorithm's (SURVEY 8d: searchsorted probes on fp64 rows); "
                        "the kernel serves them from shared memory, so the HBM-equivalent figure exceeds 1 and is a "
                        "side note only.  smem_load_frac / smem_atomic_frac: against micro-benchmarked peaks (resources)"}
    return r


def fluence_err_leg(csr, src, score, M_total):
    """"fluence err vs pc ref" (BASELINE.json metric) ON THE HEADLINE WORKLOAD: the (reduced) per-start scores of the
    last timed step against the exact M -> infinity limit of the reference's chain on this mesh (oracle/expectation:
    checker only)."""
    from oracle import expectation
    mean, var = expectation.fluence_moments(csr.indptr, csr.cols, csr.vals, csr.n_rows, csr.n_live, src, NT_EXPECT,
                                            0.0, C5["dt"], C5["v"])
    unit = C5["v"] * C5["dt"] * 3.0
    per_start = score.astype(np.float64) * unit / M_total
    st = expectation.fluence_error_stats(per_start, mean, var, M_total, unit)
    st["against"] = ("exact expectation and per-walker variance of the reference's Markov chain "
                     "(fluence_mini_app.py:279-386 matrix through the sum of :428-438), all 10 000 cells")
    return st


def multi_rank_parity_leg(eng, tabs, csr, starts, M_total, rank):
    """One sharded solve through the drop-in's own solver (nrw.fluence.solve_fluence: every rank walks its slice of
    every start's walkers, the walk's tail sums over ranks) against rank 0's UNSHARDED walk of the same seed: the
    integer scores and counters must be identical."""
    import torch
    from nrw.fluence import solve_fluence
    seed = 4242
    res = solve_fluence(csr, M_total, NT, 0, C5["dt"], C5["v"], rng="philox", seed=seed)
    out = None
    if rank == 0:
        one = eng.walk(tabs, starts, M_total, NT, seed, score=True, class_value=[0, 1, 0, 0],
                       start_id_base=starts.astype(np.int64) * M_total)
        score = one.score.cpu().numpy()
        c = one.counters_dict()
        same = bool(np.array_equal(score, res.score)) and all(res.counters[k] == c[k] for k in ("steps", "exits", "walkers"))
        out = {"sharded_equals_unsharded": same, "walkers": c["walkers"], "walker_steps": c["steps"],
               "score_sum": int(score.sum()), "kernel": res.kernel,
               "how": "solve_fluence under torchrun (fused exchange) vs one unsharded launch on rank 0, seed %d" % seed}
        assert same, "sharded solve differs from the unsharded one"
    torch.cuda.synchronize()
    return out


def e2e_leg(eng, tabs, csr, starts, Mg, id_base, rank, world, dev, args, barrier):
    import torch
    import torch.distributed as tdist
    from nrw import dist as ndist
    n_nodes = len(starts)
    at = tabs.alias_host
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    h_row, h_alias = pin(at.rowinfo), pin(at.alias)
    h_starts = pin(starts.astype(np.int32))
    h_prefix = pin(np.arange(n_nodes + 1, dtype=np.int64) * Mg)
    h_idbase = pin(id_base)
    h_score = torch.zeros(n_nodes, dtype=torch.int64).pin_memory()
    h_cnt = torch.zeros(8, dtype=torch.int64).pin_memory()

    sr = at.structured()
    h_str = pin(sr[0]) if sr is not None else None

    class _View:           # AliasTables-like view over the pinned copies
        layout, n_alias = "compact", at.n_alias
        rowinfo, alias = h_row.numpy(), h_alias.numpy()
        rowinfo_str, str_thresh, str_n = (h_str.numpy(), sr[1], sr[2]) if sr is not None else (None, 0, 0)
    ex = ndist.TallyExchange(n_nodes, eng) if world > 1 else None
    if ex is not None and not ex.fused:
        return {"value": None, "unit": METRIC, "unavailable": "no peer memory between the ranks: the host entry needs the fused exchange"}
    solve = eng.prepare_fluence_scores_host(_View, csr.n_rows, csr.n_live, h_starts.numpy(), h_prefix.numpy(),
                                            0, n_nodes * Mg, NT, [0, 1, 0, 0], h_score.numpy(), h_cnt.numpy(),
                                            start_id_base=h_idbase.numpy(), exchange=ex)
    for k in range(3):
        solve(1000 + k)
    barrier()
    n_e2e = max(min(args.steps, 20), 1)
    t0 = time.perf_counter()
    e2e_steps = 0
    for k in range(n_e2e):
        solve(2000 + k)                                          # synchronous: the (reduced) scores are on the host
        e2e_steps += int(h_cnt[0].item())                        # reduced counter: walker-steps of ALL ranks
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        tdist.all_reduce(e2e_s, op=tdist.ReduceOp.MAX)
    # structured tables: only the structured descriptors are copied (the walk reads nothing else)
    tab_bytes = h_str.numel() * 4 if (h_str is not None and "_str<" in tabs_kernel_hint(eng, tabs, starts, Mg)) else \
        h_row.numel() * 4 + h_alias.numel() * 4
    h2d = tab_bytes + h_starts.numel() * 4 + h_prefix.numel() * 8 + h_idbase.numel() * 8
    d2h = h_score.numel() * 8 + 72
    return {"value": float(e2e_steps) / float(e2e_s.item()), "unit": METRIC,
            "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "steps": n_e2e,
            "ms_per_step": 1e3 * float(e2e_s.item()) / n_e2e,
            "entry": "nrw_fluence_scores_host (pinned host buffers)%s" %
                     ("; every rank uploads its inputs, the walk sums over ranks in its tail, every rank reads the "
                      "REDUCED scores back" if world > 1 else "")}


def tabs_kernel_hint(eng, tabs, starts, Mg):
    """Name of the kernel the planner picks for the score walk on these tables (tells whether the structured draw runs)."""
    _, r = eng.prepare_walk(tabs, starts, Mg, NT, seed=1, class_value=[0, 1, 0, 0], score=True)
    return r.kernel


def run_args_leg(rank, world, barrier):
    """The call a user of the reference makes: fluence_mini_app.run(args) (fluence_mini_app.py:251-447) on the C5
    mesh with -M 100000 --rng philox (1e9 walkers), cold and warm, with the split of where the time goes."""
    import fluence_mini_app as app
    tmp = tempfile.mkdtemp(prefix="nrw_bench_")
    a = vars(app.build_parser().parse_args(
        ["-r", "pc", "-M", "100000", "-dt", "0.01", "-da", "0.02", "--rng", "philox", "--rand_seed", "1",
         "--matrix_filename", os.path.join(tmp, "curmat.csv"), "--results_filename", os.path.join(tmp, "rwSol.csv")]))
    np.random.seed(1)
    times, splits = [], []
    for i in range(6):
        barrier()
        t0 = time.perf_counter()
        app.run(a)
        times.append(time.perf_counter() - t0)
        splits.append(dict(app.last_run_split))
    if rank != 0:
        return None
    warm = float(np.median(times[1:]))
    steps = splits[-1].get("walker_steps", 0)
    return {"value": steps / warm, "unit": METRIC, "cold_s": times[0], "warm_s": warm, "warm_all_s": times[1:],
            "walker_steps": steps, "walkers": 10000 * 100000, "n_gpus": world,
            "split_cold_s": {k: v for k, v in splits[0].items() if k.endswith("_s")},
            "split_warm_s": {k: v for k, v in splits[-1].items() if k.endswith("_s")},
            "entry": "fluence_mini_app.run(args): -r pc -M 100000 -dt 0.01 -da 0.02 --rng philox; matrix file (.npz, "
                     "written in the background, skipped when unchanged) and rwSol.csv included; process start-up "
                     "(import torch ~2 s, CUDA context ~1.2 s) is outside run() and not included"}


def resource_leg(eng, tabs, starts, Mg, id_base, flush, steps_per_launch, kernel_ms, det_frac, kernel=""):
    """The memory/atomic rooflines named by BASELINE north_star: rates of the resources the walk
    touches, measured by micro-benchmarks at the walk's own launch geometry (nrw_microbench), next
    to the rates the walk kernels reach.  None of them is the binding resource (issue slots are)."""
    import torch
    peaks = {"smem_random_load_8B_per_s": eng.microbench("smem_load", 20000),
             "smem_random_atomic_per_s": eng.microbench("smem_atomic", 10001),
             "l2_random_load_8B_per_s": eng.microbench("l2_load", 158761),
             "global_random_atomic_per_s": eng.microbench("global_atomic", 10001, iters=200)}
    out = {"measured_peaks": peaks,
           "how": "nrw_microbench: one 1024-thread CTA per SM, random indices from a per-thread LCG, "
                  "dependent loads; best of 5 event-timed launches"}
    # score kernel: shared-memory look-ups per drawn step -- two dependent ones with alias tables (alias entry, then
    # the target's row descriptor), one with the structured draw (the descriptor alone)
    per_step = 1.0 if "_str<" in kernel else 2.0
    lds = per_step * steps_per_launch * (1.0 - det_frac) / (kernel_ms * 1e-3)
    out["score_kernel"] = {"smem_loads_per_s": lds, "smem_loads_per_drawn_step": per_step,
                           "frac_of_smem_load_peak": lds / peaks["smem_random_load_8B_per_s"]}
    # occupancy-histogram mode: one shared-memory atomic per walker-step (sink arrivals are counted in registers)
    hl, hres = eng.prepare_walk(tabs, starts, Mg, NT, seed=1, score=True, class_value=[0, 1, 0, 0],
                                node_hist=True, start_id_base=id_base)
    for _ in range(3):
        hl()
    ms = time_kernel(torch, hl, flush)
    hist = hres.node_hist.cpu().numpy()
    atomics = float(hist[:-1].sum())
    out["hist_kernel"] = {"kernel": hres.kernel, "kernel_ms": ms,
                          "walker_steps_per_s": steps_per_launch / (ms * 1e-3),
                          "smem_atomics_per_s": atomics / (ms * 1e-3),
                          "frac_of_smem_atomic_peak": atomics / (ms * 1e-3) / peaks["smem_random_atomic_per_s"]}
    return out


def c2_leg(eng):
    """BASELINE configs[1]: fluence_mini_app paper parameters, M=1000, Nature-Electronics source,
    one B200: time of the whole fused solve and its error against the pc reference's expectation."""
    import torch
    from nrw import mesh
    from nrw.fluence import solve_fluence
    from oracle import expectation                    # checker only
    csr = mesh.build_fluence_csr_cached(**C1)
    src = mesh.source_values(csr.meta["mesh"]["meshpoints"], mesh.source_paper)
    mean, var = expectation.fluence_moments(csr.indptr, csr.cols, csr.vals, csr.n_rows, csr.n_live, src,
                                            NT_EXPECT, 0.0, C1["dt"], C1["v"])
    out = {}
    for M in (1000, 100000):
        for _ in range(3):
            r = solve_fluence(csr, M, NT, 0, C1["dt"], C1["v"], source=mesh.source_paper, rng="philox", seed=7)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        reps = 10
        for i in range(reps):
            r = solve_fluence(csr, M, NT, 0, C1["dt"], C1["v"], source=mesh.source_paper, rng="philox",
                              seed=100 + i)
        torch.cuda.synchronize()
        dt_s = (time.perf_counter() - t0) / reps
        st = expectation.fluence_error_stats(r.per_start, mean, var, M, C1["v"] * C1["dt"] * 0.015)
        out["M%d" % M] = {"walker_steps": r.walker_steps, "seconds_per_solve_host_to_host": dt_s,
                          "walker_steps_per_s": r.walker_steps / dt_s, "fluence_rel_l2_vs_pc_expectation": st["rel_l2"],
                          "max_abs_z": st["max_abs_z"], "frac_within_3sigma": st["frac_within_3sigma"],
                          "expected_rel_l2": st["expected_rel_l2"], "kernel": r.kernel}
    out["note"] = ("fluence err vs pc ref = relative L2 of rwSol against the exact expectation of the "
                   "reference's Markov chain (oracle/expectation.py); the pc reference's own M=1000 runs "
                   "sit at ~2.4e-2 (BASELINE.md)")
    return out


def c3_leg(eng, rank, world, flush):
    """BASELINE configs[2]: simple_flux_miniapp.py -sa 2 -ss 8 -dt 0.005 --use_sinks (absorption + sink exits),
    M = 1e8 walkers per start state, nt = 1002, on 1 GPU and -- under torchrun -- sharded over 2 (every rank walks
    its slice of both states' walkers, one all-reduce of the 2 x 1002 x 3 counts).  Error: max |rwSol - E| / sigma
    against the exact propagation of the reference's 3-state chain, and the distance to the closed form."""
    import torch
    import torch.distributed as tdist
    from nrw import dist as ndist, dropin, mesh
    M, nt = 100_000_000, 1002
    q = mesh.build_simple_flux_dense(0.005, 8.0, 2.0)
    csr = mesh.csr_from_dense(q, n_live=2)
    _, tabs = dropin.device_tables(csr)
    lo, hi = ndist.shard_range(M, rank, world)
    kw = dict(count_tn=True, t_cap=nt, count_per_start=True, start_id_base=[lo, M + lo])
    pw, res = eng.prepare_walk(tabs, [0, 1], [hi - lo, hi - lo], nt, 1, **kw)
    times = []
    for rep in range(4):
        pw.set_seed(100 + rep)
        flush.fill_(1)
        pw.zero()
        if world > 1:
            tdist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        pw.launch_kernel()
        if world > 1:
            ndist.allreduce_sum_([res.count_tn, res.counters])
        b.record()
        b.synchronize()
        times.append(a.elapsed_time(b))
    ms = torch.tensor([float(np.mean(times[1:]))], dtype=torch.float64, device=eng.device)
    if world > 1:
        tdist.all_reduce(ms, op=tdist.ReduceOp.MAX)
    if rank != 0:
        return None
    ms = float(ms.item())
    c = res.counters_dict()
    cnt = res.count_tn.cpu().numpy().astype(np.float64)            # [2, nt, 3] summed over ranks
    sol = (5.0 * cnt[:, :, 0] + 3.0 * cnt[:, :, 1]) / M            # rwSol[i, t] (simple_flux_miniapp.py:327)
    p, w, zmax = np.eye(2), np.array([5.0, 3.0]), 0.0
    for j in range(1, 1000):
        p = q[:2, :2] @ p if j > 1 else q[:2, :2].copy()
        mean = p @ w
        var = p @ (w * w) - mean ** 2
        zmax = max(zmax, float((np.abs(sol[:, j] - mean) / np.sqrt(np.maximum(var, 1e-300) / M)).max()))
    t = np.arange(nt) * 0.005
    e1, e2 = np.exp(-2.0 * t), np.exp(-10.0 * t)
    true = np.stack([2.5 * (e1 + e2) + 1.5 * (e1 - e2), 2.5 * (e1 - e2) + 1.5 * (e1 + e2)])
    absorbed = cnt[:, 1:, 2].sum()
    peak_atom = eng.microbench("smem_atomic", 6012)
    return {"config": "C3 simple_flux -sa 2 -ss 8 -dt 0.005 --use_sinks, M=%d per state, nt=%d, %d GPU(s)" % (M, nt, world),
            "kernel": res.kernel, "ms": ms, "walker_steps": c["steps"], "walker_steps_per_s": c["steps"] / (ms * 1e-3),
            "roofline": {"bound": "smem_atomic", "achieved": c["steps"] / world / (ms * 1e-3), "peak": peak_atom,
                         "unit": "atomics/s", "frac": c["steps"] / world / (ms * 1e-3) / peak_atom,
                         "note": "one shared-memory atomic per walker-step on the per-CTA (t, state) histogram; peak = "
                                 "nrw_microbench random atomics on 6012 bins at the same geometry"},
            "max_abs_z_vs_chain_expectation": zmax,
            "max_abs_err_vs_closed_form": float(np.abs(sol[:, 1:1000] - true[:, 1:1000]).max()),
            "absorbed": float(absorbed), "exits": c["exits"], "survivors": c["survivors"],
            "conservation_ok": bool(c["exits"] + c["survivors"] == c["walkers"] == 2 * M)}


def c4_leg(eng, flush):
    """BASELINE configs[3]: the refined mesh (dt=0.001, da=1/63: 158 760 nodes, nnz 2e7).  (a) the fused fluence
    solve, M = 1000 per node; error on every 7th start against the exact chain propagation (nt = 300 on both sides);
    (b) the densityrw.DensityRW surface: time-resolved occupancy of 2e7 walkers from 4 nodes over 40 steps, error =
    z of every (t, node) bin with an expected count >= 50 against the propagated occupancy."""
    import torch
    from densityrw.DensityRW import CSRTransitions, DensityRW
    from nrw import dropin, mesh
    from nrw.fluence import fluence_tables, solve_fluence
    from oracle import expectation                    # checker only
    t0 = time.perf_counter()
    csr = mesh.build_fluence_csr_cached(**C4)
    t_build = time.perf_counter() - t0
    src = mesh.source_values(csr.meta["mesh"]["meshpoints"], mesh.source_default)
    out = {"mesh": "dt=0.001 da=1/63: %d nodes, nnz=%d" % (csr.n_live, csr.nnz), "csr_build_s": t_build}
    # (a) fused solve
    M = 1000
    r = solve_fluence(csr, M, NT, 0, C4["dt"], C4["v"], rng="philox", seed=1)
    times = []
    for i in range(4):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = solve_fluence(csr, M, NT, 0, C4["dt"], C4["v"], rng="philox", seed=2 + i)
        times.append(time.perf_counter() - t0)
    _, tabs, _, _, _ = fluence_tables(csr, mesh.source_default)        # the tables solve_fluence just used
    launch, res = eng.prepare_walk(tabs, np.arange(csr.n_live), M, NT, seed=1, class_value=[0, 1, 0, 0], score=True)
    for _ in range(2):
        launch()
    kms = time_kernel(torch, launch, flush, reps=5)
    ksteps = int(res.counters.cpu()[0].item())
    nt_e, Me = 300, 4000
    sub = np.arange(0, csr.n_live, 7)
    mean, var = expectation.fluence_moments(csr.indptr, csr.cols, csr.vals, csr.n_rows, csr.n_live, src, nt_e, 0.0,
                                            C4["dt"], C4["v"])
    rs = eng.walk(tabs, sub, Me, nt_e, 21, score=True, class_value=[0, 1, 0, 0])
    unit = C4["v"] * C4["dt"] * 3.0
    st = expectation.fluence_error_stats(rs.score.cpu().numpy().astype(np.float64) * unit / Me, mean[sub], var[sub], Me, unit)
    out["fused_solve"] = {"M": M, "kernel": r.kernel, "walker_steps": r.walker_steps,
                          "seconds_per_solve_host_to_host": float(np.median(times)),
                          "walker_steps_per_s_host_to_host": r.walker_steps / float(np.median(times)),
                          "kernel_ms": kms, "walker_steps_per_s": ksteps / (kms * 1e-3),
                          "steps_per_walker": r.walker_steps / (csr.n_live * M),
                          "fluence_err": dict(st, how="every 7th start node, M=%d, nt=%d on both sides" % (Me, nt_e))}
    # (b) densityrw surface
    Md, nt_d = 20_000_000, 40
    dstarts = [79443, 40000, 120000, 1000]
    walk = DensityRW(nt_d, CSRTransitions(csr), init_walkers={(s,): Md // len(dstarts) for s in dstarts})
    np.random.seed(3)
    walk.run(rng="philox")                                            # tables, first launch
    t0 = time.perf_counter()
    counts = walk.run(rng="philox")
    t_run = time.perf_counter() - t0
    _, dt_tabs = dropin.device_tables(walk.csr)
    pw, dres = eng.prepare_walk(dt_tabs, dstarts, [Md // len(dstarts)] * len(dstarts), nt_d, 50, count_tn=True,
                                t_cap=nt_d, count_path=True)
    for _ in range(2):
        pw()
    dms = time_kernel(torch, pw, flush, reps=5)
    c = dres.counters_dict()
    expect, evar = np.zeros((nt_d, csr.n_rows)), np.zeros((nt_d, csr.n_rows))
    for s_node in dstarts:                                # counts = sum over starts of Binomial(N_s, p_s(t, node))
        sd = np.zeros(csr.n_rows)
        sd[s_node] = Md // len(dstarts)
        es = expectation.occupancy_expectation(csr.indptr, csr.cols, csr.vals, csr.n_rows, csr.n_live, sd, nt_d)
        expect += es
        evar += es * (1.0 - es / (Md // len(dstarts)))
    got = dres.count_tn.cpu().numpy()[0].astype(np.float64)          # [t, node]
    big = (expect[1:] >= 50.0) & (evar[1:] > 0)
    z = np.abs(got[1:][big] - expect[1:][big]) / np.sqrt(evar[1:][big])
    peak_atom = eng.microbench("global_atomic", 1 << 22, iters=100)
    out["densityrw"] = {"config": "%d walkers from %d nodes, %d steps, time-resolved occupancy count[t, node] "
                                  "(%d bins)" % (Md, len(dstarts), nt_d, nt_d * csr.n_rows),
                        "kernel": dres.kernel, "kernel_ms": dms, "walker_steps": c["steps"],
                        "walker_steps_per_s": c["steps"] / (dms * 1e-3),
                        "DensityRW_run_seconds_host_to_host": t_run,
                        "roofline": {"bound": "global_atomic", "achieved": c["steps"] / (dms * 1e-3), "peak": peak_atom,
                                     "unit": "atomics/s", "frac": c["steps"] / (dms * 1e-3) / peak_atom,
                                     "note": "the reference algorithm tallies one (t, node) bin per walker-step; peak = "
                                             "random 64-bit global atomics on 4M bins (nrw_microbench).  frac > 1 means "
                                             "the per-CTA privatisation (path histogram / hot-bin cache) keeps most "
                                             "tallies out of global memory"},
                        "max_abs_z_bins_ge_50": float(z.max()), "frac_within_3sigma": float((z < 3.0).mean()),
                        "bins_tested": int(big.sum()), "counts_total_ok": bool(got[1:].sum() == c["steps"])}
    return out


def _protect_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version
    banner to stdout on the first collective), so file descriptor 1 is pointed at stderr for the
    duration of the run and the JSON line goes to the saved descriptor."""
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    out = os.fdopen(saved, "w")

    def emit(line):
        out.write(line + "\n")
        out.flush()
    return emit


_emit = print


def main():
    global _emit
    args = parse_args()
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    _emit = _protect_stdout()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
