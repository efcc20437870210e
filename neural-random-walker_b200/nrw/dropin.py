"""Shared machinery behind the drop-in entry points (fluence_mini_app.run_pc / run,
simple_flux_miniapp.run_pc / run, loihirw.rw_model.rad_test, densityrw.*).

The reference's back ends share one signature (fluence_mini_app.py:218-222, :399-407):
    run_X(matrix_filename, neuraltime, initial_walkers=[[node, M]], meshSize, runtime_args) -> float64[rows, nt]
and the matrix travels BY FILENAME.  Here the file is parsed once and its device tables are cached,
and two RNG modes are offered:
  * "replay" -- consume the global legacy ``np.random`` stream exactly like the reference
    (bit-identical tempdata / rwSol, stream left at the reference's position);
  * "philox" -- counter-based on-device RNG (statistically equivalent, orders of magnitude faster).
"""
from __future__ import annotations

import logging
import os

import numpy as np

from . import mesh as _mesh
from .engine import get_engine

rw_logger = logging.getLogger("NRW")

RNG_MODES = ("auto", "replay", "philox")
AUTO_REPLAY_MAX_WALKERS = 1 << 22     # "auto": bit-exact replay up to ~4M walkers, Philox beyond
SAMPLERS = ("alias", "cdf")

_table_cache: dict = {}


def resolve_rng(mode, total_walkers: int) -> str:
    mode = mode or os.environ.get("NRW_RNG", "auto")
    if mode not in RNG_MODES:
        raise ValueError("rng must be one of %s" % (RNG_MODES,))
    if mode == "auto":
        return "replay" if total_walkers <= AUTO_REPLAY_MAX_WALKERS else "philox"
    return mode


def device_tables(csr, node_class=None, key=None):
    """Upload (once) and cache the device tables of a matrix."""
    eng = get_engine()
    k = (id(csr), None if node_class is None else node_class.tobytes(), eng.device.index) \
        if key is None else (key, eng.device.index)
    hit = _table_cache.get(k)
    if hit is None:
        if len(_table_cache) > 8:
            _table_cache.clear()
        hit = (eng.upload(csr, node_class=node_class), csr)      # keep csr alive: id() stays unique
        _table_cache[k] = hit
    return eng, hit[0]


_pinned = {}


def to_host(t):
    """Device tensor -> freshly owned numpy array.  Large results (the dense tempdata the reference's
    call signature asks for is 216 MB at nt = 30000 on the default mesh) go through a cached pinned
    staging buffer: a pageable ``.cpu()`` runs at ~2 GB/s, pinned DMA plus one host copy at ~3x that."""
    nbytes = t.numel() * t.element_size()
    if nbytes < (8 << 20) or not t.is_cuda:
        return t.cpu().numpy()
    import torch
    t = t.contiguous()
    buf = _pinned.get(t.device.index)
    if buf is None or buf.numel() < nbytes:
        _pinned[t.device.index] = buf = torch.empty(int(nbytes * 1.25), dtype=torch.uint8).pin_memory()
    view = buf[:nbytes].view(t.dtype).view(t.shape)
    view.copy_(t)
    return view.numpy().copy()


def philox_seed_from_global_stream() -> int:
    """One 63-bit seed per call, drawn from the legacy global stream so that ``np.random.seed(s)``
    (fluence_mini_app.py:523) still makes Philox-mode runs reproducible."""
    return int(np.random.randint(0, 2 ** 63 - 1, dtype=np.int64))


def pc_tempdata(csr, neuraltime, initial_walkers, n_out_rows, rng=None, sampler="alias"):
    """The body of a drop-in ``run_pc``: float64 [n_out_rows, neuraltime] occupancy counts for
    ``initial_walkers[0]`` (only element 0 is honoured, fluence_mini_app.py:229-230, Q11)."""
    place = int(initial_walkers[0][0])
    amount = int(initial_walkers[0][1])
    nt = int(neuraltime)
    if not (0 <= place < csr.n_rows):
        raise IndexError("index %d is out of bounds for axis 0 with size %d" % (place, csr.n_rows))
    eng, tabs = device_tables(csr)
    mode = resolve_rng(rng, amount)
    if amount <= 0 or nt <= 1:
        out = np.zeros((n_out_rows, nt))
        if nt >= 1:
            out[place, 0] = amount
        return out
    if mode == "replay":
        rep = eng.replay(tabs, [place], [amount], nt)
        dense = eng.expand_events(rep, 0, csr.n_rows, nt)
    else:
        seed = philox_seed_from_global_stream()
        t_cap = min(nt, 1024)
        while True:
            res = eng.walk(tabs, [place], [amount], nt, seed, sampler=sampler, count_tn=True,
                           t_cap=t_cap, count_path=True)
            if res.counters_dict()["t_overflow"] == 0 or t_cap >= nt:
                break
            t_cap = min(nt, t_cap * 8)       # same seed -> same trajectories, longer tally window
        dense = eng.expand_counts(res.count_tn, res.t_cap, csr.n_rows, nt, place, amount)
    return to_host(dense[:n_out_rows])


def multi_source_tempdata(csr, neuraltime, initial_walkers, n_out_rows, rng=None, sampler="alias"):
    """Occupancy counts float64 [n_out_rows, neuraltime] for walkers started on SEVERAL nodes at
    once (``initial_walkers = [[node, M], ...]``; the ``init_walkers`` dict of the reference's
    densityrw.DensityRW, DensityRW.py:28-41, and the ``location`` list of loihirw rad_test,
    rw_model.py:70-78).  Walkers never interact, so the joint occupancy is the sum over sources:
    Philox mode walks all sources in one K1 launch; replay mode walks the sources in list order on
    the global legacy stream (there is no multi-source pc reference to be bit-exact with; a single
    source reproduces run_pc exactly)."""
    init = [(int(a), int(b)) for a, b in initial_walkers if int(b) > 0]
    nt = int(neuraltime)
    for place, _ in init:
        if not (0 <= place < csr.n_rows):
            raise IndexError("index %d is out of bounds for axis 0 with size %d" % (place, csr.n_rows))
    out = np.zeros((n_out_rows, max(nt, 0)))
    if not init or nt < 1:
        return out
    if len(init) == 1:
        return pc_tempdata(csr, nt, [list(init[0])], n_out_rows, rng=rng, sampler=sampler)
    eng, tabs = device_tables(csr)
    mode = resolve_rng(rng, sum(b for _, b in init))
    if nt == 1:
        for place, amount in init:
            if place < n_out_rows:
                out[place, 0] += amount
        return out
    nodes = [a for a, _ in init]
    counts = [b for _, b in init]
    if mode == "replay":
        rep = eng.replay(tabs, nodes, counts, nt)
        dense = None
        for s in range(len(nodes)):
            d = eng.expand_events(rep, s, csr.n_rows, nt)
            dense = d if dense is None else dense.add_(d)
    else:
        seed = philox_seed_from_global_stream()
        t_cap = min(nt, 1024)
        while True:
            res = eng.walk(tabs, nodes, counts, nt, seed, sampler=sampler, count_tn=True, t_cap=t_cap,
                           count_path=True)
            if res.counters_dict()["t_overflow"] == 0 or t_cap >= nt:
                break
            t_cap = min(nt, t_cap * 8)
        dense = eng.expand_counts(res.count_tn, res.t_cap, csr.n_rows, nt, nodes[0], 0)
        for place, amount in init:
            dense[place, 0] += amount
    return to_host(dense[:n_out_rows])


# ---- adapters of the neuromorphic dispatch entries onto the B200 surfaces ---------------------
def spike_table_to_tempdata(results, meshSize):
    """The post-processing of fluence_mini_app.run_fugu (:66-107), vectorised: a model step starts
    wherever consecutive spike times differ by more than 1; ``tempdata[node, step]`` = number of
    spikes of ``node`` in ``step`` for nodes < meshSize and steps < NT, NT = index of the last step
    (the reference sizes the array by the last step's index, so that step itself is dropped)."""
    time_array = np.asarray(results['time'], dtype=np.float64)
    if len(time_array) == 0:
        return np.zeros((meshSize, 0))
    timestep = np.zeros(len(time_array), dtype=np.int64)
    timestep[1:] = np.cumsum(np.diff(time_array) > 1)
    NT = int(timestep[-1])
    node = np.array([int(str(x).strip("(), ").split(",")[0]) if not isinstance(x, (int, np.integer))
                     else int(x) for x in results['node']], dtype=np.int64)
    tempdata = np.zeros((meshSize, NT))
    keep = (node >= 0) & (node < meshSize) & (timestep < NT)
    np.add.at(tempdata, (node[keep], timestep[keep]), 1.0)
    return tempdata


def fugu_adapter(matrix_filename=None, neuraltime=None, initial_walkers=None, meshSize=None,
                 runtime_args={}):
    """``run_fugu`` of the reference (fluence_mini_app.py:43-109) over this build's
    ``densityrw.run_miniapp``: same call, same decode, the walk itself runs on the GPU."""
    from densityrw.rw_miniapp_fugu import run_miniapp
    results = run_miniapp(matrix_filename, neuraltime, initial_walkers,
                          runtime_args.get('use_sinks', False), runtime_args.get('fugu_backend', 'snn'),
                          rng=runtime_args.get('rng'), sampler=runtime_args.get('sampler', 'alias'))
    return spike_table_to_tempdata(results, meshSize)


def loihi_adapter(matrix_filename=None, neuraltime=None, initial_walkers=None, meshSize=None,
                  runtime_args={}):
    """``run_loihi`` of the reference (fluence_mini_app.py:111-192) over this build's
    ``loihirw``: rad_test -> track-6 probes -> process_outputs_split -> count[meshSize, bins + 1]."""
    from loihirw.rw_analysis import process_outputs_split
    from loihirw.rw_model import rad_test
    assert len(initial_walkers) == 1
    i, num_walkers = initial_walkers[0][0], initial_walkers[0][1]
    results = rad_test(matrix_filename, [i], neuraltime, num_walkers=num_walkers,
                       core_density=runtime_args.get('core_density', 1.1), verbose=0,
                       input_times=runtime_args.get('input_times', [1]),
                       track=runtime_args.get('track', 6), size=(meshSize, 1),
                       rng=runtime_args.get('rng'), sampler=runtime_args.get('sampler', 'alias'))
    track_probes = results[2]
    count, num_bins = process_outputs_split(track_probes, 0, 1, meshSize)
    rw_logger.info("Number of bins: {}".format(num_bins))
    return count


def unavailable_backend(name: str, needs: str):
    """Dispatch-table entry for the neuromorphic back ends this build does not carry."""
    def _run(matrix_filename=None, neuraltime=None, initial_walkers=None, meshSize=None,
             runtime_args={}):
        raise NotImplementedError(
            "run_mode '%s' needs %s, which this B200 build does not include; "
            "use -r pc (same call signature and return value)" % (name, needs))
    _run.__name__ = "run_" + name
    return _run


def exp_time_weights(SigA, dt, nt):
    """w[j] = np.exp(-SigA*dt*j) evaluated scalar by scalar as fluence_mini_app.py:434 does (a
    vectorised np.exp may take a SIMD path with different last bits)."""
    w = np.empty(nt)
    lim = min(nt, 1 << 16)
    for j in range(lim):
        w[j] = np.exp(-SigA * dt * j)
    if lim < nt:
        w[lim:] = np.exp(-SigA * dt * np.arange(lim, nt))
    return w


def time_seed():
    """fluence_mini_app.py:519 (Q10; the '+hour*24' is the reference's, kept verbatim)."""
    from datetime import datetime
    return (datetime.now().microsecond + datetime.now().second + datetime.now().minute * 60
            + datetime.now().hour * 24)
