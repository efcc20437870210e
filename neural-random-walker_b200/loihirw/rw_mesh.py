"""``rw_mesh`` on a B200: transition-matrix file -> device CSR mesh -> probe-shaped walk output.

The reference's rw_mesh.py loads the matrix as ``networkx.DiGraph(np.genfromtxt(file))``
(rw_mesh.py:183: an edge i -> j with attribute ``weight`` for every non-zero entry), places the mesh
points on Loihi cores (:196-241), instantiates the spiking circuits and returns the spike probes.
Here the same file becomes the CSR tables of ``nrw`` (cached per file), the "network" is the GPU
walk, and ``makeNetwork`` returns probes synthesised from the walk's counts (``loihirw.probes``) in
the layout of the requested ``track`` mode, so the reference's own decoders read them unchanged.
"""
import logging
import math

import numpy as np

from nrw import dropin, mesh as _mesh
from . import probes as _probes

rw_logger = logging.getLogger("NRW")

CORES_PER_CHIP = 128


def load_mesh(custom_network, n_live=None):
    """Matrix file (dense CSV as the miniapps write it, or CSR ``.npz``) -> ``TransitionCSR``.
    Edge <=> non-zero entry, weight = entry (the DiGraph semantics of rw_mesh.py:183).  Rows with no
    outgoing edge are absorbing; by default every row may be stepped (``n_live = n_rows``)."""
    if isinstance(custom_network, _mesh.TransitionCSR):
        return custom_network
    if isinstance(custom_network, np.ndarray):
        q = np.asarray(custom_network, dtype=np.float64)
        return _mesh.csr_from_dense(q, n_live=q.shape[0] if n_live is None else n_live)
    csr = _mesh.load_matrix(custom_network)
    want = csr.n_rows if n_live is None else int(n_live)
    if csr.n_live != want:                      # derived once per file: repeated calls share device tables
        key = "_n_live_%d" % want
        if key not in csr.meta:
            csr.meta[key] = _mesh.TransitionCSR(csr.n_rows, want, csr.indptr, csr.cols, csr.vals, {})
        csr = csr.meta[key]
    return csr


def mesh_graph(custom_network):
    """The reference's ``networkx.DiGraph`` view of the mesh (needs networkx; for inspection only --
    the walk never builds it)."""
    import networkx
    csr = load_mesh(custom_network)
    g = networkx.DiGraph()
    g.add_nodes_from(range(csr.n_rows))
    rows = np.repeat(np.arange(csr.n_rows), csr.row_lengths())
    g.add_weighted_edges_from(zip(rows.tolist(), csr.cols.tolist(), csr.vals.tolist()))
    return g


def core_assignment(n_nodes, core_density=9.0, core_start=0):
    """Where the reference would have placed the mesh points on Loihi (rw_mesh.py:206-241): node k
    (in mesh order) sits on deterministic core ``core_start + 1 + k // (10*core_density)``; one more
    core follows for the stochastic neurons.  Returns (core_id[n_nodes], max_coreid, chip_count).
    Informational on a GPU -- kept so that callers that log or size by it keep working."""
    per_core = core_density * 10
    k = np.arange(n_nodes)
    if per_core == int(per_core) and per_core >= 1:
        core_id = core_start + 1 + k // int(per_core)
    else:           # a fractional density never triggers the reference's `==` roll-over test (:219)
        core_id = np.full(n_nodes, core_start + 1, dtype=np.int64)
    last = int(core_id.max(initial=core_start))
    rolls = n_nodes // int(per_core) if (per_core == int(per_core) and per_core >= 1) else 0
    final = core_start + 1 + rolls                      # `current_core` when the loop ends
    chip_count = 1 + final // CORES_PER_CHIP - (core_start + 1) // CORES_PER_CHIP
    return core_id, max(last, core_start) + 1, chip_count


def makeNetwork(superNet, x, y, custom_network, core_density=9.0,
                input_x=[(0, 0), (0, 1), (0, 2), (1, 0), (1, 1), (1, 2)], core_start=0,
                sim_time=10000, num_walkers=100, track=0, verbose=0, input_times=[1], s=0,
                spikeStreamer=0, rng=None, sampler="alias", spike_mode="auto"):
    """Signature of rw_mesh.py:137-150.  ``superNet`` / ``s`` / ``spikeStreamer`` (NxSDK objects) are
    accepted and ignored.  ``input_x`` lists the mesh points that receive walkers (ints, or the
    reference's (row, col) tuples of a x-by-y grid, flattened row-major); each receives
    ``num_walkers`` walkers per entry of ``input_times`` (define_inputs, rw_mesh.py:24-26).
    ``sim_time`` counts MODEL steps here (there are no neural ticks on a GPU).

    Returns ``(core_neuroncount, max_coreid, track_probes, chip_count, s)`` as the reference does;
    ``track_probes`` follows ``track``: 6 -> split layout, 2/4/5 -> flat layout, 1/3 -> end-state
    counters, 0 -> [] (walk still runs, nothing is read back -- the reference's probe-free rerun)."""
    csr = load_mesh(custom_network)
    n = csr.n_rows
    locs = []
    for loc in input_x:
        if isinstance(loc, (tuple, list, np.ndarray)):
            loc = int(loc[0]) * int(y) + int(loc[1]) if len(loc) == 2 else int(loc[0])
        locs.append(int(loc))
    per_loc = int(num_walkers) * max(len(input_times), 1)
    rw_logger.info("    Run starting")
    rw_logger.info("        Num walkers: {}".format(num_walkers))
    rw_logger.info("        locations: {}".format(locs))
    rw_logger.info("        time: {}".format(sim_time))
    counts = dropin.multi_source_tempdata(csr, int(sim_time), [[k, per_loc] for k in locs], n,
                                          rng=rng, sampler=sampler)
    core_id, max_coreid, chip_count = core_assignment(n, core_density, core_start)
    core_neuroncount = np.zeros((48640,), dtype=int)
    np.add.at(core_neuroncount, np.minimum(core_id, 48639), 1)
    live = int(np.count_nonzero(csr.row_lengths()[:csr.n_live]))
    if track == 6:
        track_probes, _ = _probes.split_track_probes(counts, live_rows=_live_rows(csr), spike_mode=spike_mode)
    elif track in (2, 4, 5):
        track_probes, _ = _probes.flat_track_probes(counts, spike_mode=spike_mode)
    elif track in (1, 3):
        track_probes = _probes.ends_track_probes(counts.sum(axis=1), counts.shape[1])
    else:
        track_probes = []
    if verbose:
        rw_logger.info("    {} mesh points, {} with outgoing edges, {} cores on {} chips".format(
            n, live, max_coreid, chip_count))
    return core_neuroncount, max_coreid, track_probes, chip_count, s


def _live_rows(csr):
    """Rows up to the last one that has outgoing edges (trailing absorbing rows = sinks)."""
    nz = np.nonzero(csr.row_lengths()[:csr.n_live])[0]
    return int(nz[-1]) + 1 if len(nz) else 0
