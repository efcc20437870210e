"""``DensityRW`` on a B200 (reference: densityrw/DensityRW.py).

Kept: ``load_transitions(mat_file, remove_sink_connections, verbose)`` (:177-205) with its
``{(i,): Transition}`` result and sink-removal rule, ``build_test_graph`` (:150-159), and the
constructor ``DensityRW(timesteps, transitions, init_walkers, name, coding)`` (:28-56).  The reference
class is a Fugu brick that emits a spiking network; this one converts the transitions to CSR tables
and ``run()`` walks them on the GPU, returning walker counts per (node, model step).
"""
import logging
import time

import numpy as np

from nrw import dropin, mesh as _mesh
from . import SpikingPDE

rw_logger = logging.getLogger("NRW")

test_timesteps = 50
num_test_nodes = 25


class CSRTransitions(dict):
    """``{(i,): Transition}`` view of a CSR matrix whose Transition objects are built on access --
    what ``load_transitions`` returns for ``.npz`` (CSR) files, so that meshes far beyond the
    dense-CSV limit (the 158 760-node refined mesh: 1.9e7 edges) never become Python tuples."""

    def __init__(self, csr, remove_sink_connections=True):
        super().__init__()
        self.csr = csr
        self.remove_sink_connections = remove_sink_connections

    def __len__(self):
        return self.csr.n_rows

    def __iter__(self):
        return ((i,) for i in range(self.csr.n_rows))

    def keys(self):
        return list(iter(self))

    def __contains__(self, key):
        return isinstance(key, tuple) and len(key) == 1 and 0 <= int(key[0]) < self.csr.n_rows

    def __getitem__(self, key):
        if key not in self:
            raise KeyError(key)
        i = int(key[0])
        lo, hi = int(self.csr.indptr[i]), int(self.csr.indptr[i + 1])
        nb = [((int(c),), float(p)) for c, p in zip(self.csr.cols[lo:hi], self.csr.vals[lo:hi]) if p > 0]
        if self.remove_sink_connections and len(nb) == 1 and nb[0] == ((i,), 1.0):
            nb = []
        return SpikingPDE.Transition(location=(i,), neighbors=nb)

    def items(self):
        return [(k, self[k]) for k in self]

    def values(self):
        return [self[k] for k in self]


def load_transitions(mat_file, remove_sink_connections=True, verbose=1):
    """DensityRW.py:177-205: matrix file -> ``{(i,): Transition}``; an entry > 0 is an edge; a node
    whose only edge is a probability-1 self-loop is a sink and, with ``remove_sink_connections``,
    loses it (walkers arriving there are counted once and stop).  ``.csv`` is parsed as the reference
    does; ``.mat`` needs scipy; ``.npz`` holds CSR (additive).  The result is a ``CSRTransitions``: a
    dict-like ``{(i,): Transition}`` view of the parsed matrix."""
    SpikingPDE.reset_neuron_list()
    if '.npz' in mat_file or '.csv' in mat_file:
        # parsed once per (path, mtime, size) -- the reference re-parses the CSV on every call
        # (0.5 s for the 901 x 901 default matrix, once per start node) -- and kept as CSR; the
        # Transition objects are built when a key is looked up
        csr = _mesh.load_matrix(mat_file)
        if csr.n_live != csr.n_rows:           # every row may be stepped on this surface; derived once, so
            full = csr.meta.get("_all_live")   # that repeated calls share one set of device tables
            if full is None:
                full = _mesh.TransitionCSR(csr.n_rows, csr.n_rows, csr.indptr, csr.cols, csr.vals, {})
                csr.meta["_all_live"] = full
            csr = full
        transitions = CSRTransitions(csr, remove_sink_connections)
    elif '.mat' in mat_file:
        from scipy.io import loadmat
        prob_mtx = np.asarray(loadmat(mat_file)[mat_file[:-4]], dtype=np.float64)
        transitions = CSRTransitions(_mesh.csr_from_dense(prob_mtx, n_live=prob_mtx.shape[0]),
                                     remove_sink_connections)
    else:
        raise ValueError("Matrix file incompatible")
    if verbose > 0:
        csr = transitions.csr
        for i in range(csr.n_rows):
            p = 0
            for x in csr.vals[csr.indptr[i]:csr.indptr[i + 1]]:     # sequential sum, as the reference (:192)
                p += x
            rw_logger.info("node " + str(i) + " has total probability of " + str(p))
            nb = transitions[(i,)].neighbors
            if remove_sink_connections and not nb and csr.indptr[i + 1] - csr.indptr[i] == 1:
                rw_logger.info("Removing connections from a sink at " + str((i,)))
    return transitions


def build_test_graph(num_nodes=20):
    """Ring of ``num_nodes`` nodes, probability 0.5 to either neighbour (DensityRW.py:150-159).
    Neighbour order and the last node's ``location=(9,)`` label are the reference's."""
    last = num_nodes - 1
    transitions = {}
    for i in range(num_nodes):
        nxt, prv = ((i + 1) % num_nodes,), ((i - 1) % num_nodes,)
        order = [prv, nxt] if i == last else [nxt, prv]
        transitions[(i,)] = SpikingPDE.Transition(location=(9,) if i == last else (i,),
                                                  neighbors=[(d, 0.5) for d in order])
    return transitions


def transitions_to_csr(transitions) -> _mesh.TransitionCSR:
    """``{location: Transition}`` -> CSR.  Locations are ordered by sorted key; columns ascending
    within a row (the order ``np.cumsum`` over the dense row sees); duplicate edges are summed."""
    if isinstance(transitions, CSRTransitions):
        csr = transitions.csr
        if not transitions.remove_sink_connections:
            return csr
        if "_no_sinks" in csr.meta:
            return csr.meta["_no_sinks"]
        lens = csr.row_lengths()
        rows = np.nonzero(lens == 1)[0]
        at = csr.indptr[rows]
        sinks = rows[(csr.cols[at] == rows) & (csr.vals[at] == 1.0)]
        if len(sinks) == 0:
            csr.meta["_no_sinks"] = csr
            return csr
        keep = np.ones(csr.nnz, dtype=bool)
        keep[csr.indptr[sinks]] = False
        lens = lens.copy()
        lens[sinks] = 0
        indptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        out = _mesh.TransitionCSR(csr.n_rows, csr.n_rows, indptr, csr.cols[keep], csr.vals[keep], {})
        csr.meta["_no_sinks"] = out
        return out
    if isinstance(transitions, _mesh.TransitionCSR):
        return transitions
    keys = sorted(transitions.keys())
    index = {k: i for i, k in enumerate(keys)}
    n = len(keys)
    indptr = np.zeros(n + 1, dtype=np.int64)
    cols, vals = [], []
    for i, k in enumerate(keys):
        acc = {}
        for dest, p in transitions[k].neighbors:
            if dest not in index:
                raise KeyError("transition from %r to unknown location %r" % (k, dest))
            acc[index[dest]] = acc.get(index[dest], 0.0) + float(p)
        for c in sorted(acc):
            cols.append(c)
            vals.append(acc[c])
        indptr[i + 1] = len(cols)
    csr = _mesh.TransitionCSR(n, n, indptr, np.asarray(cols, dtype=np.int32),
                              np.asarray(vals, dtype=np.float64), dict(kind="densityrw"))
    csr.meta["locations"] = keys
    return csr


class DensityRW:
    """Density random walk on an arbitrary sparse graph (constructor of DensityRW.py:28-56).

    Arguments:
        + timesteps - number of MODEL steps to walk (the reference counts neural ticks)
        + transitions - {source_location: Transition}; probabilities of a row are used as given
        + init_walkers - {start_location: number_of_walkers}
    """

    def __init__(self, timesteps, transitions, init_walkers={(10,): 30, (13,): 30}, name=None,
                 coding='Raster'):
        self.is_built = False
        self.metadata = {}
        self.output_coding = coding
        self.name = name
        self.brick_tag = 'DensityRW'
        self.transitions = transitions
        self.init_walkers = init_walkers
        self.timesteps = timesteps
        self.csr = None
        self.locations = None

    def build(self, *unused, **unused_kw):
        """Convert the transitions to CSR once (the reference's ``build`` lays a spiking network,
        DensityRW.py:58-147)."""
        if not self.is_built:
            self.csr = transitions_to_csr(self.transitions)
            self.locations = self.csr.meta.get("locations") or [(i,) for i in range(self.csr.n_rows)]
            self._index = None if "locations" not in self.csr.meta else \
                {k: i for i, k in enumerate(self.locations)}
            self.is_built = True
        return self

    def _node(self, location):
        if self._index is not None:
            return self._index[location]
        return int(location[0]) if isinstance(location, tuple) else int(location)

    def run(self, rng=None, sampler="alias"):
        """Walk; returns float64 ``counts[n_nodes, timesteps]``: walkers standing on each node at
        each model step (column 0 = ``init_walkers``).  A walker arriving on a node without outgoing
        edges is counted there once and removed."""
        self.build()
        init = [[self._node(loc), int(n)] for loc, n in self.init_walkers.items()]
        t0 = time.time()
        counts = dropin.multi_source_tempdata(self.csr, int(self.timesteps), init, self.csr.n_rows,
                                              rng=rng, sampler=sampler)
        rw_logger.debug("DensityRW: %d nodes, %d steps in %.4f s", self.csr.n_rows, self.timesteps,
                        time.time() - t0)
        return counts
