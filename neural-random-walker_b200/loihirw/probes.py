"""Result writer feeding ``rw_analysis``: walker counts -> probe-shaped spike rasters.

The reference reads NxSDK spike probes (``probe.data[neuron][time]``) and bins them into model
steps (loihirw/rw_analysis.py:142-179).  The GPU walk produces the counts directly, so this module
writes them back out in the probe layout the reference's decoders expect, such that the *unmodified*
decoders return exactly those counts:

  neural time is cut into model steps of ``period = width + 1`` ticks; the supervisor raster has a
  1 on the last tick of each step (the bin boundary), and counter neuron 1 of mesh point j fires
  once per walker standing on j during the first ``count[j, step]`` ticks of the step ("unary"
  mode) or carries the count on the first tick ("count" mode, for walker numbers whose unary
  raster would not fit in memory -- the decoders only ever sum the raster over a bin).

Rasters are generated row by row on access; nothing of size nodes x time is materialised.
"""
import numpy as np

UNARY_MAX_CELLS = 1 << 28


class LazyRaster:
    """``data`` of a probe: ``data[:]`` is the raster itself, ``data[k]`` builds neuron k's row."""

    def __init__(self, n_neurons, n_ticks, row_fn):
        self.shape = (n_neurons, n_ticks)
        self._row_fn = row_fn

    def __len__(self):
        return self.shape[0]

    def __getitem__(self, key):
        if isinstance(key, slice):
            if key == slice(None):
                return self
            return [self._row_fn(k) for k in range(*key.indices(self.shape[0]))]
        k = int(key)
        if k < 0:
            k += self.shape[0]
        if not 0 <= k < self.shape[0]:
            raise IndexError("probe has %d neurons" % self.shape[0])
        return self._row_fn(k)

    def __array__(self, dtype=None, copy=None):
        a = np.stack([self._row_fn(k) for k in range(self.shape[0])])
        return a if dtype is None else a.astype(dtype)


class Probe:
    """Stand-in for an NxSDK probe: only ``.data`` is used by the decoders."""

    def __init__(self, data):
        self.data = data


class RasterLayout:
    """Tick layout shared by every probe of one run."""

    def __init__(self, counts, spike_mode="auto"):
        counts = np.asarray(counts)
        self.counts = counts
        self.n_nodes, self.n_steps = counts.shape
        peak = int(counts.max(initial=0))
        if spike_mode == "auto":
            cells = (peak + 1) * self.n_steps * max(self.n_nodes, 1)
            spike_mode = "unary" if cells <= UNARY_MAX_CELLS else "count"
        if spike_mode not in ("unary", "count"):
            raise ValueError("spike_mode must be 'auto', 'unary' or 'count'")
        self.spike_mode = spike_mode
        self.width = max(peak, 1) if spike_mode == "unary" else 1
        self.period = self.width + 1
        self.n_ticks = self.n_steps * self.period
        self.dtype = np.uint8 if spike_mode == "unary" else np.float64

    def boundary(self):
        b = np.zeros(self.n_ticks, dtype=np.uint8)
        b[self.width::self.period] = 1
        return b

    def step_start(self):
        b = np.zeros(self.n_ticks, dtype=np.uint8)
        b[0::self.period] = 1
        return b

    def spikes(self, node):
        c = self.counts[node]
        grid = np.zeros((self.n_steps, self.period), dtype=self.dtype)
        if self.spike_mode == "unary":
            grid[:, :self.width] = np.arange(self.width)[None, :] < c[:, None]
        else:
            grid[:, 0] = c
        return grid.reshape(-1)

    def empty_from(self, live_rows):
        """1 on every tick from the first model step on which no walker stands on a live node
        (the reference's 'First empty number', fluence_mini_app.py:169-172)."""
        alive = self.counts[:live_rows].sum(axis=0) > 0
        dead = np.nonzero(~alive)[0]
        z = np.zeros(self.n_ticks, dtype=np.uint8)
        if len(dead):
            z[int(dead[0]) * self.period:] = 1
        return z

    def zeros(self):
        return np.zeros(self.n_ticks, dtype=np.uint8)


def split_track_probes(counts, live_rows=None, spike_mode="auto"):
    """Track-6 layout (rw_mesh.py:386-390, :613-615, :685): ``[super_probes, mesh_probes]`` with
    ``super_probes = [['Supervisor Global Spike', P], ['Supervisor Fire Spike', P]]`` and
    ``mesh_probes[j] = ['Mesh Node: j', P]``; feed to ``rw_analysis.process_outputs_split``."""
    lay = RasterLayout(counts, spike_mode)
    live = lay.n_nodes if live_rows is None else int(live_rows)
    sup_rows = [lay.boundary, lay.boundary]                   # start / end of the supervisor pulse
    fire_rows = [lay.boundary, lay.zeros, lambda: lay.empty_from(live)]
    super_probes = [
        ['Supervisor Global Spike', Probe(LazyRaster(2, lay.n_ticks, lambda k: sup_rows[k]()))],
        ['Supervisor Fire Spike', Probe(LazyRaster(3, lay.n_ticks, lambda k: fire_rows[k]()))],
    ]

    def mesh_rows(j):
        return lambda k: (lay.boundary() if k == 0 else lay.spikes(j) if k == 1 else lay.zeros())

    mesh_probes = [['Mesh Node: ' + str(j), Probe(LazyRaster(3, lay.n_ticks, mesh_rows(j)))]
                   for j in range(lay.n_nodes)]
    return [super_probes, mesh_probes], lay


def flat_track_probes(counts, probe_start=3, probe_stride=5, spike_mode="auto"):
    """Track-2/5 layout (rw_analysis.py:63-69): a flat list, supervisor probes first, then
    ``probe_stride`` entries per mesh point of which the first is the counter spike probe; feed to
    ``rw_analysis.process_outputs``."""
    lay = RasterLayout(counts, spike_mode)
    sup = [Probe(LazyRaster(2, lay.n_ticks, lambda k: lay.boundary()))]
    out = [sup for _ in range(probe_start)]

    def mesh_rows(j):
        return lambda k: (lay.boundary() if k == 0 else lay.spikes(j) if k == 1 else lay.zeros())

    for j in range(lay.n_nodes):
        out.append([Probe(LazyRaster(3, lay.n_ticks, mesh_rows(j)))])
        out.extend([sup] * (probe_stride - 1))
    return out, lay


def _digits(value):
    v = int(value)
    return [[v % 1000], [(v // 1000) % 1000], [v // 1000000]]


def ends_track_probes(totals, n_steps, probe_start=4, probe_stride=5):
    """Track-1/3 layout (rw_analysis.py:11-36): final counter voltages as base-1000 digits; the
    supervisor counter holds 128 x the number of model steps.  ``totals[j]`` = cumulative walker
    count of mesh point j.  Feed to ``rw_analysis.process_outputs_ends`` / ``count_steps``."""
    sup = [Probe(_digits(128 * int(n_steps)))]
    out = [[], sup, []]
    while len(out) < probe_start:
        out.append([])
    for j in range(len(totals)):
        out.append([Probe(_digits(totals[j]))])
        out.extend([[]] * (probe_stride - 1))
    return out
