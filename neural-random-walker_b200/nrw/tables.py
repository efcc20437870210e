"""Per-row sampling tables for the device: cumulative (cdf) and alias, with row-pattern dedup.

The reference samples a transition with ``np.searchsorted(np.cumsum(myQ,axis=1)[j,:], u)``
(fluence_mini_app.py:224,244).  The device keeps two equivalent forms of each row:

* ``cum``   -- the sequential fp64 prefix sums over the row's non-zeros (bit-equal to the dense
  ``np.cumsum`` at the non-zero columns).  Used by the replay kernel and the ``cdf`` sampler.
* alias tables (Walker/Vose) -- O(1) draws for the Philox walk.  Rows of the fluence matrix are
  translates of a few patterns (same relative columns, same probabilities), so patterns are
  de-duplicated and each row keeps only ``(pattern offset, length, column base, source class)``:
  the 10 001-node mesh needs 80 KB of patterns + 80 KB of row descriptors, which is what lets the
  kernel stage everything in one SM's shared memory.

Quantisation: the alias coin is compared with a 32-bit threshold.  On the compact layout slot and
coin are the high and low word of ``w * n`` for ONE 32-bit Philox word ``w`` (csrc/common.cuh, "paired
draws"): both deviate from uniform by less than n/2^32; on the wide layout the slot is
``mulhi(r0, n)`` and the coin a second word (resolved to 2^-32).  All orders of magnitude below the
Monte-Carlo error of any feasible run and stated in DESIGN.md.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .mesh import TransitionCSR

COMPACT_MAX_N = 1023
COMPACT_MAX_REL = 65535
COMPACT_MAX_ENTRIES = 1 << 15          # offsets are 15 bits: what one SM's shared memory can hold anyway
# compact row descriptor (csrc/common.cuh): w0 = off[0:15) | n[15:25) | detabs[25] | term[26] | cls[27:29) | inc[29:32),
# w1 = base[0:29) | detlast[29] | sinc[30:32)
ROW_N_SHIFT, ROW_DETABS_BIT, ROW_TERM_BIT, ROW_CLS_SHIFT, ROW_INC_SHIFT = 15, 25, 26, 27, 29
ROW_BASE_MASK, ROW_DETLAST_BIT, ROW_SINC_SHIFT = 0x1FFFFFFF, 29, 30
ROW_CLASS_UNIT = (1 << ROW_CLS_SHIFT) + (1 << ROW_INC_SHIFT)      # what one unit of source class adds to w0


def row_cum(csr: TransitionCSR) -> np.ndarray:
    """Sequential fp64 prefix sums per row (host form of kernel K4 / fluence_mini_app.py:224)."""
    cum = np.empty_like(csr.vals)
    lens = csr.row_lengths()
    for ln in np.unique(lens):
        if ln == 0:
            continue
        rows = np.nonzero(lens == ln)[0]
        at = csr.indptr[rows][:, None] + np.arange(ln)[None, :]
        cum[at] = np.cumsum(csr.vals[at], axis=1)
    return cum


def _vose(p: np.ndarray):
    """Walker/Vose alias table of a probability vector (normalised here).  Returns
    (prob[n] in [0,1], alias[n])."""
    n = len(p)
    q = p / p.sum() * n
    prob = np.ones(n)
    alias = np.arange(n)
    small = [i for i in range(n) if q[i] < 1.0]
    large = [i for i in range(n) if q[i] >= 1.0]
    q = q.copy()
    while small and large:
        s = small.pop()
        g = large.pop()
        prob[s] = q[s]
        alias[s] = g
        q[g] = (q[g] + q[s]) - 1.0
        (small if q[g] < 1.0 else large).append(g)
    for i in small + large:            # numerical leftovers: probability 1
        prob[i] = 1.0
        alias[i] = i
    return prob, alias


@dataclass
class AliasTables:
    layout: str                 # "compact" | "wide"
    rowinfo: np.ndarray         # uint32 [n_rows, 2] (compact) / [n_rows, 4] (wide)
    alias: np.ndarray           # uint32 [n_alias, 2] (compact) / [n_alias, 4] (wide)
    n_patterns: int
    pattern_of_row: np.ndarray  # int64 [n_rows], -1 for empty rows

    @property
    def n_alias(self) -> int:
        return int(self.alias.shape[0])

    def structured(self):
        """``structure_rows(self)`` computed once: (rowinfo_str, threshold, common row length) or None."""
        if not hasattr(self, "_structured"):
            self._structured = structure_rows(self)
        return self._structured


@dataclass
class FactoredRows:
    """Row descriptors of a translation-invariant mesh, factored so that they need no per-row table
    (walk_philox.cuh::fac_row): live row r = i + j * period (direction i, cell j) is

    * a *stay* row when 0 <= j + shift[i] < cells: w0 = dir_w0[i] + class * ROW_CLASS_UNIT, w1 = (j + shift[i]) * period | 1 << 30
    * an *exit* row otherwise:                      w0 = exit_w0   + class * ROW_CLASS_UNIT, w1 = exit_base (with its flag bits)

    and rows >= n_live are kept verbatim in ``tail``.  ``cls_bits`` packs the 2-bit source classes of
    the live rows 16 to a word; ``k // period == (k * magic) >> shift`` for every k < n_live."""
    period: int
    cells: int
    dirtab: np.ndarray      # uint32 [period, 2]: (dir_w0[i], shift[i] as two's complement)
    exit_w0: int
    exit_base: int
    tail: np.ndarray        # uint32 [n_rows - n_live, 2]
    cls_bits: np.ndarray    # uint32 [ceil(n_live / 16)]
    magic: int
    shift: int

    def synthesise(self, n_live: int) -> np.ndarray:
        """The uint32 [n_live, 2] descriptors the device derives -- equal to the table they replace."""
        k = np.arange(n_live, dtype=np.uint64)
        j = ((k * np.uint64(self.magic)) >> np.uint64(self.shift)).astype(np.int64)
        i = k.astype(np.int64) - j * self.period
        jn = j + self.dirtab[i, 1].astype(np.int32).astype(np.int64)
        out_of_mesh = (jn < 0) | (jn >= self.cells)
        kk = k.astype(np.int64)
        cls = (self.cls_bits[kk >> 4] >> ((kk & 15) * 2).astype(np.uint32)) & np.uint32(3)
        w0 = (np.where(out_of_mesh, np.uint32(self.exit_w0), self.dirtab[i, 0]).astype(np.uint64)
              + cls.astype(np.uint64) * np.uint64(ROW_CLASS_UNIT)) & np.uint64(0xFFFFFFFF)
        base = np.where(out_of_mesh, self.exit_base, (jn * self.period) | (1 << ROW_SINC_SHIFT))
        return np.stack([w0.astype(np.uint32), base.astype(np.uint32)], axis=1)


def structure_rows(at: "AliasTables"):
    """The alias tables of a fluence matrix have a closed form: a stay row keeps the walker's direction with
    probability 1 - pone and else scatters it uniformly over the ND-1 directions of the new cell
    (fluence_mini_app.py:368-371), so Vose's construction pairs every slot with the row's own direction `same`:
    entry s = (T, own = s, alias = same) with ONE threshold T for the whole mesh, and entry `same` always yields
    `same`.  A draw (slot, coin) then needs no table:  target = base + (coin <= T ? slot : same).

    Returns ``(rowinfo_str, T, n_draw)`` -- the compact row descriptors with `same` in the 15-bit pattern-offset
    field, the threshold, and the ONE length of every row a walker can draw from when rows with a single absorbing
    successor are stepped without a draw (all rows that carry neither the `term` bit nor length 0: the stay rows of
    the fluence matrix all hold ND-1 entries; nrw_tables.str_n) -- or None.  Nothing is assumed: every entry of every
    pattern is compared with the closed form and the lengths are compared, so the structured draw returns the alias
    draw's target for EVERY (slot, coin) -- the same trajectories bit for bit."""
    if at.layout != "compact":
        return None
    ri = at.rowinfo
    w0 = ri[:, 0]
    n = ((w0 >> np.uint32(ROW_N_SHIFT)) & np.uint32(0x3FF)).astype(np.int64)
    off = (w0 & np.uint32(0x7FFF)).astype(np.int64)
    th = at.alias[:, 0].astype(np.int64)
    own = (at.alias[:, 1] & np.uint32(0xFFFF)).astype(np.int64)
    ali = (at.alias[:, 1] >> np.uint32(16)).astype(np.int64)
    live = n > 0
    if not live.any():
        return None
    T = None
    same_of = {}
    for o, k in set(zip(off[live].tolist(), n[live].tolist())):         # one check per (pattern, length)
        e = slice(o, o + k)
        if o + k > len(th) or not np.array_equal(own[e], np.arange(k)):
            return None
        certain = (th[e] == 0xFFFFFFFF) | (ali[e] == own[e])              # entries whose outcome is their own slot
        if k == 1:
            if not certain[0]:
                return None
            same_of[(o, k)] = 0
            continue
        cand = np.nonzero(certain)[0]
        if len(cand) != 1:
            return None
        same = int(cand[0])
        rest = np.arange(k) != same
        if not (ali[e][rest] == same).all():
            return None
        ts = np.unique(th[e][rest])
        if len(ts) != 1 or (T is not None and int(ts[0]) != T):
            return None
        T = int(ts[0])
        same_of[(o, k)] = same
    if T is None:
        T = 0xFFFFFFFF                                                    # only single-successor rows: no coin matters
    draws = live & (((w0 >> np.uint32(ROW_TERM_BIT)) & np.uint32(1)) == 0)  # rows a walker is alive on (DET walk)
    lens = np.unique(n[draws])
    if len(lens) > 1:
        return None                                                       # the walk multiplies by ONE length
    n_draw = int(lens[0]) if len(lens) else 1
    same = np.zeros(len(w0), dtype=np.uint32)
    key = off * 1024 + n                                                  # one key per (pattern, length)
    keys, inv = np.unique(key[live], return_inverse=True)
    same[live] = np.array([same_of[(int(k) // 1024, int(k) % 1024)] for k in keys], dtype=np.uint32)[inv]
    # The walk carries a walker's direction instead of its descriptor word (walk_score.cuh, STR && DET): the offset a
    # walker is sent to must BE the `same` of the row it lands on, for every row it can draw from next -- on the
    # fluence matrix row = first row of the cell + direction (fluence_mini_app.py:331-381).  Checked for every
    # (row, slot) whose target is such a row.
    base = (ri[:, 1] & np.uint32(ROW_BASE_MASK)).astype(np.int64)
    src = np.nonzero(draws)[0]
    slots = np.arange(n_draw, dtype=np.int64)
    same64 = same.astype(np.int64)
    block = max(1, (1 << 22) // max(n_draw, 1))                           # ~4 M (row, slot) pairs at a time
    for lo in range(0, len(src), block):
        tgt = base[src[lo:lo + block]][:, None] + slots[None, :]
        if tgt.max() >= len(w0):
            return None
        onward = draws[tgt]
        if not np.array_equal(same64[tgt][onward], np.broadcast_to(slots[None, :], tgt.shape)[onward]):
            return None
    out = ri.copy()
    out[:, 0] = (w0 & ~np.uint32(0x7FFF)) | same
    return np.ascontiguousarray(out), T, n_draw


def factorise_rows(at: "AliasTables", n_live: int, rowinfo=None):
    """FactoredRows of compact alias tables whose row descriptors have the (direction, cell)
    structure of the fluence mesh (fluence_mini_app.py:331-381: row r = i + j (ND-1), a stay row
    reaches all directions of one other cell, an exit row reaches the sink), or None.  Nothing is
    assumed: the factored form is accepted only if it reproduces every descriptor bit for bit.
    ``rowinfo``: another descriptor array of the same tables (the structured one, ``structure_rows``)."""
    if at.layout != "compact":
        return None
    ri = at.rowinfo if rowinfo is None else rowinfo
    n_rows = ri.shape[0]
    n_tail = n_rows - n_live
    if n_live < 2 or not (1 <= n_tail <= 64):
        return None
    w0 = ri[:n_live, 0]
    w1 = ri[:n_live, 1].astype(np.int64)
    base = w1 & ROW_BASE_MASK
    cls_all = (w0 >> np.uint32(ROW_CLS_SHIFT)) & np.uint32(3)
    body = (w0.astype(np.int64) - cls_all.astype(np.int64) * ROW_CLASS_UNIT).astype(np.uint32)   # w0 without its class
    n = (w0 >> np.uint32(ROW_N_SHIFT)) & np.uint32(0x3FF)
    wide = n > 1
    if not wide.any():
        return None
    period = int(np.bincount(n[wide]).argmax())
    if period < 2 or n_live % period:
        return None
    cells = n_live // period
    s = int(np.ceil(np.log2(period)))
    L = int(n_live - 1).bit_length()
    shift = L + s
    if shift > 31:
        return None
    magic = -((-(1 << shift)) // period)                    # ceil(2^shift / period)
    if magic >= 1 << 32:
        return None
    r = np.arange(n_live, dtype=np.int64)
    i, j = r % period, r // period
    # exit rows: the most common (body, base) pair among the rows with at most one entry
    narrow = ~wide
    exit_w0 = exit_base = 0
    is_exit = np.zeros(n_live, dtype=bool)
    if narrow.any():
        key = (body[narrow].astype(np.uint64) << np.uint64(32)) | w1[narrow].astype(np.uint64)
        vals, cnts = np.unique(key, return_counts=True)
        top = int(vals[cnts.argmax()])
        exit_w0, exit_base = top >> 32, top & 0xFFFFFFFF              # exit_base keeps its flag bits
        is_exit = (body == np.uint32(exit_w0)) & (w1 == exit_base)
    stay = ~is_exit
    if (base[stay] % period).any():
        return None
    dirtab = np.zeros((period, 2), dtype=np.uint32)
    sh_all = base // period - j
    shifts = np.full(period, cells, dtype=np.int64)           # directions without a stay row always exit
    first = np.full(period, -1, dtype=np.int64)
    idx = np.nonzero(stay)[0]
    first[i[idx][::-1]] = idx[::-1]                           # first stay row of each direction
    have = first >= 0
    shifts[have] = sh_all[first[have]]
    dirtab[have, 0] = body[first[have]]
    if np.abs(shifts).max() >= 1 << 30:
        return None
    dirtab[:, 1] = shifts.astype(np.int32).view(np.uint32)
    cls = cls_all.astype(np.uint32)
    words = (n_live + 15) // 16
    padded = np.zeros(words * 16, dtype=np.uint32)
    padded[:n_live] = cls
    cls_bits = np.zeros(words, dtype=np.uint32)
    for b in range(16):
        cls_bits |= padded[b::16] << np.uint32(2 * b)
    fr = FactoredRows(period, cells, dirtab, int(exit_w0), int(exit_base), np.ascontiguousarray(ri[n_live:]),
                      cls_bits, int(magic), int(shift))
    if not np.array_equal(fr.synthesise(n_live), ri[:n_live]):
        return None
    return fr


def _unique_rows(key: np.ndarray):
    """(index of the first row of each distinct row, group index of every row) of a uint64 matrix.
    Rows are grouped by a 64-bit multiplicative hash (a lexicographic sort of a 158 760 x 252 key
    matrix takes 20 s); every group is then checked against its representative, and the exact sort
    is the fallback should two different rows ever collide."""
    mult = np.random.default_rng(0x5eed).integers(1, 2 ** 63, size=key.shape[1], dtype=np.uint64) | np.uint64(1)
    with np.errstate(over="ignore"):
        h = (key * mult[None, :]).sum(axis=1, dtype=np.uint64)
        h ^= h >> np.uint64(29)
    _, first, inv = np.unique(h, return_index=True, return_inverse=True)
    inv = inv.reshape(-1)
    if not np.array_equal(key[first][inv], key):               # hash collision
        _, first, inv = np.unique(key, axis=0, return_index=True, return_inverse=True)
        inv = inv.reshape(-1)
    return first, inv


def dedup_rows(csr: TransitionCSR):
    """Group rows whose (relative columns, values) coincide.  Returns (pattern_of_row,
    representative_row_of_pattern, base_col_of_row)."""
    lens = csr.row_lengths()
    n = csr.n_rows
    pat = np.full(n, -1, dtype=np.int64)
    base = np.zeros(n, dtype=np.int64)
    reps = []
    for ln in np.unique(lens):
        if ln == 0:
            continue
        rows = np.nonzero(lens == ln)[0]
        at = csr.indptr[rows][:, None] + np.arange(ln)[None, :]
        c = csr.cols[at].astype(np.int64)
        b = c[:, 0]
        rel = c - b[:, None]
        base[rows] = b
        key = np.concatenate([rel.astype(np.uint64), csr.vals[at].view(np.uint64)], axis=1)
        first, inv = _unique_rows(key)
        pat[rows] = len(reps) + inv
        reps.extend(rows[first].tolist())
    return pat, np.array(reps, dtype=np.int64), base


def build_alias(csr: TransitionCSR, node_class: np.ndarray | None = None,
                layout: str = "auto") -> AliasTables:
    """Alias tables + row descriptors in the device layouts of include/nrw.h."""
    n = csr.n_rows
    lens = csr.row_lengths()
    if node_class is None:
        node_class = np.zeros(n, dtype=np.uint8)
    node_class = np.asarray(node_class).astype(np.uint32)
    if node_class.shape != (n,) or node_class.max(initial=0) > 3:
        raise ValueError("node_class must hold one class 0..3 per row")
    pat, reps, base = dedup_rows(csr)

    # alias tables per exact pattern, then a second dedup on the QUANTISED tables: rows that differ
    # only in the last fp64 bits of the reference's pairwise normalisation (54 bit
    # patterns for 31 structural ones at the defaults) draw from identical 32-bit tables.
    merged: dict = {}
    merged_of = np.zeros(len(reps), dtype=np.int64)
    th_all, a_all, b_all, sizes = [], [], [], []
    max_rel = 0
    for pi, r in enumerate(reps):
        lo, hi = int(csr.indptr[r]), int(csr.indptr[r + 1])
        rel = csr.cols[lo:hi].astype(np.int64) - int(csr.cols[lo])
        prob, alias = _vose(csr.vals[lo:hi])
        k = np.rint(prob * 4294967296.0).astype(np.int64)      # P(own) = k / 2^32
        k = np.clip(k, 0, 4294967296)
        a = rel.copy()
        b = rel[alias]
        never = k == 0
        a[never] = b[never]                                    # coin can never pick "own"
        th = np.where(never, 0, k - 1).astype(np.uint32)       # own  <=>  r1 <= th
        key = (th.tobytes(), a.tobytes(), b.tobytes())
        mi = merged.get(key)
        if mi is None:
            mi = len(sizes)
            merged[key] = mi
            th_all.append(th); a_all.append(a); b_all.append(b); sizes.append(hi - lo)
            max_rel = max(max_rel, int(rel.max()))
        merged_of[pi] = mi
    offs = np.zeros(len(sizes) + 1, dtype=np.int64)
    offs[1:] = np.cumsum(sizes)
    pat = np.where(pat >= 0, merged_of[np.maximum(pat, 0)], -1)
    n_patterns = len(sizes)
    n_alias = int(offs[-1])
    th_all = np.concatenate(th_all) if th_all else np.zeros(0, np.uint32)
    a_all = np.concatenate(a_all) if a_all else np.zeros(0, np.int64)
    b_all = np.concatenate(b_all) if b_all else np.zeros(0, np.int64)

    fits = (n_alias <= COMPACT_MAX_ENTRIES and int(lens.max(initial=0)) <= COMPACT_MAX_N
            and max_rel <= COMPACT_MAX_REL)
    if layout == "auto":
        layout = "compact" if fits else "wide"
    if layout == "compact" and not fits:
        raise ValueError("matrix does not fit the compact alias layout "
                         "(entries=%d, max row=%d, max relative column=%d)"
                         % (n_alias, int(lens.max(initial=0)), max_rel))

    row_off = np.where(pat >= 0, offs[np.maximum(pat, 0)], 0).astype(np.uint32)
    row_n = lens.astype(np.uint32)
    row_n[csr.n_live:] = 0                                     # rows >= n_live are never stepped
    row_base = base.astype(np.uint32)
    if layout == "compact":
        # deterministic exits: one successor, and that successor is never stepped (absorbing)
        first = csr.cols[np.minimum(csr.indptr[:-1], max(csr.nnz - 1, 0))].astype(np.int64) if csr.nnz else np.zeros(n, np.int64)
        detabs = (row_n == 1) & (row_n[first] == 0)
        detcls = np.where(detabs, node_class[first], 0).astype(np.uint32)
        detlast = detabs & (first == n - 1)                    # ... and it is the last row (the fluence sink)
        if int(row_base.max(initial=0)) > ROW_BASE_MASK:
            raise ValueError("compact layout: column bases must fit 29 bits")
        # inc / sinc: what ARRIVING on the row adds to score and step count when rows with an absorbing successor
        # are stepped without a draw (csrc/common.cuh): inc = cls + class of that successor, sinc = 1 + detabs
        w0 = (row_off | (row_n << np.uint32(ROW_N_SHIFT))
              | (detabs.astype(np.uint32) << np.uint32(ROW_DETABS_BIT))
              | ((detabs | (row_n == 0)).astype(np.uint32) << np.uint32(ROW_TERM_BIT))
              | (node_class << np.uint32(ROW_CLS_SHIFT))
              | ((node_class + detcls) << np.uint32(ROW_INC_SHIFT)))
        w1 = (row_base | (detlast.astype(np.uint32) << np.uint32(ROW_DETLAST_BIT))
              | ((np.uint32(1) + detabs.astype(np.uint32)) << np.uint32(ROW_SINC_SHIFT)))
        rowinfo = np.stack([w0, w1], axis=1).astype(np.uint32)
        ab = (a_all.astype(np.uint32) & np.uint32(0xFFFF)) | (b_all.astype(np.uint32) << np.uint32(16))
        alias = np.stack([th_all, ab], axis=1).astype(np.uint32)
        if alias.shape[0] == 0:
            alias = np.zeros((1, 2), np.uint32)
    else:
        rowinfo = np.stack([row_off, row_n, row_base, node_class], axis=1).astype(np.uint32)
        alias = np.stack([th_all, a_all.astype(np.uint32), b_all.astype(np.uint32),
                          np.zeros(n_alias, np.uint32)], axis=1).astype(np.uint32)
        if alias.shape[0] == 0:
            alias = np.zeros((1, 4), np.uint32)
    return AliasTables(layout=layout, rowinfo=np.ascontiguousarray(rowinfo),
                       alias=np.ascontiguousarray(alias), n_patterns=n_patterns,
                       pattern_of_row=pat)


def alias_row_distribution(csr: TransitionCSR, at: AliasTables, row: int) -> dict:
    """Exact distribution the device draws from for ``row`` (for tests): {column: probability}."""
    if at.layout == "compact":
        w0, base = int(at.rowinfo[row, 0]), int(at.rowinfo[row, 1]) & ROW_BASE_MASK
        off, n = w0 & 0x7FFF, (w0 >> ROW_N_SHIFT) & 0x3FF
    else:
        off, n, base = (int(x) for x in at.rowinfo[row, :3])
    out: dict = {}
    for s in range(n):
        e = at.alias[off + s]
        if at.layout == "compact":
            th, a, b = int(e[0]), int(e[1]) & 0xFFFF, int(e[1]) >> 16
        else:
            th, a, b = int(e[0]), int(e[1]), int(e[2])
        p_own = (th + 1) / 4294967296.0
        out[base + a] = out.get(base + a, 0.0) + p_own / n
        out[base + b] = out.get(base + b, 0.0) + (1.0 - p_own) / n
    return {k: v for k, v in out.items() if v > 0}


def classify_source(srcvals: np.ndarray, n_rows: int):
    """Map per-node source values to <= 3 non-zero classes: (node_class uint8[n_rows],
    class_value float64[4]).  The fused score tally adds class_value[class(node)] per visit."""
    vals = np.unique(srcvals[srcvals != 0])
    if len(vals) > 3:
        raise ValueError("the fused score tally supports at most 3 distinct non-zero source "
                         "values (got %d); use the occupancy tally instead" % len(vals))
    node_class = np.zeros(n_rows, dtype=np.uint8)
    class_value = np.zeros(4)
    for ci, x in enumerate(vals, start=1):
        node_class[:len(srcvals)][srcvals == x] = ci
        class_value[ci] = x
    return node_class, class_value


def likely_paths(csr: TransitionCSR, start_nodes, t_cap: int) -> np.ndarray:
    """int32 [n_starts, t_cap]: the most likely trajectory of every start -- path[s, 0] = the start node,
    path[s, t] = the most probable successor of path[s, t-1] (first one on ties), -1 once the path has reached an
    absorbing row.  Walkers of one start move along it almost in lock step (the fluence matrix keeps a walker's
    direction with probability 1 - pone, fluence_mini_app.py:368-371), which is what makes the time-resolved
    tally contended; the walk tallies walkers on this path through a leave-time histogram instead
    (include/nrw.h, nrw_walk.count_path).  Any choice of path gives the same counts; the most likely one removes
    the most atomics."""
    start_nodes = np.asarray(start_nodes, dtype=np.int64).reshape(-1)
    path = np.full((len(start_nodes), int(t_cap)), -1, dtype=np.int32)
    path[:, 0] = start_nodes
    main: dict = {}

    def successor(r: int) -> int:
        if r < 0 or r >= csr.n_live:
            return -1
        hit = main.get(r)
        if hit is None:
            lo, hi = int(csr.indptr[r]), int(csr.indptr[r + 1])
            hit = main[r] = int(csr.cols[lo + int(np.argmax(csr.vals[lo:hi]))]) if hi > lo else -1
        return hit
    for t in range(1, int(t_cap)):
        for s in range(len(start_nodes)):
            path[s, t] = successor(int(path[s, t - 1]))
    return path


def build_tree_tables(csr: TransitionCSR):
    """The reference's probability circuit (loihirw/rw_core.py:11-156) as a sampler table: for every row the
    firing thresholds floor(128 * p) of its binary tree of Bernoulli neurons -- the 8-bit weights ``128*p + 36`` of
    rw_core.py:151-153 without their offset -- in heap order (root first).  Returns (thresholds uint8[total],
    tree_ptr int32[n_rows + 1]); row r owns n2(r) - 1 entries, n2 = next power of two of its fan-out.  Rows are
    de-duplicated by their probabilities while building (a fluence mesh has a few dozen patterns)."""
    from loihirw.rw_core import build_tree
    lens = csr.row_lengths().astype(np.int64)
    lens[csr.n_live:] = 0
    if lens.max(initial=0) > 512:
        raise ValueError("the tree sampler spends 7 bits per level of one 64-bit draw: rows of <= 512 entries")
    n2 = np.where(lens > 1, 1 << np.ceil(np.log2(np.maximum(lens, 2))).astype(np.int64), 1)
    ptr = np.zeros(csr.n_rows + 1, dtype=np.int64)
    ptr[1:] = np.cumsum(n2 - 1)
    thr = np.zeros(int(ptr[-1]), dtype=np.uint8)
    seen: dict = {}
    for r in range(csr.n_live):
        n = int(lens[r])
        if n <= 1:
            continue
        vals = csr.vals[csr.indptr[r]:csr.indptr[r + 1]]
        key = vals.tobytes()
        heap = seen.get(key)
        if heap is None:
            _, fire_p, _, _, _ = build_tree(list(vals))       # neurons level by level, leaves' parents first, root last
            q = np.floor(128.0 * fire_p + 36.0) - 36.0         # the programmed weight, offset removed: floor(128 p)
            m = int(n2[r])
            heap = np.zeros(m - 1, dtype=np.uint8)
            at, width = 0, m // 2
            while width >= 1:                                  # level with `width` neurons sits at heap [width-1, 2*width-1)
                heap[width - 1:2 * width - 1] = q[at:at + width]
                at += width
                width //= 2
            seen[key] = heap
        thr[ptr[r]:ptr[r + 1]] = heap
    return thr, ptr.astype(np.int32)
