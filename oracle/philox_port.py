"""numpy port of the DEVICE sampler (Philox2x32-10 + alias / cdf draw) -- TEST INFRASTRUCTURE.

This is *not* a restatement of the reference (whose RNG is MT19937, restated.py); it restates
walk_philox.cuh so that the Philox kernel's integer tallies can be checked bit for bit on the
same (seed, walker id) streams, independent of grid shape and sharding.  Statistical agreement
with the reference is checked separately against ``expectation.py``.
"""
from __future__ import annotations

import numpy as np

M2 = np.uint64(0xD256D193)
W32 = np.uint64(0x9E3779B9)
MASK = np.uint64(0xFFFFFFFF)


def philox2x32_10(c0, c1, key):
    """Vectorised Philox2x32-10 (Random123).  uint32 arrays in, (uint32, uint32) out."""
    c0 = np.asarray(c0, dtype=np.uint64) & MASK
    c1 = np.asarray(c1, dtype=np.uint64) & MASK
    k = np.uint64(int(key) & 0xFFFFFFFF)
    for _ in range(10):
        prod = M2 * c0
        hi, lo = prod >> np.uint64(32), prod & MASK
        c0 = (hi ^ k ^ c1) & MASK
        c1 = lo
        k = (k + W32) & MASK
    return c0.astype(np.uint32), c1.astype(np.uint32)


def seed_to_key(seed: int) -> int:
    """splitmix64 finaliser folded to 32 bits (common.cuh::seed_to_key)."""
    m = (1 << 64) - 1
    z = (int(seed) + 0x9E3779B97F4A7C15) & m
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & m
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & m
    z = z ^ (z >> 31)
    return (z ^ (z >> 32)) & 0xFFFFFFFF


def u53(a, b):
    return ((a >> np.uint32(5)).astype(np.float64) * 67108864.0
            + (b >> np.uint32(6)).astype(np.float64)) * (1.0 / 9007199254740992.0)


def _decode_alias(at):
    """AliasTables -> (off[n_rows], n[n_rows], base[n_rows], cls[n_rows], thresh, a, b)."""
    ri = at.rowinfo
    if at.layout == "compact":
        off = (ri[:, 0] & 0x7FFF).astype(np.int64)
        n = ((ri[:, 0] >> 15) & 0x3FF).astype(np.int64)
        cls = ((ri[:, 0] >> 27) & 3).astype(np.int64)          # w0 = off | n << 15 | ... | cls << 27 | inc << 29
        base = (ri[:, 1] & 0x1FFFFFFF).astype(np.int64)       # w1 = base | detlast << 29 | sinc << 30
        th = at.alias[:, 0].astype(np.uint32)
        a = (at.alias[:, 1] & 0xFFFF).astype(np.int64)
        b = (at.alias[:, 1] >> 16).astype(np.int64)
    else:
        off, n, base, cls = (ri[:, i].astype(np.int64) for i in range(4))
        th = at.alias[:, 0].astype(np.uint32)
        a, b = at.alias[:, 1].astype(np.int64), at.alias[:, 2].astype(np.int64)
    return off, n, base, cls, th, a, b


def walk_port(csr, start_nodes, start_counts, nt, seed, *, alias=None, cum=None, node_class=None,
              sampler="alias", walker_range=None, class_value=None, time_weight=None, t_cap=None,
              count_per_start=False, start_id_base=None, tree=None):
    """Simulate every walker of walk_philox.cuh; returns a dict with the same tallies:
    score[n_starts] (int64), score_sq[n_starts] (float64, summed in id order -- compare with
    tolerance), node_hist[n_rows], count_tn[groups, t_cap, n_rows], counters."""
    start_nodes = np.asarray(start_nodes, dtype=np.int64)
    n_starts = len(start_nodes)
    counts = np.broadcast_to(np.asarray(start_counts, dtype=np.int64), (n_starts,))
    prefix = np.concatenate([[0], np.cumsum(counts)])
    lo, hi = (0, int(prefix[-1])) if walker_range is None else walker_range
    g = np.arange(lo, hi, dtype=np.int64)
    s_of = np.searchsorted(prefix, g, side="right") - 1
    key = seed_to_key(seed)
    n_rows, n_live = csr.n_rows, csr.n_live
    if sampler == "alias":
        off, nrow, base, cls, th, aa, bb = _decode_alias(alias)
    else:
        nrow = np.diff(csr.indptr).astype(np.int64)
        nrow[n_live:] = 0
        cls = (np.zeros(n_rows, np.int64) if node_class is None
               else np.asarray(node_class).astype(np.int64))
    cv = np.zeros(4) if class_value is None else np.asarray(class_value, dtype=np.float64)
    t_cap = nt if t_cap is None else min(t_cap, nt)
    groups = n_starts if count_per_start else 1
    out = dict(score=np.zeros(n_starts, np.int64), score_sq=np.zeros(n_starts),
               node_hist=np.zeros(n_rows, np.int64),
               count_tn=np.zeros((groups, t_cap, n_rows), np.int64))
    k = start_nodes[s_of].copy()
    x = np.zeros(len(g), np.int64)
    alive = (nrow[k] > 0) & (nt - 1 >= 1)
    exits = int((nrow[k] == 0).sum())
    surv = int(((nrow[k] > 0) & ~alive).sum())
    steps = 0
    sid = g if start_id_base is None else (np.asarray(start_id_base, np.int64)[s_of] + (g - prefix[s_of]))
    c0 = (sid & 0xFFFFFFFF).astype(np.uint32)
    chi = ((sid >> 32) << 24).astype(np.uint32)
    for t in range(1, nt):
        idx = np.nonzero(alive)[0]
        if len(idx) == 0:
            break
        steps += len(idx)
        paired = sampler == "alias" and alias.layout == "compact"      # common.cuh: "paired draws"
        r0, r1 = philox2x32_10(c0[idx], chi[idx] | np.uint32(((t - 1) | 1) if paired else t), key)
        kk = k[idx]
        if paired:
            v = (r0 if (t & 1) else r1).astype(np.uint64) * nrow[kk].astype(np.uint64)
            e = off[kk] + (v >> np.uint64(32)).astype(np.int64)
            knew = base[kk] + np.where((v & MASK).astype(np.uint32) <= th[e], aa[e], bb[e])
        elif sampler == "alias":
            slot = ((r0.astype(np.uint64) * nrow[kk].astype(np.uint64)) >> np.uint64(32)).astype(np.int64)
            e = off[kk] + slot
            knew = base[kk] + np.where(r1 <= th[e], aa[e], bb[e])
        elif sampler == "tree":
            # walk_philox.cuh TAB_TREE: 7 bits per level of (r0 | r1 << 32), left iff u7 < threshold (heap order)
            thr, ptr = tree
            bits = r0.astype(np.uint64) | (r1.astype(np.uint64) << np.uint64(32))
            n = nrow[kk]
            depth = np.where(n > 1, np.ceil(np.log2(np.maximum(n, 2))).astype(np.int64), 0)
            node = np.zeros(len(idx), np.int64)
            for l in range(int(depth.max(initial=0))):
                act = depth > l
                t7 = thr[np.minimum(ptr[kk].astype(np.int64) + node, max(len(thr) - 1, 0))].astype(np.int64)
                fire = (bits & np.uint64(127)).astype(np.int64) < t7
                node = np.where(act, 2 * node + np.where(fire, 1, 2), node)
                bits = bits >> np.uint64(7)
            leaf = np.minimum(node - ((1 << depth) - 1), n - 1)
            knew = csr.cols[csr.indptr[kk] + leaf].astype(np.int64)
        else:
            u = u53(r0, r1)
            knew = np.empty(len(idx), np.int64)
            for r in np.unique(kk):
                sel = kk == r
                a0, a1 = csr.indptr[r], csr.indptr[r + 1]
                j = np.minimum(np.searchsorted(cum[a0:a1], u[sel], side="left"), a1 - a0 - 1)
                c = csr.cols[a0 + j].astype(np.int64)
                knew[sel] = np.where(u[sel] <= 0.0, 0, c)
        k[idx] = knew
        if t <= nt - 2:
            if time_weight is None:
                x[idx] += cv[cls[knew]].astype(np.int64)
            else:
                x[idx] += np.rint(time_weight[t] * cv[cls[knew]]).astype(np.int64)
            np.add.at(out["node_hist"], knew, 1)
        if t < t_cap:
            gq = s_of[idx] if count_per_start else np.zeros(len(idx), np.int64)
            np.add.at(out["count_tn"], (gq, np.full(len(idx), t), knew), 1)
        absorbed = nrow[knew] == 0
        done = absorbed | (t >= nt - 1)
        exits += int(absorbed.sum())
        surv += int((done & ~absorbed).sum())
        alive[idx[done]] = False
    np.add.at(out["score"], s_of, x)
    np.add.at(out["score_sq"], s_of, x.astype(np.float64) ** 2)
    out["counters"] = dict(steps=steps, exits=exits, survivors=surv, walkers=len(g))
    return out
