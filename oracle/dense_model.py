"""Lane-level model of csrc/walk_dense.cuh (K1d) -- TEST INFRASTRUCTURE, never imported by the product.

walk_dense.cuh re-orders the work of the score walk (fresh batches, two survivor rings per warp,
drains per segment) and moves the accounting around; the tallies must not change.  This model executes
the SAME schedule statement by statement in numpy -- one warp at a time, 32 lanes as arrays, the rings as
64-entry arrays indexed modulo 64, 32-bit wrapping arithmetic where the kernel has it -- so that the
schedule, the ring-capacity argument and the accounting can be checked against ``philox_port.walk_port``
on the CPU (tests/test_dense_model.py), before and independently of the GPU parity tests.
"""
from __future__ import annotations

import numpy as np

from .philox_port import philox2x32_10, seed_to_key

ROW_OFF_MASK, ROW_N_MASK, ROW_DETABS, ROW_TERM = 0x7FFF, 0x01FF8000, 0x08000000, 0x10000000
DQ = 64
U32 = 0xFFFFFFFF


def _row_len(w0):
    return (w0 >> 15) & 0x3FF


def _detcls(w0):
    return (w0 >> 25) & 3


class _Warp:
    """Per-warp state of the kernel: per-lane tallies and the two rings."""

    def __init__(self):
        self.x = np.zeros(32, np.int64)          # uint32 in the kernel; wrapped when flushed
        self.nst = np.zeros(32, np.int64)
        self.nsurv = np.zeros(32, np.int64)
        self.q1 = np.zeros((DQ, 4), np.int64)
        self.q0 = np.zeros((DQ, 4), np.int64)
        self.max_fill = 0


def walk_dense_model(alias, start_nodes, start_counts, nt, seed, *, class_value, walker_range=None,
                     start_id_base=None, range_len=96, n_warps=3):
    """Returns dict(score[n_starts], counters) computed with the schedule of walk_dense_kernel.
    ``alias``: tables.AliasTables (compact layout).  ``class_value``: 4 integers in [0, 255]."""
    assert alias.layout == "compact"
    rowinfo = alias.rowinfo.astype(np.int64)
    th = alias.alias[:, 0].astype(np.int64)
    own = (alias.alias[:, 1] & 0xFFFF).astype(np.int64)
    oth = (alias.alias[:, 1] >> 16).astype(np.int64)
    cw = np.asarray(class_value, dtype=np.int64)
    start_nodes = np.asarray(start_nodes, dtype=np.int64)
    n_starts = len(start_nodes)
    counts = np.broadcast_to(np.asarray(start_counts, dtype=np.int64), (n_starts,))
    prefix = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    lo, hi = (0, int(prefix[-1])) if walker_range is None else walker_range
    key = seed_to_key(seed)
    t_last = nt - 1
    assert t_last >= 1
    lane = np.arange(32)
    score = np.zeros(n_starts, np.int64)
    counters = dict(steps=0, exits=0, survivors=0, walkers=0)
    warps = [_Warp() for _ in range(n_warps)]
    work_next = 0
    turn = 0

    def step(w, act, ra, rb, word, t):
        """The kernel's step lambda on the active lanes; returns (alive, knew)."""
        v = word.astype(np.uint64) * _row_len(ra).astype(np.uint64)
        tn = t + 1
        e = (ra & ROW_OFF_MASK) + (v >> np.uint64(32)).astype(np.int64)
        e = np.where(act, e, 0)
        coin = (v & np.uint64(U32)).astype(np.int64)
        knew = rb + np.where(coin <= th[e], own[e], oth[e])
        knew = np.where(act, knew, 0)
        row = rowinfo[knew, 0]
        w.x += np.where(act, cw[row >> 30], 0)
        terminal = (row & ROW_TERM) != 0
        ends = act & (terminal | (tn >= t_last))
        common = ends & (tn + 1 < t_last)
        detabs = (row & ROW_DETABS) != 0
        w.x += np.where(common & detabs, cw[_detcls(row)], 0)
        w.nst += np.where(common, tn + np.where(detabs, 1, 0), 0)
        rare = ends & ~common
        if rare.any():
            tl = tn.copy()
            absorbed = (row & ROW_N_MASK) == 0
            wl = cw[row >> 30].copy()
            det_exit = rare & ~absorbed & (tn < t_last)
            wl = np.where(det_exit, cw[_detcls(row)], wl)
            w.x += np.where(det_exit, wl, 0)
            tl = np.where(det_exit, tn + 1, tl)
            absorbed = absorbed | det_exit
            out = rare & (tl >= t_last)
            w.x -= np.where(out, wl, 0)
            w.nsurv += np.where(out & ~absorbed, 1, 0)
            w.nst += np.where(rare, tl, 0)
        return act & ~ends, knew

    def push(w, q, tail, alive, c0, w1, k, t):
        rank = np.cumsum(alive) - alive
        idx = (tail + rank[alive]) % DQ
        q[idx, 0], q[idx, 1], q[idx, 2], q[idx, 3] = c0[alive], w1[alive], k[alive], t[alive]
        return tail + int(alive.sum())

    while True:
        # warps take ranges round-robin (any assignment gives the same tallies)
        nb = lo + work_next
        if nb >= hi:
            break
        work_next += range_len
        w = warps[turn % n_warps]
        turn += 1
        end = min(range_len, hi - nb)
        counters["walkers"] += end
        cur = 0
        s = -1
        s_end = 0
        while cur < end:
            g = nb + cur
            if s < 0 or g >= s_end:
                s = int(np.searchsorted(prefix, g, side="right") - 1)
                s_end = int(prefix[s + 1])
                srow = rowinfo[start_nodes[s]]
                sid_off = 0 if start_id_base is None else int(start_id_base[s]) - int(prefix[s])
            seg_end = min(end, s_end - nb)
            sid0 = g + sid_off
            c0_first = sid0 & U32
            chi = ((sid0 >> 32) << 24) & U32
            to_wrap = (-c0_first) & U32
            if to_wrap != 0 and to_wrap < seg_end - cur:
                seg_end = cur + to_wrap
            n_seg = seg_end - cur
            assert n_seg >= 1
            if _row_len(srow[0]) == 0:
                cur = seg_end
                continue
            if srow[0] & ROW_DETABS:
                k = (n_seg - lane + 31) >> 5
                if t_last >= 2:
                    w.x += cw[_detcls(srow[0])] * k
                w.nst += k
            else:
                h1 = t1 = h0 = t0 = 0
                fresh = cur
                while True:
                    n1, n0 = t1 - h1, t0 - h0
                    w.max_fill = max(w.max_fill, n1, n0)
                    assert n1 <= DQ and n0 <= DQ, "ring overflow"
                    if n1 >= 32 or (n0 < 32 and fresh >= seg_end and n1 != 0):
                        cnt = min(n1, 32)
                        act = lane < cnt
                        e = w.q1[(h1 + lane) % DQ]
                        row = rowinfo[np.where(act, e[:, 2], 0)]
                        alive, knew = step(w, act, row[:, 0], row[:, 1], e[:, 1], e[:, 3])
                        h1 += cnt
                        t0 = push(w, w.q0, t0, alive, e[:, 0], np.zeros(32, np.int64), knew, e[:, 3] + 1)
                    elif n0 >= 32 or (fresh >= seg_end and n0 != 0):
                        cnt = min(n0, 32)
                        act = lane < cnt
                        e = w.q0[(h0 + lane) % DQ]
                        row = rowinfo[np.where(act, e[:, 2], 0)]
                        tn = e[:, 3] + 1
                        assert (tn[act] & 1).all()
                        r0, r1 = philox2x32_10(e[:, 0], chi | tn, key)
                        alive, knew = step(w, act, row[:, 0], row[:, 1], r0.astype(np.int64), e[:, 3])
                        h0 += cnt
                        t1 = push(w, w.q1, t1, alive, e[:, 0], r1.astype(np.int64), knew, tn)
                    elif fresh < seg_end:
                        cnt = min(seg_end - fresh, 32)
                        act = lane < cnt
                        c0 = (c0_first + (fresh - cur) + lane) & U32
                        r0, r1 = philox2x32_10(c0, np.full(32, chi | 1), key)
                        ra = np.full(32, srow[0])
                        rb = np.full(32, srow[1])
                        alive, knew = step(w, act, ra, rb, r0.astype(np.int64), np.zeros(32, np.int64))
                        fresh += cnt
                        t1 = push(w, w.q1, t1, alive, c0, r1.astype(np.int64), knew, np.ones(32, np.int64))
                    else:
                        break
            cur = seg_end
            xs = int(w.x.sum()) & U32            # REDUX.SUM over the uint32 lanes
            w.x[:] = 0
            score[s] += xs
    for w in warps:
        counters["steps"] += int(w.nst.sum())
        counters["survivors"] += int(w.nsurv.sum())
    counters["exits"] = counters["walkers"] - counters["survivors"]
    counters["max_ring_fill"] = max(w.max_fill for w in warps)
    return dict(score=score, counters=counters)
