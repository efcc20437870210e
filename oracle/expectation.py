"""Deterministic M -> infinity limit of the reference's fluence estimate -- TEST INFRASTRUCTURE.

For a walker started on node i the reference accumulates (fluence_mini_app.py:428-438)
    X = sum_{t=1}^{nt-2} w(t) * v * S(k_t) * dt,      w(t) = exp(-SigA*dt*t),
over the nodes k_t < meshSize it visits, and rwSol = mean of X over M walkers.  With
G_t(i) = E[ sum_{tau>t} w(tau) S(k_tau) | k_t = i ] and H_t its second moment,
    G_t = Q (w_{t+1} S + G_{t+1}),
    H_t = Q ((w_{t+1} S)^2 + 2 w_{t+1} S * G_{t+1} + H_{t+1}),      G_T = H_T = 0, T = nt-2,
so E[rwSol[i]] = v*dt*G_0(i) and Var[X_i] = (v*dt)^2 (H_0(i) - G_0(i)^2) exactly (backward
Markov-chain propagation; Q is the reference's own matrix).  Used for the 3-sigma and relative-L2
acceptance tests of the Philox mode.
"""
from __future__ import annotations

import numpy as np
import scipy.sparse as sp


def csr_to_scipy(indptr, cols, vals, n_rows):
    return sp.csr_matrix((vals, cols, indptr), shape=(n_rows, n_rows))


def fluence_moments(indptr, cols, vals, n_rows, n_live, srcvals, nt, SigA, dt, v):
    """Returns (mean[n_live], var[n_live]) of the per-walker score X_i (rwSol units)."""
    Q = csr_to_scipy(indptr, cols, vals, n_rows)
    S = np.zeros(n_rows)
    S[:n_live] = srcvals[:n_live]
    G = np.zeros(n_rows)
    H = np.zeros(n_rows)
    T = nt - 2
    for t in range(T - 1, -1, -1):
        w = np.exp(-SigA * dt * (t + 1))
        wS = w * S
        Hn = Q @ (wS * wS + 2.0 * wS * G + H)
        Gn = Q @ (wS + G)
        G, H = Gn, Hn
        G[n_live:] = 0.0       # absorbing rows: never stepped
        H[n_live:] = 0.0
    scale = v * dt
    mean = scale * G[:n_live]
    var = scale * scale * np.maximum(H[:n_live] - G[:n_live] ** 2, 0.0)
    return mean, var


def expected_steps(indptr, cols, vals, n_rows, n_live, nt):
    """E[number of uniform draws] per walker started on each live node (walker-steps metric)."""
    Q = csr_to_scipy(indptr, cols, vals, n_rows)
    alive = np.zeros(n_rows)
    alive[:n_live] = (np.diff(indptr)[:n_live] > 0).astype(float)
    L = np.zeros(n_rows)
    for _ in range(nt - 1):
        L = alive * (1.0 + Q @ L)
    return L[:n_live]


def occupancy_expectation(indptr, cols, vals, n_rows, n_live, start_dist, nt):
    """E[count[k, t]] for t < nt given expected initial counts start_dist[k] (dense [nt, n_rows])."""
    QT = csr_to_scipy(indptr, cols, vals, n_rows).T.tocsr()
    out = np.zeros((nt, n_rows))
    cur = np.asarray(start_dist, dtype=np.float64).copy()
    out[0] = cur
    for t in range(1, nt):
        live = cur.copy()
        live[n_live:] = 0.0
        cur = QT @ live
        out[t] = cur
    return out


def fluence_error_stats(per_start, mean, var, M, unit, min_expected_count=50.0):
    """"fluence err vs pc ref" (BASELINE.json metric) of one Philox-mode solve against the exact M -> infinity
    limit of the reference's chain: relative L2 over all cells, and the per-cell z statistics
    |rwSol - E| / sqrt(Var/M) over the cells whose expected tally M * E / unit is at least
    ``min_expected_count`` (below that the tally is a small count, not a normal variable; those cells stay in
    the L2 figure).  ``unit`` = rwSol units per tallied hit (v * dt * S0).  Cells with zero variance must be
    reproduced exactly."""
    per_start, mean, var = (np.asarray(a, dtype=np.float64) for a in (per_start, mean, var))
    sigma = np.sqrt(var / float(M))
    exact = sigma == 0
    test = (~exact) & (float(M) * mean / float(unit) >= min_expected_count)
    z = np.abs(per_start[test] - mean[test]) / sigma[test]
    return {"M": int(M), "rel_l2": float(np.linalg.norm(per_start - mean) / np.linalg.norm(mean)),
            "max_abs_z": float(z.max()) if z.size else 0.0,
            "frac_within_3sigma": float((z < 3.0).mean()) if z.size else 1.0,
            "cells": int(per_start.size), "cells_z_tested": int(test.sum()), "cells_zero_variance": int(exact.sum()),
            "zero_variance_cells_exact": bool(np.array_equal(per_start[exact], mean[exact])),
            "expected_rel_l2": float(np.sqrt(var.sum() / float(M)) / np.linalg.norm(mean))}
