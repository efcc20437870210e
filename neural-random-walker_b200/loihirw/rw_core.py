"""``rw_core`` on a B200: the sampling-table builder of the reference's loihirw/rw_core.py.

The reference decomposes each row's categorical distribution into a binary tree of Bernoulli
"probability neurons" (``get_probCircuitValues``, rw_core.py:11-156) and instantiates it in NxSDK
(``probCircuit``, :159-339 -- Loihi circuitry, out of scope).  The GPU walk samples rows with alias
tables instead (``nrw.tables``); this module keeps ``get_probCircuitValues`` -- same argument, same
three return arrays -- for callers that inspect or export the Loihi circuit parameters, plus the
distribution those parameters realise (``tree_distribution``) so that the effect of the 8-bit weight
quantisation ``128*p + 36`` can be compared with the exact row.
"""
import math

import numpy as np


def _normalise(prob):
    """rw_core.py:27-33: >1 -> rescale; <1 -> the deficit goes to output 0 if it is unused,
    otherwise rescale."""
    total = np.sum(prob)
    if total > 1:
        return prob / total
    if total < 1 and prob[0] == 0:
        prob = prob.copy()
        prob[0] = 1 - total
        return prob
    if total < 1:
        return prob / total
    return prob


def build_tree(prob):
    """Complete binary tree over the outputs (padded with zero-probability leaves to a power of
    two).  Returns (prob_padded, fire_p, mass, left_leaves, right_leaves): neuron k fires with
    ``fire_p[k]`` = mass of its left sub-tree / mass of both; level 0 first, root last
    (the neuron order of rw_core.py:47-108)."""
    prob = list(prob)
    n2 = 2 ** math.ceil(math.log(len(prob), 2)) if len(prob) > 1 else 1
    prob = _normalise(np.asarray(prob + [0] * (n2 - len(prob))))
    fire_p, mass, left, right = [], [], [], []
    level = [([i], prob[i]) for i in range(n2)]              # (leaves, mass) per sub-tree
    while len(level) > 1:
        nxt = []
        for a in range(0, len(level), 2):
            (ll, lm), (rl, rm) = level[a], level[a + 1]
            tot = lm + rm
            fire_p.append(lm / tot if tot > 0 else 0.0)
            mass.append(tot)
            left.append(ll)
            right.append(rl)
            nxt.append((ll + rl, tot))
        level = nxt
    return prob, np.asarray(fire_p), np.asarray(mass), left, right


def get_probCircuitValues(prob):
    """Synaptic parameters of the probability circuit for one fan-out (rw_core.py:11-156).

    Returns ``(conn, relay2out_weight, weight)``: ``conn[o, 1 + k]`` is the weight from probability
    neuron k to output o (+120 shared among the neurons that must fire for o, -100 for each that must
    stay silent; column 0 is the relay input), ``relay2out_weight`` marks outputs driven by the relay
    alone, ``weight[k] = 128 * p + 36`` is the 8-bit firing threshold weight of neuron k-1.  Like the
    reference, mutates nothing but may raise IndexError on fan-outs whose pruned circuit is not
    square (e.g. ``[0, .6, .4]``)."""
    prob, p, _, left, right = build_tree(prob)
    n2 = len(prob)
    conn = np.zeros((n2, n2))
    for k in range(len(p)):
        conn[left[k], k + 1] = 1
        conn[right[k], k + 1] = -1
    # prune: neurons whose whole sub-tree is impossible, then impossible outputs (:113-124)
    used = np.array([any(prob[j] > 0 for j in left[k] + right[k]) for k in range(len(p))], dtype=bool)
    conn = conn[:, np.concatenate([[True], used, np.ones(n2 - 1 - len(p), dtype=bool)])]
    p = p[used]
    conn = conn[prob > 0, :]
    # prune neurons that inhibit nobody: their outputs fire on the relay alone (:128-133)
    for i in reversed(range(1, conn.shape[1])):
        if np.sum(conn[conn[:, i] < 0, i]) == 0:
            conn = np.delete(conn, i, 1)
            p = np.delete(p, i - 1)
    n_out, n_col = conn.shape
    for i in range(n_out):                                   # (:139-141)
        ones = conn[i, :] == 1
        conn[i, ones] = 120 / max(np.count_nonzero(ones), 1)
        conn[i, conn[i, :] == -1] = -100
    relay2out_weight = np.zeros((n_col, 1))
    for i in range(n_col):                                   # the reference indexes rows by column count
        if np.max(conn[i, :]) <= 0:
            relay2out_weight[i] = 120
    weight = np.zeros((n_out, 1))
    for i in range(1, n_out):
        weight[i] = 128 * p[i - 1] + 36
    return conn, relay2out_weight, weight


def tree_distribution(prob, quantise=False):
    """Distribution over the outputs realised by walking the tree from the root with one Bernoulli
    draw per level.  ``quantise=True`` rounds every firing probability to the 8-bit weight the
    reference programs (``w = 128 p + 36`` -> ``p' = (floor(w) - 36) / 128``)."""
    prob, p, _, left, right = build_tree(prob)
    if quantise:
        p = (np.floor(128 * p + 36) - 36) / 128
    out = np.ones(len(prob))
    for k in range(len(p)):
        out[left[k]] *= p[k]
        out[right[k]] *= 1 - p[k]
    return out
