"""CPU restatement (numpy) of the reference's `--run_mode pc` path.

TEST INFRASTRUCTURE (see oracle/__init__.py) -- never imported by the product.

Every function cites the reference lines it follows (paths relative to
/root/reference).  Pinning: `tests/test_oracle_golden.py` checks these
functions bit-for-bit against fixtures produced by the real reference
(`oracle/make_golden.py` -> `tests/golden/*.npz`), and, when /root/reference is
present, against the live reference.

Two forms of the hot loop are kept:
  * ``run_pc_literal``  -- the reference's triple loop, line for line in meaning
    (one ``np.random.uniform()`` + one ``np.searchsorted`` per walker-step).  This
    is what `bench.py` times as the CPU baseline (kind "port").
  * ``run_pc_counts``   -- a count-based, vectorised form over CSR rows that
    consumes the *same* uniform stream in the *same* order and therefore returns
    bit-identical ``tempdata`` and leaves the RNG at the same stream position.
    It is the checker for meshes the literal loop cannot finish on.
"""
from __future__ import annotations

import math

import numpy as np


# --------------------------------------------------------------------------
# mesh + transition matrix (fluence_mini_app.py:279-386)
# --------------------------------------------------------------------------

def check_fluence_params(dt, v, SigS):
    """fluence_mini_app.py:283-291 -- reject > 5 % multi-scatter probability."""
    pone = v * SigS * dt * (np.exp(-v * SigS * dt))
    pzero = np.exp(-v * SigS * dt)
    if (1 - pone - pzero) > 0.05:
        raise ValueError("The probability of more than one scattering event occurring "
                         "in a single time step is too large.")
    return pone


def fluence_mesh(dt, da, v, L):
    """fluence_mini_app.py:279, :297-324 -- (space, direction) mesh.

    row = i + j*(ND-1)  <->  (cps[j], directions[i]).
    """
    ds = da * v * dt / 2
    NS = int((2 * L) / ds) + 1
    space = np.linspace(-L, L, NS)
    cps = (space + (ds / 2))[0:(NS - 1)]
    ND = int(2 / da) + 1
    dirs = np.linspace(-1, 1, ND)
    directions = (dirs + (da / 2))[0:(ND - 1)]
    meshSize = (NS - 1) * (ND - 1)
    meshpoints = np.zeros((meshSize, 2))
    meshpoints[:, 0] = np.repeat(cps, ND - 1)
    meshpoints[:, 1] = np.tile(directions, NS - 1)
    return dict(ds=ds, NS=NS, ND=ND, space=space, dirs=dirs, cps=cps,
                directions=directions, meshSize=meshSize, meshpoints=meshpoints)


def fluence_matrix_dense(dt, da, v, SigS, L):
    """fluence_mini_app.py:331-381 -- dense myQ[(meshSize+1)^2], literal fill.

    Kept as a per-row loop with the same numpy calls as the reference (the
    ``np.where(np.round(...)==0)`` cell lookup, the dense ``sum(axis=1)``
    normalisation whose pairwise order fixes the last bits -- SURVEY Q8).
    """
    pone = check_fluence_params(dt, v, SigS)
    m = fluence_mesh(dt, da, v, L)
    ND, cps, directions = m["ND"], m["cps"], m["directions"]
    meshSize, meshpoints = m["meshSize"], m["meshpoints"]
    freshprob = pone / (ND - 1)
    returnprob = freshprob + (1 - pone)
    myQ = np.zeros((meshSize + 1, meshSize + 1))
    myQ[meshSize, meshSize] = 1
    for i in range(0, meshSize):
        newpos = meshpoints[i, 0] - v * meshpoints[i, 1] * dt
        if abs(newpos) > L:
            myQ[i, meshSize] = 1
        else:
            newindex = np.where((np.round(cps - newpos, 8)) == 0)[0][0]
            myQ[i, newindex * (ND - 1):(newindex + 1) * (ND - 1)] = freshprob
            samedir = np.where((np.round(directions - meshpoints[i, 1], 8)) == 0)[0][0]
            myQ[i, samedir + (newindex * (ND - 1))] = returnprob
    myQ = myQ / myQ.sum(axis=1)[:, None]
    myQ[meshSize, meshSize] = 0
    return myQ, m


def dense_to_csr(myQ):
    """CSR view of a dense transition matrix: non-zeros in column order, plus
    the per-row sequential fp64 prefix sums that equal the non-zero positions of
    ``np.cumsum(myQ, axis=1)`` (fluence_mini_app.py:224) bit for bit (adding an
    exact 0.0 never changes a partial sum)."""
    n = myQ.shape[0]
    indptr = np.zeros(n + 1, dtype=np.int64)
    cols, vals, cum = [], [], []
    for r in range(n):
        nz = np.nonzero(myQ[r])[0]
        indptr[r + 1] = indptr[r] + len(nz)
        cols.append(nz)
        vals.append(myQ[r, nz])
        cum.append(np.cumsum(myQ[r, nz]))
    return (indptr, np.concatenate(cols).astype(np.int64),
            np.concatenate(vals), np.concatenate(cum))


# --------------------------------------------------------------------------
# simple_flux builder (simple_flux_miniapp.py:230-285)
# --------------------------------------------------------------------------

def simple_flux_matrix(dt, SigS, SigA):
    pone = (SigS + SigA) * dt * (np.exp(-(SigA + SigS) * dt))
    pzero = np.exp(-(SigA + SigS) * dt)
    if (1 - pone - pzero) > 0.01:
        raise ValueError("The probability of more than one scattering event occurring "
                         "in a single time step is too large.")
    pscatter = pone * (SigS / (SigA + SigS))
    pabsorb = pone * (SigA / (SigA + SigS))
    myQ = np.zeros((3, 3))
    myQ[0, 0] = pscatter * 0.5 + 1 - pone
    myQ[0, 1] = pscatter * 0.5
    myQ[0, 2] = pabsorb
    myQ[1, 0] = pscatter * 0.5
    myQ[1, 1] = pscatter * 0.5 + 1 - pone
    myQ[1, 2] = pabsorb
    myQ[2, 2] = 1
    myQ = myQ / myQ.sum(axis=1)[:, None]
    myQ[2, 2] = 0
    return myQ


# --------------------------------------------------------------------------
# hot loop: fluence run_pc (fluence_mini_app.py:218-247)
# --------------------------------------------------------------------------

def run_pc_literal(myQ, neuraltime, initial_walkers, meshSize, uniform=None):
    """Line-for-line meaning of fluence_mini_app.py:224-247 (the CSV parse of
    :223 is the caller's business).  ``uniform`` defaults to the global legacy
    ``np.random.uniform`` exactly as the reference uses it (:243)."""
    if uniform is None:
        uniform = np.random.uniform
    probvecs = np.cumsum(myQ, axis=1)
    tempdata = np.zeros((meshSize + 1, neuraltime))
    place = initial_walkers[0][0]
    amount = initial_walkers[0][1]
    tempdata[place, 0] = amount
    for i in range(1, neuraltime):
        for j in range(0, meshSize):
            for k in range(0, int(tempdata[j, i - 1])):
                testrand = uniform()
                tnsp = np.searchsorted(probvecs[j, :], testrand)
                tempdata[tnsp, i] += 1
    return tempdata


def run_pc_counts(indptr, cols, cum, neuraltime, initial_walkers, meshSize,
                  draw=None, dense=True, clamp=True):
    """Count-based restatement of fluence_mini_app.py:236-245 over CSR rows.

    Per step t the alive walkers, ordered (node ascending, slot), consume the
    next ``n_alive`` uniforms of the stream in that order -- the order of the
    reference's loops (:241-243).  ``draw(n)`` must return the next n uniforms;
    the default is ``np.random.uniform(size=n)`` which continues the global
    legacy stream exactly like n scalar calls.

    Returns (tempdata or event list, draws_used).  Events are
    (t, node, count) triples sorted by (t, node) -- the sparse form of tempdata.
    """
    if draw is None:
        draw = lambda n: np.random.uniform(size=n)
    place, amount = int(initial_walkers[0][0]), int(initial_walkers[0][1])
    nodes = np.array([place], dtype=np.int64)
    cnts = np.array([amount], dtype=np.int64)
    events = [(0, place, amount)] if amount else []
    used = 0
    rowlen = np.diff(indptr)
    for t in range(1, neuraltime):
        # sink row is never stepped (:241); a row < meshSize WITHOUT entries makes the reference
        # raise IndexError -- on such graphs (densityrw surface) walkers are absorbed instead
        alive = (nodes < meshSize) & (rowlen[np.minimum(nodes, len(rowlen) - 1)] > 0)
        nodes_a, cnts_a = nodes[alive], cnts[alive]
        n = int(cnts_a.sum())
        if n == 0:
            break
        u = draw(n)
        used += n
        src = np.repeat(nodes_a, cnts_a)              # (node asc, slot) order
        tgt = np.empty(n, dtype=np.int64)
        # searchsorted(left) inside each walker's own row (:244)
        for r in np.unique(src):
            sel = np.nonzero(src == r)[0]
            lo, hi = indptr[r], indptr[r + 1]
            k = np.searchsorted(cum[lo:hi], u[sel], side="left")
            if clamp:
                k = np.minimum(k, hi - lo - 1)        # documented deviation from IndexError
            elif (k >= hi - lo).any():
                raise IndexError("uniform above the row's last cumulative value")
            c = cols[lo + k]
            # Q6: u == 0.0 lands on dense column 0 even if Q[r,0] == 0
            c = np.where(u[sel] <= 0.0, 0, c)
            tgt[sel] = c
        nodes, cnts = np.unique(tgt, return_counts=True)
        cnts = cnts.astype(np.int64)
        events.extend((t, int(a), int(b)) for a, b in zip(nodes, cnts))
    if not dense:
        return events, used
    tempdata = np.zeros((meshSize + 1, neuraltime))
    for (t, k, c) in events:
        tempdata[k, t] = c
    return tempdata, used


# --------------------------------------------------------------------------
# hot loop: simple_flux run_pc (simple_flux_miniapp.py:174-203)
# --------------------------------------------------------------------------

def simple_run_pc_literal(myQ, neuraltime, initial_walkers, uniform=None):
    """simple_flux_miniapp.py:182-203, literal."""
    if uniform is None:
        uniform = np.random.uniform
    tempdata = np.zeros((2, neuraltime))
    place = initial_walkers[0][0]
    amount = initial_walkers[0][1]
    tempdata[place, 0] = amount
    for i in range(1, neuraltime):
        for j in range(0, int(tempdata[0, i - 1])):
            testrand = uniform()
            if testrand <= myQ[0, 0]:
                tempdata[0, i] = tempdata[0, i] + 1
            elif testrand <= (myQ[0, 0] + myQ[0, 1]):
                tempdata[1, i] = tempdata[1, i] + 1
        for j in range(0, int(tempdata[1, i - 1])):
            testrand = uniform()
            if testrand <= myQ[1, 0]:
                tempdata[0, i] = tempdata[0, i] + 1
            elif testrand <= (myQ[1, 0] + myQ[1, 1]):
                tempdata[1, i] = tempdata[1, i] + 1
    return tempdata


def simple_run_pc_counts(myQ, neuraltime, initial_walkers, draw=None):
    """Vectorised simple_flux_miniapp.py:189-201: state-0 walkers first, then
    state-1 walkers (Q9); ``u <= a`` -> 0, ``u <= a+b`` -> 1, else absorbed."""
    if draw is None:
        draw = lambda n: np.random.uniform(size=n)
    tempdata = np.zeros((2, neuraltime))
    place, amount = initial_walkers[0][0], initial_walkers[0][1]
    tempdata[place, 0] = amount
    used = 0
    th = [(myQ[0, 0], myQ[0, 0] + myQ[0, 1]), (myQ[1, 0], myQ[1, 0] + myQ[1, 1])]
    for i in range(1, neuraltime):
        n0, n1 = int(tempdata[0, i - 1]), int(tempdata[1, i - 1])
        if n0 + n1 == 0:
            break
        u = draw(n0 + n1)
        used += n0 + n1
        for s, uu in ((0, u[:n0]), (1, u[n0:])):
            a, b = th[s]
            to0 = uu <= a
            to1 = (~to0) & (uu <= b)
            tempdata[0, i] += to0.sum()
            tempdata[1, i] += to1.sum()
    return tempdata, used


# --------------------------------------------------------------------------
# solution assembly
# --------------------------------------------------------------------------

def source_term_default(x, y):
    """fluence_mini_app.py:30-41 (Q1: the square binds to 7/16 only)."""
    val = (x ** 2) + ((y + (7.0 / 16.0) ** 2))
    return 3.0 if val <= 0.25 else 0.0


def source_term_paper(x, y):
    """Nature-Electronics source quoted in fluence_mini_app.py:488-489 and
    README.md:72: S(x, w) = 0.015 for |x| < 0.5 (BASELINE.json configs[1])."""
    return 0.015 if abs(x) < 0.5 else 0.0


def result_index(i, ND):
    """fluence_mini_app.py:415-419 (Q2) -> (posind, aind)."""
    aind = i % (ND - 1)
    if aind == 0:
        aind = ND - 1
    aind = aind - 1
    posind = int((i - aind) / (ND - 1))
    return posind, aind


def assemble_start_literal(tempdata, meshSize, meshpoints, SigA, dt, v, M, source=source_term_default):
    """fluence_mini_app.py:422-438 literal: sequential fp64 sum in (t, k) order,
    t = 1 .. nt-2 (Q3), rows < meshSize (Q4), zero counts skipped (:433)."""
    tempval = 0
    for j in range(1, tempdata.shape[1] - 1):
        for k in range(0, meshSize):
            if (tempdata[k, j] != 0):
                tempval = tempval + tempdata[k, j] * np.exp(-SigA * dt * j) * v * source(
                    meshpoints[k, 0], meshpoints[k, 1]) * dt
    return tempval / M


def assemble_start_events(events, nt, meshSize, srcvals, SigA, dt, v, M):
    """Same sum as ``assemble_start_literal`` driven by the sparse (t, k, count)
    list (sorted by (t, k)), so it costs O(events) instead of O(nt*meshSize).
    ``srcvals[k]`` = source(meshpoints[k]).  Operation order per term is the
    reference's: (((count*exp)*v)*S)*dt, then added to the running sum."""
    tempval = 0
    for (t, k, c) in events:
        if 1 <= t <= nt - 2 and k < meshSize and c != 0:
            tempval = tempval + float(c) * np.exp(-SigA * dt * t) * v * srcvals[k] * dt
    return tempval / M


def fluence_run(args, source=source_term_default, literal=False, matrix=None):
    """fluence_mini_app.py:251-447 restated end to end for run_mode 'pc'
    (without file I/O).  Uses the global legacy numpy stream like the reference
    (seed it first as fluence_mini_app.py:523 does).  Returns (rwSol, total_draws)."""
    M, dt, da, v = args["num_walks"], args["timestep"], args["anglestep"], args["velocity"]
    SigS, SigA, L = args["SigS"], args["SigA"], args["interval_length"]
    nt = args["neural_timesteps"]
    if matrix is None:
        myQ, m = fluence_matrix_dense(dt, da, v, SigS, L)
    else:
        myQ, m = matrix
    meshSize, ND, NS, meshpoints = m["meshSize"], m["ND"], m["NS"], m["meshpoints"]
    rwSol = np.zeros((NS - 1, ND - 1))
    total = 0
    if literal:
        for i in range(meshSize):
            td = run_pc_literal(myQ, nt, [[i, M]], meshSize)
            total += int(td[:meshSize, :nt - 1].sum())
            p, a = result_index(i, ND)
            rwSol[p, a] = assemble_start_literal(td, meshSize, meshpoints, SigA, dt, v, M, source)
        return rwSol, total
    indptr, cols, vals, cum = dense_to_csr(myQ)
    srcvals = np.array([source(meshpoints[k, 0], meshpoints[k, 1]) for k in range(meshSize)])
    for i in range(meshSize):
        ev, used = run_pc_counts(indptr, cols, cum, nt, [[i, M]], meshSize, dense=False)
        total += used
        p, a = result_index(i, ND)
        rwSol[p, a] = assemble_start_events(ev, nt, meshSize, srcvals, SigA, dt, v, M)
    return rwSol, total


def simple_flux_run(args, literal=False):
    """simple_flux_miniapp.py:207-359 restated for run_mode 'pc' (no file I/O)."""
    M, dt = args["num_walks"], args["timestep"]
    initp, initn = args["init_pos"], args["init_neg"]
    SigS, SigA = args["SigS"], args["SigA"]
    max_time, nt = args["max_time"], args["neural_timesteps"]
    myQ = simple_flux_matrix(dt, SigS, SigA)
    maxT = min(nt, int(max_time / dt)) + 2            # :296-297
    rwSol = np.zeros((2, maxT))
    total = 0
    for i in range(2):
        if literal:
            td = simple_run_pc_literal(myQ, nt, [[i, M]])
            used = int(td[:, :nt - 1].sum())
        else:
            td, used = simple_run_pc_counts(myQ, nt, [[i, M]])
        total += used
        maxT = min((td.shape[1] - 2), maxT)           # :324
        for j in range(0, maxT):
            rwSol[i, j] = (initp * td[0, j] + initn * td[1, j]) / M   # :327
    return rwSol, total, myQ


def simple_flux_true_solution(args, maxT):
    """simple_flux_miniapp.py:338-343 -- the closed form (only known answer in
    the reference)."""
    dt, ip, in_ = args["timestep"], args["init_pos"], args["init_neg"]
    SigS, SigA = args["SigS"], args["SigA"]
    t = np.arange(maxT) * dt
    e1, e2 = np.exp(-SigA * t), np.exp(-(SigA + SigS) * t)
    return np.stack([(ip / 2) * (e1 + e2) + (in_ / 2) * (e1 - e2),
                     (ip / 2) * (e1 - e2) + (in_ / 2) * (e1 + e2)])
