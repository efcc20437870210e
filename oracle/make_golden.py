"""Generate tests/golden/*.npz from the UNMODIFIED reference (build container only).

TEST INFRASTRUCTURE.  Run as ``python -m oracle.make_golden [--which ...]`` from the
repo root in a container that has /root/reference.  The reference ships no
golden vectors, so these fixtures -- outputs of the reference's
own ``run_pc`` / ``run`` under fixed seeds -- are what pins the oracle and, through
it, the CUDA path.  The GPU box has no /root/reference; it reads only the
committed ``.npz`` files.

Fixtures (all produced by calling the reference's code, never the restatement):
  fluence_c1_matrix.npz   CSR of the matrix CSV written by reference ``run`` (C1 params)
  fluence_c1_tempdata.npz sparse ``tempdata`` of reference ``run_pc`` for 12 start
                          nodes x 2 seeds at nt=64, M=100, the stream position after
                          each call (the next uniform), and one nt=200 case
  fluence_c1_rwsol.npz    full ``rwSol`` of reference ``run`` for C1, nt=64, seed 1
  fluence_c2_rwsol.npz    same at M=1000, nt=64, seed 1, with the module-level
                          ``sourceTerm`` swapped for the Nature-Electronics source
                          (the reference hard-codes its source, fluence_mini_app.py:30-41;
                          the swap rebinds that one global and changes no other line)
  simple_flux.npz         reference simple_flux ``run_pc`` tempdata (non-zero prefix)
                          and ``run`` rwSol at the script defaults, seeds 1 and 2, and
                          a ``-sa 2 -ss 8 -dt 0.005 -mt 5 -nt 1002`` case (C3 parity)
"""
import argparse
import os
import sys
import tempfile
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import ref_import  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
C1 = dict(timestep=0.01, anglestep=1.0 / 15.0, velocity=200, SigS=0.15, SigA=0,
          interval_length=1)
STARTS = [0, 1, 14, 15, 16, 29, 30, 435, 449, 450, 464, 899]


def _tmpdir():
    return tempfile.mkdtemp(prefix="nrw_gold_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)


def _sparse(td):
    k, t = np.nonzero(td)
    order = np.lexsort((k, t))
    return t[order].astype(np.int32), k[order].astype(np.int32), td[k[order], t[order]].astype(np.int64)


def _dense_to_csr(q):
    n = q.shape[0]
    indptr = np.zeros(n + 1, dtype=np.int64)
    cols, vals = [], []
    for r in range(n):
        nz = np.nonzero(q[r])[0]
        indptr[r + 1] = indptr[r] + len(nz)
        cols.append(nz)
        vals.append(q[r, nz])
    return indptr, np.concatenate(cols).astype(np.int32), np.concatenate(vals)


def _write_matrix_via_reference(ref, d, **over):
    """Let the reference's run() build and save its matrix, then stop it at the
    first back-end call (the matrix is written before the walk starts, :384)."""
    class _Stop(Exception):
        pass

    def _stopper(**kw):
        raise _Stop()
    mat = os.path.join(d, "curmat.csv")
    args = ref_import.fluence_default_args(matrix_filename=mat,
                                           results_filename=os.path.join(d, "rwSol.csv"),
                                           neural_timesteps=4, **C1)
    args.update(over)
    saved = ref.run_functions["pc"]
    ref.run_functions["pc"] = _stopper
    try:
        ref.run(args)
    except _Stop:
        pass
    finally:
        ref.run_functions["pc"] = saved
    return mat


def gold_matrix():
    ref = ref_import.fluence()
    d = _tmpdir()
    mat = _write_matrix_via_reference(ref, d)
    q = np.genfromtxt(mat, delimiter=",")
    indptr, cols, vals = _dense_to_csr(q)
    np.savez_compressed(os.path.join(GOLD, "fluence_c1_matrix.npz"),
                        indptr=indptr, cols=cols, vals=vals, shape=np.array(q.shape),
                        rowsum_bits=q.sum(axis=1).view(np.uint64))
    print("matrix", q.shape, "nnz", len(cols))


def gold_tempdata():
    ref = ref_import.fluence()
    d = _tmpdir()
    mat = _write_matrix_via_reference(ref, d)
    out = {}
    cases = [(s, seed, 64, 100) for seed in (1, 123) for s in STARTS] + [(435, 7, 200, 250)]
    meta = []
    for ci, (start, seed, nt, M) in enumerate(cases):
        np.random.seed(seed)
        td = ref.run_pc(matrix_filename=mat, neuraltime=nt, initial_walkers=[[start, M]],
                        meshSize=900, runtime_args={})
        nxt = np.random.uniform()
        t, k, c = _sparse(td)
        out["t_%d" % ci], out["k_%d" % ci], out["c_%d" % ci] = t, k, c
        meta.append((start, seed, nt, M, int(td[:900, :nt - 1].sum())))
        out["next_u_%d" % ci] = np.array([nxt])
        print("tempdata case", ci, meta[-1])
    # one chained pair: two run_pc calls on one seeded stream (Q5: stream continues)
    np.random.seed(99)
    td_a = ref.run_pc(matrix_filename=mat, neuraltime=64, initial_walkers=[[14, 100]],
                      meshSize=900, runtime_args={})
    td_b = ref.run_pc(matrix_filename=mat, neuraltime=64, initial_walkers=[[435, 100]],
                      meshSize=900, runtime_args={})
    for nm, td in (("chain_a", td_a), ("chain_b", td_b)):
        t, k, c = _sparse(td)
        out[nm + "_t"], out[nm + "_k"], out[nm + "_c"] = t, k, c
    out["meta"] = np.array(meta, dtype=np.int64)
    np.savez_compressed(os.path.join(GOLD, "fluence_c1_tempdata.npz"), **out)


def gold_rwsol(tag, M, source_swap):
    ref = ref_import.fluence()
    d = _tmpdir()
    args = ref_import.fluence_default_args(
        matrix_filename=os.path.join(d, "curmat.csv"),
        results_filename=os.path.join(d, "rwSol.csv"),
        neural_timesteps=64, num_walks=M, rand_seed=1, **C1)
    saved = ref.sourceTerm
    if source_swap:
        ref.sourceTerm = lambda x, y: 0.015 if abs(x) < 0.5 else 0.0
    t0 = time.time()
    try:
        np.random.seed(1)              # fluence_mini_app.py:523
        ref.run(args)
        nxt = np.random.uniform()
    finally:
        ref.sourceTerm = saved
    sol = np.genfromtxt(args["results_filename"], delimiter=",")
    np.savez_compressed(os.path.join(GOLD, "fluence_%s_rwsol.npz" % tag), rwSol=sol,
                        next_u=np.array([nxt]), M=np.array([M]), nt=np.array([64]),
                        seed=np.array([1]), seconds=np.array([time.time() - t0]))
    print(tag, "rwSol done in %.1f s" % (time.time() - t0), sol.shape, sol.sum())


def gold_simple():
    ref = ref_import.simple_flux()
    d = _tmpdir()
    out = {}
    for seed in (1, 2):
        args = ref_import.simple_flux_default_args(
            matrix_filename=os.path.join(d, "m.csv"), results_filename=os.path.join(d, "r.csv"),
            rand_seed=seed)
        np.random.seed(seed)
        ref.run(args)
        nxt = np.random.uniform()
        out["rwSol_seed%d" % seed] = np.genfromtxt(args["results_filename"], delimiter=",")
        out["next_u_seed%d" % seed] = np.array([nxt])
        out["myQ"] = np.genfromtxt(args["matrix_filename"], delimiter=",")
        # raw run_pc tempdata, both start states, on a fresh stream
        np.random.seed(seed)
        for i in (0, 1):
            td = ref.run_pc(matrix_filename=args["matrix_filename"], neuraltime=500000,
                            initial_walkers=[[i, 500]], meshSize=2, runtime_args={})
            last = int(np.nonzero(td.sum(axis=0))[0].max()) + 2
            out["td_seed%d_start%d" % (seed, i)] = td[:, :last].astype(np.int64)
        print("simple seed", seed, "done")
    # C3 parity shape: nt=1002
    args = ref_import.simple_flux_default_args(
        matrix_filename=os.path.join(d, "m.csv"), results_filename=os.path.join(d, "r3.csv"),
        rand_seed=5, neural_timesteps=1002, use_sinks=True)
    np.random.seed(5)
    ref.run(args)
    out["rwSol_c3_seed5"] = np.genfromtxt(args["results_filename"], delimiter=",")
    out["next_u_c3_seed5"] = np.array([np.random.uniform()])
    np.savez_compressed(os.path.join(GOLD, "simple_flux.npz"), **out)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--which", default="matrix,tempdata,simple",
                    help="comma list of matrix,tempdata,simple,rwsol_c1,rwsol_c2")
    a = ap.parse_args()
    os.makedirs(GOLD, exist_ok=True)
    for w in a.which.split(","):
        {"matrix": gold_matrix, "tempdata": gold_tempdata, "simple": gold_simple,
         "rwsol_c1": lambda: gold_rwsol("c1", 100, False),
         "rwsol_c2": lambda: gold_rwsol("c2", 1000, True)}[w]()
