"""Transition-matrix builders: (space, direction) fluence mesh and the 2-state flux chain,
built DIRECTLY as CSR (never the dense (meshSize+1)^2 matrix of the reference).

Replaces fluence_mini_app.py:279-386 and simple_flux_miniapp.py:230-292 of the reference.
The CSR values carry the reference's bits: the reference normalises each row by
``myQ.sum(axis=1)`` (fluence_mini_app.py:378), a NumPy *pairwise* sum over the dense row whose
last bits depend on where the non-zeros sit; ``_pairwise_rowsum_window`` evaluates
exactly that summation tree on the sparse row, so ``Q`` -- and through it every cumulative table
and every replayed trajectory -- is bit-identical without ever allocating the dense matrix.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field

import numpy as np

_PW_BLOCK = 128   # numpy's PW_BLOCKSIZE (loops_utils.h.src): leaves of the pairwise tree


@dataclass
class TransitionCSR:
    """Row-stochastic transition matrix in CSR, columns ascending within a row.

    Rows ``>= n_live`` are absorbing and never stepped (the sink row of
    fluence_mini_app.py:241,381; the absorbed state of simple_flux_miniapp.py:285).
    """
    n_rows: int
    n_live: int
    indptr: np.ndarray          # int64 [n_rows + 1]
    cols: np.ndarray            # int32 [nnz]
    vals: np.ndarray            # float64 [nnz]
    meta: dict = field(default_factory=dict)

    @property
    def nnz(self) -> int:
        return int(self.indptr[-1])

    def row_lengths(self) -> np.ndarray:
        return np.diff(self.indptr)

    def to_dense(self) -> np.ndarray:
        q = np.zeros((self.n_rows, self.n_rows))
        rows = np.repeat(np.arange(self.n_rows), self.row_lengths())
        q[rows, self.cols] = self.vals
        return q


# ---------------------------------------------------------------------------------------
# numpy pairwise summation on a sparse row
# ---------------------------------------------------------------------------------------

def _pairwise_rowsum_window(n: int, c0: int, block: np.ndarray) -> np.ndarray:
    """``dense.sum(axis=1)`` for rows of length ``n`` that are zero outside columns
    ``[c0, c0 + W)`` and equal ``block[r, :]`` inside, evaluated with NumPy's pairwise tree
    (n < 8: sequential; n <= 128: 8 strided accumulators; else split at ``n/2`` rounded down to
    a multiple of 8).  Leaves are summed by NumPy itself on small dense tiles, sub-trees with no
    non-zero contribute an exact 0.0 and are skipped (x + 0.0 == x)."""
    rows, W = block.shape
    hi = c0 + W

    def rec(start: int, size: int):
        if start + size <= c0 or start >= hi:
            return None
        if size <= _PW_BLOCK:
            tile = np.zeros((rows, size))
            a, b = max(start, c0), min(start + size, hi)
            tile[:, a - start:b - start] = block[:, a - c0:b - c0]
            return np.add.reduce(tile, axis=1)
        n2 = size // 2
        n2 -= n2 % 8
        left, right = rec(start, n2), rec(start + n2, size - n2)
        if left is None:
            return right
        if right is None:
            return left
        return left + right

    out = rec(0, n)
    return np.zeros(rows) if out is None else out


# ---------------------------------------------------------------------------------------
# fluence mesh (fluence_mini_app.py:279-324) and matrix (:331-381)
# ---------------------------------------------------------------------------------------

def check_fluence_physics(dt, v, SigS):
    """fluence_mini_app.py:283-291: more than 5 % chance of a second scattering -> ValueError."""
    pone = v * SigS * dt * (np.exp(-v * SigS * dt))
    pzero = np.exp(-v * SigS * dt)
    if (1 - pone - pzero) > 0.05:
        raise ValueError("The probability of more than one scattering event occurring in a "
                         "single time step is too large.")
    return pone


def fluence_mesh(dt, da, v, L):
    """fluence_mini_app.py:279, :297-324.  row = i + j*(ND-1) <-> (cps[j], directions[i])."""
    ds = da * v * dt / 2
    NS = int((2 * L) / ds) + 1
    space = np.linspace(-L, L, NS)
    cps = (space + (ds / 2))[0:(NS - 1)]
    ND = int(2 / da) + 1
    dirs = np.linspace(-1, 1, ND)
    directions = (dirs + (da / 2))[0:(ND - 1)]
    meshSize = (NS - 1) * (ND - 1)
    meshpoints = np.empty((meshSize, 2))
    meshpoints[:, 0] = np.repeat(cps, ND - 1)
    meshpoints[:, 1] = np.tile(directions, NS - 1)
    return dict(ds=ds, NS=NS, ND=ND, space=space, dirs=dirs, cps=cps, directions=directions,
                meshSize=meshSize, meshpoints=meshpoints)


def _locate_cells(cps, newpos):
    """Vectorised ``np.where(np.round(cps - newpos, 8) == 0)[0][0]`` (fluence_mini_app.py:361-362):
    the first cell whose centre equals ``newpos`` to 8 decimals; IndexError if none (Q15)."""
    ds = cps[1] - cps[0] if len(cps) > 1 else 1.0
    guess = np.rint((newpos - cps[0]) / ds).astype(np.int64)
    out = np.full(newpos.shape, -1, dtype=np.int64)
    for d in (-1, 0, 1):                      # ascending: keeps "first match" semantics
        j = guess + d
        ok = (j >= 0) & (j < len(cps)) & (out < 0)
        jj = np.clip(j, 0, len(cps) - 1)
        hit = ok & (np.round(cps[jj] - newpos, 8) == 0)
        out[hit] = jj[hit]
    if (out < 0).any():
        raise IndexError("index 0 is out of bounds for axis 0 with size 0 "
                         "(a walker step does not land on a cell centre; see SURVEY Q15)")
    return out


def build_fluence_csr(dt, da, v, SigS, L, mesh=None) -> TransitionCSR:
    """The reference's transition matrix (fluence_mini_app.py:331-381) as CSR, bit-identical.

    Exit rows (|x - v*mu*dt| > L) hold a single 1.0 in the sink column; stay rows hold
    ``freshprob`` on every direction of the new cell and ``returnprob`` on the unchanged direction,
    divided by the row's NumPy-pairwise sum; the sink row is empty (:381)."""
    pone = check_fluence_physics(dt, v, SigS)
    m = mesh if mesh is not None else fluence_mesh(dt, da, v, L)
    ND, NS, cps, meshSize, mp = m["ND"], m["NS"], m["cps"], m["meshSize"], m["meshpoints"]
    nd1 = ND - 1
    n = meshSize + 1
    freshprob = pone / (ND - 1)
    returnprob = freshprob + (1 - pone)

    newpos = mp[:, 0] - v * mp[:, 1] * dt                     # :353
    exits = np.abs(newpos) > L                                # :355
    stay = np.nonzero(~exits)[0]
    newindex = np.zeros(meshSize, dtype=np.int64)
    if len(stay):
        newindex[stay] = _locate_cells(cps, newpos[stay])
    samedir = np.arange(meshSize) % nd1                       # :370-371 (directions[i] == mp[i,1])

    lengths = np.where(exits, 1, nd1).astype(np.int64)
    indptr = np.zeros(n + 1, dtype=np.int64)
    indptr[1:meshSize + 1] = np.cumsum(lengths)
    indptr[n] = indptr[meshSize]                              # sink row: no entries
    nnz = int(indptr[-1])
    cols = np.empty(nnz, dtype=np.int32)
    vals = np.empty(nnz)

    ex = np.nonzero(exits)[0]
    cols[indptr[ex]] = meshSize
    vals[indptr[ex]] = 1.0                                    # 1/1 after normalisation

    if len(stay):
        # group stay rows by target cell: they share the window [c0, c0 + nd1)
        order = stay[np.argsort(newindex[stay], kind="stable")]
        bounds = np.nonzero(np.diff(newindex[order], prepend=-1))[0]
        for gi, b in enumerate(bounds):
            e = bounds[gi + 1] if gi + 1 < len(bounds) else len(order)
            rows = order[b:e]
            c0 = int(newindex[rows[0]]) * nd1
            block = np.full((len(rows), nd1), freshprob)
            block[np.arange(len(rows)), samedir[rows]] = returnprob
            rowsum = _pairwise_rowsum_window(n, c0, block)
            q = block / rowsum[:, None]                       # :378
            at = indptr[rows][:, None] + np.arange(nd1)[None, :]
            cols[at] = (c0 + np.arange(nd1))[None, :]
            vals[at] = q
    return TransitionCSR(n_rows=n, n_live=meshSize, indptr=indptr, cols=cols, vals=vals,
                         meta=dict(kind="fluence", mesh=m, pone=pone))


# ---------------------------------------------------------------------------------------
# simple_flux (simple_flux_miniapp.py:230-285)
# ---------------------------------------------------------------------------------------

def build_simple_flux_dense(dt, SigS, SigA) -> np.ndarray:
    """3x3 chain (+dir, -dir, absorbed) -- simple_flux_miniapp.py:230-285."""
    pone = (SigS + SigA) * dt * (np.exp(-(SigA + SigS) * dt))
    pzero = np.exp(-(SigA + SigS) * dt)
    if (1 - pone - pzero) > 0.01:
        raise ValueError("The probability of more than one scattering event occurring in a "
                         "single time step is too large.")
    pscatter = pone * (SigS / (SigA + SigS))
    pabsorb = pone * (SigA / (SigA + SigS))
    q = np.zeros((3, 3))
    q[0, 0] = pscatter * 0.5 + 1 - pone
    q[0, 1] = pscatter * 0.5
    q[0, 2] = pabsorb
    q[1, 0] = pscatter * 0.5
    q[1, 1] = pscatter * 0.5 + 1 - pone
    q[1, 2] = pabsorb
    q[2, 2] = 1
    q = q / q.sum(axis=1)[:, None]
    q[2, 2] = 0
    return q


def csr_from_dense(q: np.ndarray, n_live: int | None = None, meta=None) -> TransitionCSR:
    """Non-zeros of a dense matrix in column order (what ``np.cumsum(myQ, axis=1)`` sees)."""
    n = q.shape[0]
    rows, cols = np.nonzero(q)                                # row-major: cols ascending per row
    indptr = np.zeros(n + 1, dtype=np.int64)
    np.add.at(indptr, rows + 1, 1)
    indptr = np.cumsum(indptr)
    if n_live is None:
        n_live = n - 1
    return TransitionCSR(n_rows=n, n_live=n_live, indptr=indptr, cols=cols.astype(np.int32),
                         vals=q[rows, cols].astype(np.float64), meta=meta or {})


# ---------------------------------------------------------------------------------------
# matrix files: the reference hands the matrix to its back ends BY FILENAME
# (np.savetxt fluence_mini_app.py:384 -> np.genfromtxt :223, '%.18e' round-trips fp64 exactly)
# ---------------------------------------------------------------------------------------

MAX_DENSE_CSV_ROWS = 6000       # 6000^2 * 25 B = 0.9 GB of text; beyond that only .npz is written


def save_matrix_csv(path: str, csr: TransitionCSR) -> None:
    """Byte-identical to ``np.savetxt(path, dense, delimiter=',')`` without the dense matrix."""
    zero = "%.18e" % 0.0
    n = csr.n_rows
    with open(path, "w") as f:
        for r in range(n):
            fields = [zero] * n
            lo, hi = csr.indptr[r], csr.indptr[r + 1]
            for c, x in zip(csr.cols[lo:hi], csr.vals[lo:hi]):
                fields[c] = "%.18e" % x
            f.write(",".join(fields))
            f.write("\n")


def save_array_csv(path: str, a) -> None:
    """Byte-identical to ``np.savetxt(path, a, delimiter=',')`` for a real 1-D or 2-D array (the result files
    fluence_mini_app.py:446, simple_flux_miniapp.py:358): '%.18e' per value, one row per line.  ONE C-level format
    call for the whole file instead of savetxt's Python loop over rows -- a third of its time, and on the 10 000-cell
    mesh the result file was half of a warm ``run(args)``."""
    a = np.asarray(a, dtype=np.float64)
    if a.ndim == 0 or a.ndim > 2 or a.size == 0:
        np.savetxt(path, a, delimiter=",")
        return
    ncols = 1 if a.ndim == 1 else a.shape[1]
    nrows = a.size // ncols
    text = _format_csv_native(np.ascontiguousarray(a), nrows, ncols) if a.size >= 2048 else None
    if text is None:                         # small arrays, or libnrw.so not built: one C-level format call in Python
        row = ",".join(["%.18e"] * ncols) + "\n"
        text = ((row * nrows) % tuple(a.ravel().tolist())).encode("ascii")
    with open(path, "wb") as f:
        f.write(text)


def _format_csv_native(a: np.ndarray, nrows: int, ncols: int):
    """The same text from libnrw.so (nrw_format_csv_e18: the rows converted by a few host threads), or None when the
    library is not there.  A correctly rounded 19-digit conversion costs ~0.3 us per value in any libc and in CPython;
    on the 10 000-cell mesh that was 2.6 ms of a 5.9 ms run(args)."""
    import ctypes as C
    try:
        from .engine import load_library
        lib = load_library()
    except Exception:
        return None
    cap = nrows * ncols * 28
    out = C.create_string_buffer(cap)
    n = C.c_int64(0)
    threads = max(1, min(8, (os.cpu_count() or 1)))
    if lib.nrw_format_csv_e18(C.c_void_p(a.ctypes.data), nrows, ncols, out, cap, C.byref(n), threads) != 0:
        return None
    return out.raw[:n.value]


def save_matrix_npz(path: str, csr: TransitionCSR) -> None:
    """CSR as an uncompressed .npz (zlib on the 6 MB of the 10 001-row mesh took 0.17 s -- forty times the solve)."""
    np.savez(path, n_rows=csr.n_rows, n_live=csr.n_live, indptr=csr.indptr, cols=csr.cols, vals=csr.vals)


_csr_cache: dict = {}


def build_fluence_csr_cached(dt, da, v, SigS, L) -> TransitionCSR:
    """``build_fluence_csr`` memoised by the physics parameters: a driver that calls ``run(args)`` repeatedly
    (parameter studies, bench.py) gets the same TransitionCSR object back, so the device tables and launch
    descriptors cached on it (nrw/dropin.py, nrw/fluence.py) are reused as well."""
    key = (float(dt), float(da), float(v), float(SigS), float(L))
    hit = _csr_cache.get(key)
    if hit is None:
        if len(_csr_cache) > 4:
            _csr_cache.clear()
        hit = _csr_cache[key] = build_fluence_csr(dt, da, v, SigS, L)
    return hit


class MatrixFileWriter:
    """The transition-matrix file of run(args) (fluence_mini_app.py:384) written OFF the critical path: a
    background thread writes it while the GPU walks; ``join()`` before run() returns, so the file is complete
    when the caller sees the result.  The pc path itself never reads it (it walks the in-memory CSR); it is
    kept for the other back ends and for users.  A file this process already wrote for the same matrix object
    and that is still unchanged on disk is not rewritten."""

    _written: dict = {}

    def __init__(self, path: str, csr: TransitionCSR):
        import threading
        self.path, self.csr, self.error, self.skipped = path, csr, None, False
        tag = self._written.get(os.path.abspath(path))
        try:
            st = os.stat(path)
            if tag is not None and tag == (id(csr), st.st_mtime_ns, st.st_size):
                self.skipped = True
        except OSError:
            pass
        self._thr = None
        if not self.skipped:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def _run(self):
        try:
            if self.path.endswith(".npz"):
                save_matrix_npz(self.path, self.csr)
            else:
                save_matrix_csv(self.path, self.csr)
            st = os.stat(self.path)
            self._written[os.path.abspath(self.path)] = (id(self.csr), st.st_mtime_ns, st.st_size)
        except Exception as exc:                                   # re-raised in join()
            self.error = exc

    def join(self):
        if self._thr is not None:
            self._thr.join()
        if self.error is not None:
            raise self.error


_matrix_cache: dict = {}


def load_matrix(path: str, n_live: int | None = None) -> TransitionCSR:
    """CSV (dense, as the reference writes it) or .npz (CSR) -> TransitionCSR, cached by
    (path, mtime, size) so that a driver calling ``run_pc`` once per start node does not re-parse
    the file every time (the reference does: 0.53 s x 900 calls)."""
    st = os.stat(path)
    key = (os.path.abspath(path), st.st_mtime_ns, st.st_size, n_live)
    hit = _matrix_cache.get(key)
    if hit is not None:
        return hit
    if path.endswith(".npz"):
        z = np.load(path)
        csr = TransitionCSR(int(z["n_rows"]), int(z["n_live"]) if n_live is None else n_live,
                            z["indptr"].astype(np.int64), z["cols"].astype(np.int32),
                            z["vals"].astype(np.float64))
    else:
        q = np.loadtxt(path, delimiter=",", ndmin=2)
        csr = csr_from_dense(q, n_live=n_live)
    csr.meta["path"] = path
    if len(_matrix_cache) > 8:
        _matrix_cache.clear()
    _matrix_cache[key] = csr
    return csr


# ---------------------------------------------------------------------------------------
# source terms (fluence_mini_app.py:30-41 and the paper source quoted at :488-489)
# ---------------------------------------------------------------------------------------

def source_default(x, y):
    """fluence_mini_app.py:36-41 -- note the square binds to 7/16 only."""
    val = (x ** 2) + ((y + (7.0 / 16.0) ** 2))
    return 3.0 if val <= 0.25 else 0.0


def source_paper(x, y):
    """S(x, w) = 0.015 for |x| < 0.5: the Nature-Electronics source (fluence_mini_app.py:488-489)."""
    return 0.015 if abs(x) < 0.5 else 0.0


SOURCES = {"default": source_default, "paper": source_paper}


def source_values(meshpoints: np.ndarray, source) -> np.ndarray:
    return np.array([source(float(meshpoints[k, 0]), float(meshpoints[k, 1]))
                     for k in range(meshpoints.shape[0])])


def result_index(i: int, ND: int):
    """fluence_mini_app.py:415-419: start node -> (posind, aind) of rwSol."""
    aind = i % (ND - 1)
    if aind == 0:
        aind = ND - 1
    aind = aind - 1
    posind = int((i - aind) / (ND - 1))
    return posind, aind
