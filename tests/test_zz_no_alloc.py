"""nrw_fluence_scores_host_static (include/nrw.h): the host-buffer entry that allocates nothing -- all device-side
buffers in one `static __device__` array of NRW_STATIC_SCRATCH_BYTES inside libnrw.so.  It has no counterpart in the
reference, so it is held to the existing entry point,
nrw_fluence_scores_host, on the same inputs: integer tallies, so the tolerance is zero.  Called through the C ABI
(ctypes), twice in a row and more -- the array keeps the previous call's contents and the call must clear what has
to start at zero.  Collected last on purpose; passes alone in a fresh process and
after the other tests."""
import ctypes as C

import numpy as np
import pytest

from conftest import C1

pytestmark = pytest.mark.gpu

STATIC_BYTES = 1 << 20


def _problem(par, starts, M):
    from nrw import mesh, tables
    csr = mesh.build_fluence_csr(par["dt"], par["da"], par["v"], par["SigS"], par["L"])
    src = mesh.source_values(csr.meta["mesh"]["meshpoints"], mesh.source_default)
    nc, _ = tables.classify_source(src, csr.n_rows)
    at = tables.build_alias(csr, node_class=nc, layout="compact")
    starts = np.ascontiguousarray(starts(csr.n_live), dtype=np.int32)
    counts = np.broadcast_to(np.asarray(M, dtype=np.int64), starts.shape)
    prefix = np.zeros(len(starts) + 1, dtype=np.int64)
    prefix[1:] = np.cumsum(counts)
    return csr, at, starts, prefix


def _call(lib, fn_name, handle, csr, at, starts, prefix, id_base, seed, nt, cv, lo=None, hi=None, use_str=True):
    score = np.full(len(starts), -7, dtype=np.int64)            # poisoned: the call must overwrite every word
    counters = np.full(8, 2 ** 62, dtype=np.uint64)
    cvv = (C.c_double * 4)(*cv)
    sr = at.structured() if use_str else None                  # the structured form of the tables, or without it
    head = (handle, csr.n_rows, csr.n_live, at.n_alias, C.c_void_p(at.rowinfo.ctypes.data),
            C.c_void_p(at.alias.ctypes.data), C.c_void_p(sr[0].ctypes.data) if sr else None, sr[1] if sr else 0,
            sr[2] if sr else 0,
            len(starts), C.c_void_p(starts.ctypes.data),
            C.c_void_p(prefix.ctypes.data), C.c_void_p(id_base.ctypes.data) if id_base is not None else None,
            int(prefix[0] if lo is None else lo), int(prefix[-1] if hi is None else hi), seed, nt, C.byref(cvv),
            C.c_void_p(score.ctypes.data), C.c_void_p(counters.ctypes.data))
    if fn_name == "nrw_fluence_scores_host":
        rc = lib.nrw_fluence_scores_host(*head, None, None)     # reduce = NULL (one GPU), stream = NULL
    else:
        rc = lib.nrw_fluence_scores_host_static(*head, None)
    assert rc == 0, lib.nrw_last_error().decode()
    return score, counters


def test_static_scratch_entry_equals_the_allocating_entry_on_calls_in_a_row():
    from nrw.engine import get_engine
    eng = get_engine()
    lib, h = eng.lib, eng.handle
    c5 = dict(dt=0.01, da=0.02, v=200, SigS=0.15, L=1)
    cases = [
        # default mesh, every start, identity weights (the IDW kernel) ...
        (_problem(C1, lambda n: np.arange(n), 700), None, [0, 1, 0, 0], 64),
        # ... the bench mesh (10 001 rows: 0.42 MiB of the 1 MiB array), a strided start list, sharded stream ids
        (_problem(c5, lambda n: np.arange(0, n, 3), 900), "ids", [0, 1, 0, 0], 30000),
        # ... packed (non-identity) weights, ragged counts incl. an empty start, short horizon (survivors)
        (_problem(C1, lambda n: np.array([435, 0, 14, 14, 899, 450]), [3000, 0, 1, 333, 64, 500]), None,
         [0, 5, 0, 0], 4),
    ]
    for (csr, at, starts, prefix), ids, cv, nt in cases:
        id_base = (starts.astype(np.int64) * 100000 + 17) if ids else None
        # two calls in a row and a third with the first seed again: nothing of a call may leak into the next
        for seed in (11, 12, 11):
            want = _call(lib, "nrw_fluence_scores_host", h, csr, at, starts, prefix, id_base, seed, nt, cv)
            got = _call(lib, "nrw_fluence_scores_host_static", h, csr, at, starts, prefix, id_base, seed, nt, cv)
            assert np.array_equal(got[0], want[0]) and np.array_equal(got[1][:5], want[1][:5]), (seed, nt)
            assert got[1][4] == prefix[-1] and (got[0] >= 0).all()
            # ... and on the alias tables alone (no structured form passed): the same integers again
            plain = _call(lib, "nrw_fluence_scores_host_static", h, csr, at, starts, prefix, id_base, seed, nt, cv,
                          use_str=False)
            assert np.array_equal(plain[0], want[0]) and np.array_equal(plain[1][:5], want[1][:5])
        # a sub-range of the ids (what a rank walks) and an empty one
        lo, hi = int(prefix[-1]) // 3, int(prefix[-1]) // 2
        for a, b in ((lo, hi), (lo, lo)):
            want = _call(lib, "nrw_fluence_scores_host", h, csr, at, starts, prefix, id_base, 5, nt, cv, a, b)
            got = _call(lib, "nrw_fluence_scores_host_static", h, csr, at, starts, prefix, id_base, 5, nt, cv, a, b)
            assert np.array_equal(got[0], want[0]) and np.array_equal(got[1][:5], want[1][:5])
            assert got[1][4] == b - a


def test_static_scratch_entry_refuses_what_does_not_fit():
    """Sizes beyond NRW_STATIC_SCRATCH_BYTES: NRW_ENOSPC with both numbers in the message, nothing touched (the
    size test comes before any buffer is read: the tiny arrays passed here are never dereferenced)."""
    from nrw.engine import get_engine
    eng = get_engine()
    lib, h = eng.lib, eng.handle
    tiny = np.zeros(16, dtype=np.int64)
    cv = (C.c_double * 4)(0, 1, 0, 0)
    p = C.c_void_p(tiny.ctypes.data)
    # the refined mesh (158 761 rows: 1.27 MB of descriptors alone) does not fit
    rc = lib.nrw_fluence_scores_host_static(h, 158761, 158760, 8001, p, p, None, 0, 0, 100, p, p, None, 0, 100, 1, 64,
                                            C.byref(cv), p, p, None)
    msg = lib.nrw_last_error().decode()
    assert rc == -3 and str(STATIC_BYTES) in msg and "158761" in msg, (rc, msg)
    # ... nor does a start list of 60 000 entries on a small mesh
    rc = lib.nrw_fluence_scores_host_static(h, 901, 900, 930, p, p, None, 0, 0, 60000, p, p, p, 0, 100, 1, 64,
                                            C.byref(cv), p, p, None)
    assert rc == -3, lib.nrw_last_error().decode()
    # bad arguments are EINVAL, as in the allocating entry
    assert lib.nrw_fluence_scores_host_static(None, 901, 900, 930, p, p, None, 0, 0, 10, p, p, None, 0, 10, 1, 64,
                                              C.byref(cv), p, p, None) == -1
