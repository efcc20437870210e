"""The C-ABI library loads on a CPU-only box and exports every symbol include/nrw.h declares.
No compute calls here (no GPU): nrw_create must fail cleanly with a message, never crash."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT

HEADER = os.path.join(ROOT, "include", "nrw.h")


def _declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"NRW_API\s+[\w\s\*]+?\b(nrw_\w+)\s*\(", text)))


def test_header_declares_the_documented_entry_points():
    syms = _declared_symbols()
    assert len(syms) >= 13
    for s in ("nrw_create", "nrw_walk_philox", "nrw_walk_replay", "nrw_assemble_events",
              "nrw_expand_counts", "nrw_fluence_scores_host", "nrw_last_error"):
        assert s in syms


def test_library_exports_every_declared_symbol():
    from nrw import engine
    lib = engine.load_library()
    for s in _declared_symbols():
        assert hasattr(lib, s), "libnrw.so does not export %s" % s
    assert set(_declared_symbols()) == set(engine.EXPORTS)
    assert lib.nrw_abi_version() == 5


def test_struct_mirrors_match_header_field_order():
    """ctypes mirrors list the same fields, in order, as the C structs."""
    from nrw import engine
    text = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    for cname, mirror in (("nrw_tables", engine._Tables), ("nrw_walk", engine._Walk),
                          ("nrw_replay", engine._Replay), ("nrw_reduce", engine._Reduce)):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (cname, cname), text, re.S).group(1)
        fields = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            for part in decl.split(","):
                name = re.findall(r"(\w+)\s*(?:\[\d+\])?\s*$", part.strip())[0]
                fields.append(name)
        assert fields == [f[0] for f in mirror._fields_], cname


def test_create_without_gpu_fails_with_a_message():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from nrw import engine
    lib = engine.load_library()
    h = C.c_void_p()
    rc = lib.nrw_create(0, C.byref(h))
    assert rc < 0 and not h.value
    assert len(lib.nrw_last_error()) > 0
    assert lib.nrw_destroy(None) == 0
    # NULL-argument calls return EINVAL instead of crashing
    assert lib.nrw_walk_philox(None, None, None, None) == -1
    assert lib.nrw_walk_replay(None, None, None, None) == -1
    assert lib.nrw_device_info(None, None, None) == -1


def _phases(lib, n, rng, warps, range_tail, tail_ranges):
    import numpy as np
    start = np.zeros(9, np.int64); ticket = np.zeros(9, np.int64); r = np.zeros(8, np.int64)
    k = C.c_int32()
    rc = lib.nrw_guided_phases(n, rng, warps, range_tail, tail_ranges, start.ctypes.data, ticket.ctypes.data,
                               r.ctypes.data, C.byref(k))
    assert rc == 0, lib.nrw_last_error()
    return start[:k.value + 1], ticket[:k.value + 1], r[:k.value]


def test_guided_range_phases_tile_the_id_space():
    """The walk hands out its walker ids as tickets -> ranges of guided size (walk_philox.cuh::fetch_range,
    planner nrw_guided_phases): every id must belong to exactly one ticket, whatever the sizes."""
    import numpy as np
    from hypothesis import given, settings, strategies as st
    from nrw import engine
    lib = engine.load_library()

    # the bench launch: 1.25e8 ids, 148 x 32 warps, big ranges of 4096 ids (six per warp, capped), tail down to 1024
    start, ticket, r = _phases(lib, 125_000_000, 4096, 4736, 1024, 2)
    assert list(r) == [4096, 2048, 1024] and start[0] == 0 and start[-1] == 125_000_000
    assert start[1] == 125_000_000 - 2 * 4096 * 4736                     # the tail: two big ranges per warp
    assert ticket[-1] < 12 * 4736                                        # ~10 tickets per warp

    @settings(max_examples=300, deadline=None)
    @given(n=st.integers(0, 200_000), rng=st.integers(1, 5000), warps=st.integers(1, 64),
           rt=st.integers(0, 6000), tr=st.integers(-1, 5))
    def check(n, rng, warps, rt, tr):
        start, ticket, r = _phases(lib, n, rng, warps, rt, tr)
        k = len(r)
        assert 1 <= k <= 8 and start[0] == 0 and start[-1] == n and ticket[0] == 0
        assert (np.diff(start) >= 0).all() and (np.diff(r) <= 0).all() and (r >= 1).all()
        covered = np.zeros(n, np.int32)
        for t in range(int(ticket[-1])):                                 # fetch_range's mapping
            j = int(np.searchsorted(ticket, t, side="right") - 1)
            a = start[j] + (t - ticket[j]) * r[j]
            assert a < start[j + 1]
            covered[a:min(a + r[j], start[j + 1])] += 1
        assert (covered == 1).all()
    check()
    assert lib.nrw_guided_phases(-1, 32, 1, 32, 2, None, None, None, None) == -1
