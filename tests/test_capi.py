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
    assert lib.nrw_abi_version() == 2


def test_struct_mirrors_match_header_field_order():
    """ctypes mirrors list the same fields, in order, as the C structs."""
    from nrw import engine
    text = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    for cname, mirror in (("nrw_tables", engine._Tables), ("nrw_walk", engine._Walk),
                          ("nrw_replay", engine._Replay)):
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
