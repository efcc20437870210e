"""Engine hygiene the C ABI promises (include/nrw.h, "threading"): every launch owns its ticket slot, the walk's
tail hands the slot back / zeroes the next tally buffer / (N > 1) sums over ranks, and the measurement switches
are resolved once per engine."""
import ctypes as C

import numpy as np
import pytest

from conftest import C1

pytestmark = pytest.mark.gpu


def _c1_tables(engine, c1_csr):
    from nrw import mesh, tables
    src = mesh.source_values(c1_csr.meta["mesh"]["meshpoints"], mesh.source_default)
    nc, _ = tables.classify_source(src, c1_csr.n_rows)
    return engine.upload(c1_csr, node_class=nc)


def test_two_streams_on_one_engine_do_not_share_tickets(engine, c1_csr):
    """Two launches of ONE engine overlapping on two streams (round 1 shared one ticket counter per engine: the
    second launch's reset rewound the first one's id hand-out).  Long launches so that they really overlap;
    tallies must equal the serial ones bit for bit."""
    import torch
    tabs = _c1_tables(engine, c1_csr)
    starts = np.arange(900)
    M = 20000                                              # 1.8e7 walkers per launch: ~0.1 ms each
    jobs = []
    for seed in (11, 12, 13, 14):
        pw, res = engine.prepare_walk(tabs, starts, M, 30000, seed, score=True, class_value=[0, 1, 0, 0])
        pw()
        torch.cuda.synchronize()
        jobs.append((pw, res, res.score.clone(), res.counters.clone()))
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for rep in range(6):
        for pw, res, _, _ in jobs:
            pw.zero()
        torch.cuda.synchronize()
        for i, (pw, res, _, _) in enumerate(jobs):
            with torch.cuda.stream(s1 if i % 2 == 0 else s2):
                pw.launch_kernel()
        torch.cuda.synchronize()
        for pw, res, score, counters in jobs:
            assert torch.equal(res.score, score), rep
            assert torch.equal(res.counters, counters), rep


def test_ticket_slots_are_handed_back(engine, c1_csr):
    """More launches than the ring has slots (NRW_TICKET_SLOTS = 1024): a slot that its launch did not reset would
    start the next launch on it with a stale ticket and walk fewer ids."""
    import torch
    tabs = _c1_tables(engine, c1_csr)
    pw, res = engine.prepare_walk(tabs, np.arange(0, 900, 9), 40, 64, 3, score=True, class_value=[0, 1, 0, 0])
    pw()
    want = (res.score.clone(), res.counters.clone())
    for _ in range(2100):
        pw.launch_kernel()
    torch.cuda.synchronize()
    assert torch.equal(res.score, want[0] * 2101) and torch.equal(res.counters, want[1] * 2101)


def test_tally_exchange_single_rank_alternates_and_zeroes(engine, c1_csr):
    """nrw.dist.TallyExchange at world = 1: the walk's tail zeroes the buffer of the NEXT launch, so a loop of
    solves needs no memset and every solve equals the plain walk."""
    import torch
    from nrw import dist as ndist
    tabs = _c1_tables(engine, c1_csr)
    starts = np.arange(0, 900, 3)
    ex = ndist.TallyExchange(len(starts), engine)
    assert ex.fused
    pws = [engine.prepare_walk(tabs, starts, 500, 64, 0, score=ex.scores(h), class_value=[0, 1, 0, 0],
                               counters=ex.counters(h))[0] for h in (0, 1)]
    for k, seed in enumerate((5, 6, 7, 8, 9)):
        for pw in pws:
            pw.set_seed(seed)
        out = ex.launch(pws).clone()
        ref = engine.walk(tabs, starts, 500, 64, seed, score=True, class_value=[0, 1, 0, 0])
        assert torch.equal(out[:len(starts)], ref.score), k
        assert torch.equal(out[len(starts):], ref.counters), k
        torch.cuda.synchronize()
        assert not ex.pair.half((k & 1) ^ 1).any()         # the other half is clean for the next solve


def test_switches_are_read_once_per_engine(c1_csr, monkeypatch):
    """nrw_create resolves the NRW_* switches; changing the environment afterwards changes nothing."""
    from nrw.engine import Engine
    monkeypatch.delenv("NRW_K1", raising=False)
    eng = Engine(0)
    general = Engine(0, switches={"NRW_K1": "general", "NRW_DET": "0"})
    assert "k1=auto" in eng.config and "det=1" in eng.config
    assert "k1=general" in general.config and "det=0" in general.config
    monkeypatch.setenv("NRW_K1", "general")
    tabs = _c1_tables(eng, c1_csr)
    a = eng.walk(tabs, [435, 14], [300, 200], 64, 2, score=True, class_value=[0, 1, 0, 0])
    b = general.walk(tabs, [435, 14], [300, 200], 64, 2, score=True, class_value=[0, 1, 0, 0])
    assert "walk_score" in a.kernel and "walk_philox<" in b.kernel
    assert np.array_equal(a.score.cpu().numpy(), b.score.cpu().numpy())
    assert a.counters_dict() == b.counters_dict()


def test_reduce_descriptor_is_validated(engine, c1_csr):
    """score / counters outside the exchange buffer are refused (the tail would sum something else)."""
    import torch
    from nrw import dist as ndist
    from nrw.engine import NrwError
    tabs = _c1_tables(engine, c1_csr)
    ex = ndist.TallyExchange(3, engine)
    pw, _ = engine.prepare_walk(tabs, [1, 2, 3], 10, 64, 0, score=True, class_value=[0, 1, 0, 0])
    h, rp = ex.arm()
    pw.w.reduce = rp
    with pytest.raises(NrwError):
        pw.launch_kernel()
    pw.w.reduce = C.POINTER(type(rp.contents))()
    pw.launch_kernel()
    torch.cuda.synchronize()
