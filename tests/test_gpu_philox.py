"""GPU parity, statistical mode (K1 walk_philox):
  (a) integer tallies bit-exact against the numpy port of the device sampler on the same
      (seed, walker id) streams -- every table location (smem / hybrid / global / cdf) and tally;
  (b) agreement with the reference's algorithm through the exact Markov-chain expectation of its
      matrix: per-cell |rwSol - E| <= 4.5 sigma/sqrt(M) (all cells; 3 sigma for >= 99 %), and
      relative L2 <= 1e-2 (BASELINE north_star tolerance) once M is large enough for that to be
      above the Monte-Carlo floor (M = 1e5: expected 2.4e-3);
  (c) size-independent properties at the full BASELINE mesh sizes (conservation, checksum of
      checksums, invariance under sharding)."""
import numpy as np
import pytest

from conftest import C1

pytestmark = pytest.mark.gpu


def _classes(csr, source):
    from nrw import mesh, tables
    src = mesh.source_values(csr.meta["mesh"]["meshpoints"], source)
    nc, cv = tables.classify_source(src, csr.n_rows)
    return src, nc, cv


def _cmp(res, port, score=True, hist=True, count=False):
    if score:
        assert np.array_equal(res.score.cpu().numpy(), port["score"])
    if hist:
        assert np.array_equal(res.node_hist.cpu().numpy(), port["node_hist"])
    if count:
        got = res.count_tn.cpu().numpy().astype(np.int64)
        ref = port["count_tn"].copy()
        ref[:, 0, :] = 0
        assert np.array_equal(got, ref)
    c = res.counters_dict()
    for k in ("steps", "exits", "survivors", "walkers"):
        assert c[k] == port["counters"][k], k


def test_smem_tables_score_hist_bit_exact(engine, c1_csr):
    from nrw import mesh
    from oracle import philox_port
    _, nc, _ = _classes(c1_csr, mesh.source_default)
    tabs = engine.upload(c1_csr, node_class=nc)
    assert tabs.alias_host.layout == "compact"
    starts = np.arange(900)
    for seed, M, nt in ((1, 37, 64), (2 ** 40 + 5, 5, 6)):
        res = engine.walk(tabs, starts, M, nt, seed, score=True, score_sq=True,
                          class_value=[0, 1, 0, 0], node_hist=True)
        assert "TAB_SMEM" in res.kernel
        port = philox_port.walk_port(c1_csr, starts, M, nt, seed, alias=tabs.alias_host,
                                     class_value=[0, 1, 0, 0])
        _cmp(res, port)
        assert np.allclose(res.score_sq.cpu().numpy(), port["score_sq"], rtol=1e-12)


def test_ragged_starts_and_walker_ranges(engine, c1_csr):
    """Ragged / empty start counts and arbitrary [lo, hi) id windows (the sharding primitive)."""
    from oracle import philox_port
    tabs = engine.upload(c1_csr)
    starts = np.array([435, 0, 900, 14, 14, 899, 450])          # 900 = the sink itself
    counts = np.array([1000, 0, 17, 1, 333, 64, 5])
    total = int(counts.sum())
    for rng_ in ((0, total), (0, 1), (999, 1001), (1017, 1018), (total - 1, total), (5, 5)):
        res = engine.walk(tabs, starts, counts, 50, 9, walker_range=rng_, count_tn=True, t_cap=20,
                          count_per_start=True)
        port = philox_port.walk_port(c1_csr, starts, counts, 50, 9, alias=tabs.alias_host,
                                     walker_range=rng_, t_cap=20, count_per_start=True)
        _cmp(res, port, score=False, hist=False, count=True)
        # tallies dropped beyond t_cap are counted, not lost silently
        c = res.counters_dict()
        assert c["t_overflow"] == port["counters"]["steps"] - int(port["count_tn"][:, 1:, :].sum())


def test_time_weighted_score_bit_exact(engine, c1_csr):
    from nrw import mesh
    from nrw.dropin import exp_time_weights
    from oracle import philox_port
    _, nc, cv = _classes(c1_csr, mesh.source_paper)
    tabs = engine.upload(c1_csr, node_class=nc)
    nt = 40
    tw = exp_time_weights(0.7, 0.01, nt)
    cvq = cv * 4294967296.0
    starts = np.arange(0, 900, 3)
    res = engine.walk(tabs, starts, 64, nt, 77, score=True, class_value=cvq, time_weight=tw)
    port = philox_port.walk_port(c1_csr, starts, 64, nt, 77, alias=tabs.alias_host, class_value=cvq,
                                 time_weight=tw)
    _cmp(res, port, hist=False)


def test_cdf_sampler_bit_exact(engine, c1_csr):
    """The reference's own draw (searchsorted-left on fp64 prefix sums, u from 53 random bits)."""
    from nrw import mesh, tables
    from oracle import philox_port
    _, nc, _ = _classes(c1_csr, mesh.source_default)
    tabs = engine.upload(c1_csr, node_class=nc)
    starts = np.arange(0, 900, 5)
    res = engine.walk(tabs, starts, 40, 64, 3, sampler="cdf", score=True, class_value=[0, 1, 0, 0],
                      node_hist=True)
    assert "TAB_CDF" in res.kernel
    port = philox_port.walk_port(c1_csr, starts, 40, 64, 3, sampler="cdf", cum=tables.row_cum(c1_csr),
                                 node_class=nc, class_value=[0, 1, 0, 0])
    _cmp(res, port)


def test_wide_global_tables_bit_exact(engine, c1_csr):
    from nrw import mesh
    from oracle import philox_port
    _, nc, _ = _classes(c1_csr, mesh.source_default)
    tabs = engine.upload(c1_csr, node_class=nc, alias_layout="wide")
    starts = np.arange(0, 900, 5)
    res = engine.walk(tabs, starts, 40, 64, 4, score=True, class_value=[0, 1, 0, 0], node_hist=True)
    assert "TAB_GLB" in res.kernel
    port = philox_port.walk_port(c1_csr, starts, 40, 64, 4, alias=tabs.alias_host, class_value=[0, 1, 0, 0])
    _cmp(res, port)


def test_hybrid_tables_bit_exact(engine, monkeypatch):
    """38 441-row mesh: alias patterns fit shared memory, the row descriptors (300 KB) do not: they stay
    in L2 (TAB_HYB, forced here) or are derived from the factored tables (TAB_FAC, the planner's choice)."""
    from nrw import mesh
    from oracle import philox_port
    csr = mesh.build_fluence_csr(0.001, 1.0 / 31.0, 200, 0.15, 1)
    _, nc, _ = _classes(csr, mesh.source_default)
    tabs = engine.upload(csr, node_class=nc)
    starts = np.arange(0, csr.n_live, 97)
    port = philox_port.walk_port(csr, starts, 16, 400, 6, alias=tabs.alias_host, class_value=[0, 1, 0, 0])
    for mode, want in (("hyb", "TAB_HYB"), (None, "TAB_FAC")):
        if mode:
            monkeypatch.setenv("NRW_TAB", mode)
        else:
            monkeypatch.delenv("NRW_TAB", raising=False)
        res = engine.walk(tabs, starts, 16, 400, 6, score=True, class_value=[0, 1, 0, 0], node_hist=True)
        assert want in res.kernel, res.kernel
        _cmp(res, port)


def test_simple_flux_counts_smem_bit_exact(engine):
    """2-state chain with absorption (simple_flux_miniapp.py:174-203): time-resolved counts in the CTA's shared-memory
    histogram, against the numpy port -- the trimmed count walk with plain atomics (long windows: the C3 shape), the
    general kernel on the same launch (NRW_K1=general), and the general kernel's __match_any_sync aggregation (windows
    of <= 256 steps); per-start and pooled bins, a window shorter than the walk (overflow counter), survivors."""
    from nrw import mesh
    from nrw.engine import Engine
    from oracle import philox_port
    csr = mesh.csr_from_dense(mesh.build_simple_flux_dense(0.005, 8.0, 2.0), n_live=2)
    tabs = engine.upload(csr)
    general = Engine(0, switches={"NRW_K1": "general"})
    for nt, t_cap, per_start, want in ((300, 300, True, "walk_count_smem<"), (1002, 1002, False, "walk_count_smem<"),
                                       (400, 290, True, "walk_count_smem<"), (12, 300, True, ",16>"),
                                       (200, 200, True, ",16>")):
        port = philox_port.walk_port(csr, [0, 1], [5000, 3000], nt, 8, alias=tabs.alias_host, t_cap=t_cap,
                                     count_per_start=per_start)
        res = engine.walk(tabs, [0, 1], [5000, 3000], nt, 8, count_tn=True, t_cap=t_cap, count_per_start=per_start)
        assert want in res.kernel, res.kernel
        _cmp(res, port, score=False, hist=False, count=True)
        res_g = general.walk(tabs, [0, 1], [5000, 3000], nt, 8, count_tn=True, t_cap=t_cap, count_per_start=per_start)
        assert ",16>" in res_g.kernel, res_g.kernel
        _cmp(res_g, port, score=False, hist=False, count=True)


def test_sharded_by_start_slices_equals_unsharded(engine, c1_csr):
    """What nrw.dist does on N GPUs, emulated on one: each 'rank' walks a slice of every start's
    walkers with start_id_base keeping each walker's stream; integer sums must match exactly."""
    from nrw import dist, mesh
    _, nc, _ = _classes(c1_csr, mesh.source_default)
    tabs = engine.upload(c1_csr, node_class=nc)
    starts, M, nt, seed = np.arange(900), 50, 64, 21
    full = engine.walk(tabs, starts, M, nt, seed, score=True, class_value=[0, 1, 0, 0], node_hist=True)
    acc_s, acc_h, acc_c = 0, 0, 0
    for world in (3,):
        for rank in range(world):
            lo, hi = dist.shard_range(M, rank, world)
            r = engine.walk(tabs, starts, hi - lo, nt, seed, score=True, class_value=[0, 1, 0, 0],
                            node_hist=True, start_id_base=starts * M + lo)
            acc_s = acc_s + r.score
            acc_h = acc_h + r.node_hist
            acc_c = acc_c + r.counters
    assert bool((acc_s == full.score).all()) and bool((acc_h == full.node_hist).all())
    assert bool((acc_c == full.counters).all())


def test_fluence_philox_matches_reference_expectation(engine, c1_csr):
    """North-star acceptance: rwSol per (x, mu) cell within the Monte-Carlo error of the
    reference's Markov chain; relative L2 <= 1e-2."""
    from nrw import mesh
    from nrw.fluence import solve_fluence
    from oracle import expectation
    nt = 200
    for source in (mesh.source_default, mesh.source_paper):
        src = mesh.source_values(c1_csr.meta["mesh"]["meshpoints"], source)
        mean, var = expectation.fluence_moments(c1_csr.indptr, c1_csr.cols, c1_csr.vals, c1_csr.n_rows,
                                                c1_csr.n_live, src, nt, 0.0, C1["dt"], C1["v"])
        for M, l2_max in ((1000, 5e-2), (100000, 1e-2)):
            res = solve_fluence(c1_csr, M, nt, 0, C1["dt"], C1["v"], source=source, rng="philox",
                                seed=1234 + M, want_stderr=True)
            sigma = np.sqrt(var / M)
            z = np.abs(res.per_start - mean) / np.where(sigma > 0, sigma, 1.0)
            assert (res.per_start[sigma == 0] == mean[sigma == 0]).all()
            assert z.max() < 4.5, z.max()
            assert (z < 3.0).mean() > 0.99
            rel_l2 = np.linalg.norm(res.per_start - mean) / np.linalg.norm(mean)
            assert rel_l2 < l2_max, (M, rel_l2)
            # the tallied second moment reproduces the exact per-walker variance
            if M >= 100000:
                big = sigma > 0
                ratio = res.stderr[big] / sigma[big]
                assert abs(np.median(ratio) - 1.0) < 0.01 and np.abs(ratio - 1.0).max() < 0.25


def test_fused_solve_path_equals_the_general_launch(engine, c1_csr):
    """solve_fluence's prepared path (one launch descriptor, scores + counters in one buffer, pinned copy
    back; taken when no per-walker variance is asked for) returns what the per-call path returns, and a
    re-used descriptor carries nothing over from the previous seed."""
    from nrw import mesh
    from nrw.fluence import solve_fluence
    args = (c1_csr, 700, 64, 0, C1["dt"], C1["v"])
    for source in (mesh.source_default, mesh.source_paper):
        ref = {seed: solve_fluence(*args, source=source, rng="philox", seed=seed, want_stderr=True) for seed in (3, 4)}
        for seed in (3, 4, 3):
            res = solve_fluence(*args, source=source, rng="philox", seed=seed)
            assert res.stderr is None and "walk_score" in res.kernel
            assert np.array_equal(res.per_start, ref[seed].per_start)
            assert np.array_equal(res.rwSol, ref[seed].rwSol)
            assert res.walker_steps == ref[seed].walker_steps
            for k in ("steps", "exits", "survivors", "walkers"):
                assert res.counters[k] == ref[seed].counters[k], k


def test_fluence_philox_time_weighted_matches_expectation(engine, c1_csr):
    from nrw import mesh
    from nrw.fluence import solve_fluence
    from oracle import expectation
    nt, SigA, M = 120, 0.9, 50000
    src = mesh.source_values(c1_csr.meta["mesh"]["meshpoints"], mesh.source_default)
    mean, var = expectation.fluence_moments(c1_csr.indptr, c1_csr.cols, c1_csr.vals, c1_csr.n_rows,
                                            c1_csr.n_live, src, nt, SigA, C1["dt"], C1["v"])
    res = solve_fluence(c1_csr, M, nt, SigA, C1["dt"], C1["v"], rng="philox", seed=5)
    sigma = np.sqrt(var / M)
    z = np.abs(res.per_start - mean) / np.where(sigma > 0, sigma, 1.0)
    assert z.max() < 4.5
    assert np.linalg.norm(res.per_start - mean) / np.linalg.norm(mean) < 1e-2


def test_simple_flux_philox_matches_closed_form(tmp_path):
    """The reference's only known answer (simple_flux_miniapp.py:338-343).  With M = 2e6 the walk
    must sit within 5 sigma of the exact propagation of the reference's chain, and within the
    chain's own discretisation bias (0.0675) + noise of the continuous closed form."""
    import simple_flux_miniapp as app
    from nrw import mesh
    M = 2000000
    a = vars(app.build_parser().parse_args(["-r", "pc", "--rng", "philox", "-M", str(M)]))
    a.update(matrix_filename=str(tmp_path / "m.csv"), results_filename=str(tmp_path / "r.csv"))
    np.random.seed(4)
    sol = app.run(a)
    assert sol.shape == (2, 1002)
    t, true = app.true_solution(a, sol.shape[1])
    assert np.abs(sol - true).max() < 0.08
    q = mesh.build_simple_flux_dense(0.005, 8.0, 2.0)[:2, :2]
    p = np.eye(2)
    w = np.array([5.0, 3.0])
    for j in range(1002):
        mean = p @ w
        var = p @ (w * w) - mean ** 2            # per-walker variance of ip*1[+] + in*1[-]
        assert (np.abs(sol[:, j] - mean) <= 5.0 * np.sqrt(np.maximum(var, 0) / M) + 1e-12).all(), j
        p = q @ p
    # --use_sinks must not change rwSol (Q14)
    a["use_sinks"] = True
    np.random.seed(4)
    assert np.array_equal(app.run(a), sol)


def test_full_size_mesh_properties(engine):
    """BASELINE config C5 mesh (10 001 rows, nnz 505 000) at 1e7 walkers: conservation and a
    checksum of checksums (sum of per-start scores == source-weighted occupancy histogram)."""
    from nrw import mesh
    csr = mesh.build_fluence_csr(0.01, 0.02, 200, 0.15, 1)
    assert csr.n_rows == 10001 and csr.nnz == 505000
    _, nc, _ = _classes(csr, mesh.source_default)
    tabs = engine.upload(csr, node_class=nc)
    starts, M, nt = np.arange(10000), 1000, 30000
    res = engine.walk(tabs, starts, M, nt, 1, score=True, class_value=[0, 1, 0, 0], node_hist=True)
    assert "TAB_SMEM" in res.kernel
    c = res.counters_dict()
    assert c["walkers"] == 10000 * M
    assert c["exits"] + c["survivors"] == c["walkers"]
    hist = res.node_hist.cpu().numpy()
    assert hist.sum() == c["steps"] and c["survivors"] == 0
    assert hist[10000] == c["exits"]
    assert int(res.score.sum().item()) == int(hist[:10000][nc[:10000] == 1].sum())
    mean_steps = c["steps"] / c["walkers"]
    assert 2.0 < mean_steps < 2.6          # 2.23 steps per walker


def test_run_pc_philox_dropin_shapes(tmp_path, c1_csr):
    """The drop-in signature in Philox mode: float64 [meshSize+1, nt], column 0 = initial walkers,
    every column conserves walkers until they exit."""
    import fluence_mini_app as app
    from nrw import mesh
    p = str(tmp_path / "curmat.csv")
    mesh.save_matrix_csv(p, c1_csr)
    np.random.seed(0)
    td = app.run_pc(matrix_filename=p, neuraltime=3000, initial_walkers=[[435, 5000]], meshSize=900,
                    runtime_args={"rng": "philox"})
    assert td.shape == (901, 3000) and td.dtype == np.float64
    assert td[435, 0] == 5000 and td[:, 0].sum() == 5000
    alive = td[:900].sum(axis=0)
    exits = td[900]
    assert np.array_equal(alive[1:] + exits[1:], alive[:-1])
    assert alive[-1] == 0


def test_errors(engine, c1_csr, tmp_path):
    import fluence_mini_app as app
    from nrw.engine import NrwError
    with pytest.raises(ValueError):
        a = vars(app.build_parser().parse_args(["-r", "pc", "-dt", "0.05"]))
        a.update(matrix_filename=str(tmp_path / "m.csv"), results_filename=str(tmp_path / "r.csv"))
        app.run(a)
    with pytest.raises(NotImplementedError):
        app.run_functions["spinnaker"](matrix_filename="x", neuraltime=3, initial_walkers=[[0, 1]], meshSize=900)
    tabs = engine.upload(c1_csr)
    with pytest.raises(IndexError):
        engine.walk(tabs, [901], 1, 10, 0, node_hist=True)
    with pytest.raises(NrwError):
        engine.walk(tabs, [0], 1, 10, 0)                       # no tally requested
    with pytest.raises(NrwError):
        engine.walk(tabs, [0], 1, 1 << 25, 0, node_hist=True)  # nt beyond the counter's 24 bits


# ---------------------------------------------------------------- K1s: the score-only fast path
@pytest.fixture
def k1_env(monkeypatch):
    def set_variant(variant):
        monkeypatch.delenv("NRW_K1", raising=False)
        monkeypatch.delenv("NRW_DET", raising=False)
        monkeypatch.delenv("NRW_CPS", raising=False)
        # "fast*": the trimmed kernel on the alias tables (NRW_STR=0); "str*": its structured draw, which the
        # planner takes by default wherever the tables have the closed form of the fluence matrix
        if variant.startswith("str"):
            monkeypatch.delenv("NRW_STR", raising=False)
        else:
            monkeypatch.setenv("NRW_STR", "0")
        if variant == "str_nodet":
            monkeypatch.setenv("NRW_DET", "0")
        if variant == "fast_cps2":
            monkeypatch.setenv("NRW_CPS", "2")
        if variant == "general":
            monkeypatch.setenv("NRW_K1", "general")
        elif variant == "fast_nodet":
            monkeypatch.setenv("NRW_DET", "0")
    return set_variant


@pytest.mark.parametrize("variant", ["fast", "fast_cps2", "fast_nodet", "general", "str", "str_nodet"])
def test_score_fast_path_bit_exact_all_variants(engine, c1_csr, k1_env, variant):
    """walk_score_kernel and the general kernel produce the same integers as the numpy port:
    identity and packed class weights, ragged starts with the sink and empty starts in the list,
    sub-ranges, sharded stream ids (incl. a 2^32 carry inside a start), nt edge cases."""
    from nrw import mesh
    from oracle import philox_port
    k1_env(variant)
    _, nc1, _ = _classes(c1_csr, mesh.source_default)                    # classes {0, 1}
    nc3 = (np.arange(c1_csr.n_rows) % 4).astype(np.uint8)                # classes {0, 1, 2, 3}
    nc3[900] = 2                                                         # a weighted sink
    want = "walk_philox<" if variant == "general" else ("walk_score_str<TAB_SMEM" if variant.startswith("str") else "walk_score<TAB_SMEM")
    starts = np.array([435, 0, 900, 14, 14, 899, 450, 16, 464])
    counts = np.array([3000, 0, 17, 1, 333, 64, 5, 2500, 1200])
    total = int(counts.sum())
    carry = np.arange(9) * 2 ** 33 + (2 ** 32 - 1500)                    # ids cross a multiple of 2^32
    for nc, cv, idw in ((nc1, [0, 1, 0, 0], True), (nc1, [0, 7, 0, 0], False), (nc3, [0, 1, 2, 3], True),
                        (nc3, [5, 0, 255, 1], False)):
        tabs = engine.upload(c1_csr, node_class=nc)
        for nt, rng_, idb in ((64, None, None), (64, (2999, 3400), None), (3, None, None), (2, None, None),
                              (1, None, None), (40, None, carry), (64, (total - 1, total), None)):
            res = engine.walk(tabs, starts, counts, nt, 5, score=True, class_value=cv, walker_range=rng_,
                              start_id_base=idb)
            if nt >= 2:
                assert want in res.kernel, res.kernel
                if variant != "general":
                    assert (",true," if idw else ",false,") in res.kernel, res.kernel
                    assert (",false,cps=" if variant in ("fast_nodet", "str_nodet") else ",true,cps=") in res.kernel, res.kernel
                    assert res.kernel.endswith("cps=2>" if variant == "fast_cps2" else "cps=1>"), res.kernel
            else:
                assert "walk_philox<" in res.kernel                      # nt = 1: no step at all
            port = philox_port.walk_port(c1_csr, starts, counts, nt, 5, alias=tabs.alias_host, class_value=cv,
                                         walker_range=rng_, start_id_base=idb)
            _cmp(res, port, hist=False)
            res = engine.walk(tabs, starts, counts, nt, 5, score=True, class_value=cv, walker_range=rng_,
                              start_id_base=idb, node_hist=True)                  # + occupancy histogram
            if nt >= 2:
                assert ("walk_philox<" if variant == "general" else "walk_score_hist_str<TAB_SMEM"
                        if variant.startswith("str") else "walk_score_hist<TAB_SMEM") in res.kernel
            _cmp(res, port, hist=True)
    # a long uniform start list: many range fetches and start changes per warp
    tabs = engine.upload(c1_csr, node_class=nc1)
    res = engine.walk(tabs, np.arange(900), 777, 64, 6, score=True, class_value=[0, 1, 0, 0])
    port = philox_port.walk_port(c1_csr, np.arange(900), 777, 64, 6, alias=tabs.alias_host,
                                 class_value=[0, 1, 0, 0])
    _cmp(res, port, hist=False)


def test_score_fast_path_hybrid_tables_and_survivors(engine, k1_env, monkeypatch):
    """Row descriptors in L2 (TAB_HYB) or derived (TAB_FAC) + walkers that outlive nt (survivors; the
    step t = nt-1 is outside the solution sum, fluence_mini_app.py:428)."""
    from nrw import mesh
    from oracle import philox_port
    k1_env("fast")
    csr = mesh.build_fluence_csr(0.001, 1.0 / 31.0, 200, 0.15, 1)
    _, nc, _ = _classes(csr, mesh.source_default)
    tabs = engine.upload(csr, node_class=nc)
    starts = np.arange(0, csr.n_live, 97)
    for nt in (400, 9):                                  # nt = 9: most walkers survive
        port = philox_port.walk_port(csr, starts, 48, nt, 6, alias=tabs.alias_host, class_value=[0, 1, 0, 0])
        for mode, want in (("hyb", "walk_score<TAB_HYB"), (None, "walk_score<TAB_FAC")):
            if mode:
                monkeypatch.setenv("NRW_TAB", mode)
            else:
                monkeypatch.delenv("NRW_TAB", raising=False)
            for st in ("0", None):                       # alias tables, then the structured draw on the same modes
                if st:
                    monkeypatch.setenv("NRW_STR", st)
                else:
                    monkeypatch.delenv("NRW_STR", raising=False)
                res = engine.walk(tabs, starts, 48, nt, 6, score=True, class_value=[0, 1, 0, 0])
                assert (want if st else want.replace("walk_score<", "walk_score_str<")) in res.kernel, res.kernel
                _cmp(res, port, hist=False)
            monkeypatch.setenv("NRW_STR", "0")
    assert port["counters"]["survivors"] > 0


def _full_size_scores(engine, k1_env, variants):
    from nrw import mesh
    csr = mesh.build_fluence_csr(0.01, 0.02, 200, 0.15, 1)
    _, nc, _ = _classes(csr, mesh.source_default)
    tabs = engine.upload(csr, node_class=nc)
    out = {}
    for variant in variants:
        k1_env(variant)
        r = engine.walk(tabs, np.arange(10000), 1000, 30000, 3, score=True, class_value=[0, 1, 0, 0])
        out[variant] = (r.score.cpu().numpy(), r.counters_dict(), r.kernel)
    return out


def test_score_fast_path_equals_general_kernel_at_full_size(engine, k1_env):
    """C5 mesh, 1e7 walkers: the trimmed kernel and the general kernel agree on every score."""
    out = _full_size_scores(engine, k1_env, ("fast", "fast_nodet", "general", "str", "str_nodet"))
    assert "walk_score<" in out["fast"][2] and "walk_philox<" in out["general"][2] and "walk_score_str<" in out["str"][2]
    for v in ("fast", "fast_nodet", "str", "str_nodet"):
        assert np.array_equal(out[v][0], out["general"][0])
        assert out[v][1] == out["general"][1]


def test_deterministic_rows_on_a_general_graph(engine, k1_env):
    """Single-successor rows of every kind: successor absorbing (folded without a draw), successor
    live (stepped normally), chains of them, a deterministic start, weighted absorbing rows."""
    from nrw import mesh, tables
    from oracle import philox_port
    rng = np.random.default_rng(12)
    n = 60
    q = np.zeros((n, n))
    for i in range(n):
        k = int(rng.integers(2, 7))
        cols = rng.choice(n, size=k, replace=False)
        q[i, cols] = rng.dirichlet(np.ones(k))
    q[50:] = 0                                   # rows 50..59 absorbing
    for r, tgt in ((5, 55), (6, 5), (7, 6), (8, 59), (9, 20), (10, 10)):
        q[r] = 0; q[r, tgt] = 1.0                # 5 -> sink, 6 -> 5 -> sink, 7 -> 6 -> ..., 9 -> live row, 10 self loop
    csr = mesh.csr_from_dense(q, n_live=50)
    nc = (np.arange(n) % 4).astype(np.uint8)
    at = tables.build_alias(csr, node_class=nc)
    w0 = at.rowinfo[:, 0]
    detabs = (w0 >> 25) & 1
    assert detabs[5] == 1 and detabs[8] == 1 and detabs[6] == 0 and detabs[9] == 0 and detabs[10] == 0
    cls, inc = (w0 >> 27) & 3, w0 >> 29                      # inc = cls + class of the absorbing successor
    assert np.array_equal(cls, nc) and inc[5] - cls[5] == nc[55] and inc[8] - cls[8] == nc[59] and inc[6] == cls[6]
    assert np.array_equal(at.rowinfo[:, 1] >> 30, 1 + detabs)  # sinc: steps an arrival accounts for
    tabs = engine.upload(csr, node_class=nc)
    starts = np.array([0, 5, 6, 7, 8, 9, 10, 3, 55, 21])
    counts = np.array([4000, 300, 300, 300, 77, 300, 50, 4000, 9, 2000])
    for variant in ("fast", "fast_nodet", "general"):
        k1_env(variant)
        for cv in ([0, 1, 2, 3], [3, 0, 200, 7]):
            for nt in (40, 2, 3, 4):
                res = engine.walk(tabs, starts, counts, nt, 31, score=True, class_value=cv)
                port = philox_port.walk_port(csr, starts, counts, nt, 31, alias=tabs.alias_host, class_value=cv)
                _cmp(res, port, hist=False)


@pytest.mark.parametrize("seed", list(range(16)))
def test_random_graphs_differential(engine, k1_env, seed):
    """Differential test on random sparse graphs: both kernels == numpy port for random row lengths,
    absorbing rows anywhere, single-successor rows (absorbing and live targets, self loops), random
    classes / weights, ragged start lists, id windows and stream-id bases."""
    from nrw import mesh
    from oracle import philox_port
    rng = np.random.default_rng(1000 + seed)
    n = int(rng.integers(5, 200))
    q = np.zeros((n, n))
    for i in range(n):
        kind = rng.random()
        if kind < 0.15:
            continue                                           # absorbing row
        if kind < 0.35:
            q[i, rng.integers(0, n)] = 1.0                     # single successor (maybe itself)
            continue
        k = int(rng.integers(2, min(n, 12) + 1))
        cols = rng.choice(n, size=k, replace=False)
        q[i, cols] = rng.dirichlet(np.ones(k) * rng.choice([0.3, 1.0, 5.0]))
    n_live = int(rng.integers(max(n // 2, 1), n + 1))
    csr = mesh.csr_from_dense(q, n_live=n_live)
    nc = rng.integers(0, 4, size=n).astype(np.uint8)
    tabs = engine.upload(csr, node_class=nc)
    n_starts = int(rng.integers(1, 12))
    starts = rng.integers(0, n, size=n_starts)
    counts = rng.integers(0, 3000, size=n_starts)
    counts[rng.random(n_starts) < 0.2] = 0
    total = int(counts.sum())
    if total == 0:
        counts[0] = total = 50
    cv = [0, 1, 2, 3] if rng.random() < 0.5 else [int(x) for x in rng.integers(0, 256, size=4)]
    nt = int(rng.integers(2, 50))
    lo = int(rng.integers(0, total))
    window = None if rng.random() < 0.5 else (lo, int(rng.integers(lo, total + 1)))
    idb = None if rng.random() < 0.4 else rng.integers(0, 2 ** 39, size=n_starts)
    if idb is not None and rng.random() < 0.6:                 # stream ids that cross a multiple of 2^32 inside a start
        idb = (np.arange(1, n_starts + 1, dtype=np.int64) << 32) - rng.integers(0, np.maximum(counts, 1) + 1)
    port = philox_port.walk_port(csr, starts, counts, nt, seed, alias=tabs.alias_host, class_value=cv,
                                 walker_range=window, start_id_base=idb)
    for variant in ("fast", "fast_nodet", "general"):
        k1_env(variant)
        res = engine.walk(tabs, starts, counts, nt, seed, score=True, class_value=cv, walker_range=window,
                          start_id_base=idb)
        _cmp(res, port, hist=False)
        # the same walk with the occupancy histogram (walk_score_hist / general kernel), and histogram only
        res = engine.walk(tabs, starts, counts, nt, seed, score=True, class_value=cv, walker_range=window,
                          start_id_base=idb, node_hist=True)
        assert ("walk_philox<" if variant == "general" else "walk_score_hist<") in res.kernel, res.kernel
        _cmp(res, port, hist=True)
        res = engine.walk(tabs, starts, counts, nt, seed, node_hist=True, walker_range=window, start_id_base=idb)
        _cmp(res, port, score=False, hist=True)


def test_microbench_rates_are_sane(engine):
    """nrw_microbench: the roofline denominators bench.py reports (operations per second)."""
    for kind, n, iters in (("smem_load", 20000, 200), ("smem_atomic", 10001, 200), ("l2_load", 158761, 200),
                           ("global_atomic", 10001, 50)):
        rate = engine.microbench(kind, n, iters=iters, reps=2)
        assert 1e9 < rate < 1e14, (kind, rate)
    from nrw.engine import NrwError
    with pytest.raises(NrwError):
        engine.microbench("smem_load", 10 ** 6)                  # 8 MB does not fit shared memory


def test_host_buffer_entry_matches_device_walk(engine, c1_csr):
    """nrw_fluence_scores_host (host buffers in, host buffers out: the call a C / ctypes user makes):
    pinned and pageable buffers, repeated solves over the same buffers: all equal the device-resident
    walk for every seed, id window and weight vector."""
    import torch
    from nrw import mesh
    _, nc, _ = _classes(c1_csr, mesh.source_default)
    tabs = engine.upload(c1_csr, node_class=nc)
    at = tabs.alias_host
    starts = np.array([435, 0, 900, 14, 899, 450], dtype=np.int32)
    counts = np.array([3000, 0, 17, 333, 64, 2500])
    prefix = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    total = int(prefix[-1])

    def pinned(a):
        return torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()

    class Pinned:
        layout, n_alias = "compact", at.n_alias
        rowinfo, alias = pinned(at.rowinfo), pinned(at.alias)

    def check(view, h_starts, h_prefix, score, cnt, lo, hi, nt, cv, seeds):
        solve = engine.prepare_fluence_scores_host(view, c1_csr.n_rows, c1_csr.n_live, h_starts, h_prefix, lo, hi,
                                                   nt, cv, score, cnt)
        for seed in seeds:
            score[:] = -1
            solve(seed)
            ref = engine.walk(tabs, starts, counts, nt, seed, score=True, class_value=cv, walker_range=(lo, hi))
            assert np.array_equal(score, ref.score.cpu().numpy()), (seed, lo, hi, nt, cv)
            c = ref.counters.cpu().numpy()
            assert [int(v) for v in cnt[:5]] == [int(v) for v in c[:5]]

    p_starts, p_prefix = pinned(starts), pinned(prefix)
    p_score, p_cnt = pinned(np.zeros(len(starts), np.int64)), pinned(np.zeros(8, np.uint64))
    check(Pinned, p_starts, p_prefix, p_score, p_cnt, 0, total, 64, [0, 1, 0, 0], (1, 2, 3))
    check(Pinned, p_starts, p_prefix, p_score, p_cnt, 100, total - 7, 64, [0, 1, 0, 0], (4,))          # id window
    check(Pinned, p_starts, p_prefix, p_score, p_cnt, 0, total, 5, [0, 9, 0, 0], (5, 6))               # other kernel
    check(Pinned, p_starts, p_prefix, p_score, p_cnt, 50, 50, 64, [0, 1, 0, 0], (7,))                  # empty window
    # the structured form of the tables beside them (only it is copied), pinned
    class PinnedStr(Pinned):
        rowinfo_str, str_thresh, str_n = pinned(at.structured()[0]), at.structured()[1], at.structured()[2]
    check(PinnedStr, p_starts, p_prefix, p_score, p_cnt, 0, total, 64, [0, 1, 0, 0], (1, 2))
    check(PinnedStr, p_starts, p_prefix, p_score, p_cnt, 100, total - 7, 5, [0, 9, 0, 0], (4,))
    # pageable buffers and fresh result buffers (an AliasTables object: its structured form is found by itself)
    check(at, starts.copy(), prefix.copy(), np.zeros(len(starts), np.int64), np.zeros(8, np.uint64), 0, total, 64,
          [0, 1, 0, 0], (8, 9))
    other_score = pinned(np.zeros(len(starts), np.int64))
    check(Pinned, p_starts, p_prefix, other_score, p_cnt, 0, total, 64, [0, 1, 0, 0], (10,))


def test_factored_row_descriptors_equal_the_table(engine, c1_csr, k1_env, monkeypatch):
    """TAB_FAC: row descriptors derived from the (direction, cell) structure instead of loaded.  Forced
    on the default mesh (where the table would fit shared memory) and compared with the numpy port and
    with the table-driven kernels: score (fast and general kernel), histogram, time-resolved counts."""
    from nrw import mesh
    from oracle import philox_port
    _, nc1, _ = _classes(c1_csr, mesh.source_default)
    nc3 = (np.arange(c1_csr.n_rows) % 4).astype(np.uint8)
    starts = np.array([435, 0, 900, 14, 14, 899, 450, 16, 464])
    counts = np.array([3000, 0, 17, 1, 333, 64, 5, 2500, 1200])
    for nc, cv in ((nc1, [0, 1, 0, 0]), (nc3, [5, 0, 255, 1])):
        tabs = engine.upload(c1_csr, node_class=nc)
        assert tabs.factored and tabs.struct.fac_period == 30 and tabs.struct.fac_cells == 30
        for variant in ("fast", "fast_nodet", "general", "str", "str_nodet"):
            k1_env(variant)
            monkeypatch.setenv("NRW_TAB", "fac")
            for nt in (64, 3, 2):
                res = engine.walk(tabs, starts, counts, nt, 5, score=True, class_value=cv)
                assert "TAB_FAC" in res.kernel, res.kernel
                port = philox_port.walk_port(c1_csr, starts, counts, nt, 5, alias=tabs.alias_host, class_value=cv)
                _cmp(res, port, hist=False)
            monkeypatch.delenv("NRW_TAB")
    k1_env("fast")
    monkeypatch.setenv("NRW_TAB", "fac")
    tabs = engine.upload(c1_csr, node_class=nc1)
    res = engine.walk(tabs, starts, counts, 50, 8, score=True, node_hist=True, class_value=[0, 1, 0, 0])
    assert "TAB_FAC" in res.kernel
    port = philox_port.walk_port(c1_csr, starts, counts, 50, 8, alias=tabs.alias_host, class_value=[0, 1, 0, 0])
    _cmp(res, port, hist=True)
    res = engine.walk(tabs, [435, 14], [5000, 700], 40, 9, count_tn=True, t_cap=40, count_per_start=True)
    assert "TAB_FAC" in res.kernel
    port = philox_port.walk_port(c1_csr, [435, 14], [5000, 700], 40, 9, alias=tabs.alias_host, t_cap=40,
                                 count_per_start=True)
    _cmp(res, port, score=False, hist=False, count=True)


def test_refined_mesh_factored_equals_l2_resident_descriptors(engine, monkeypatch):
    """C4 (158 760 nodes): the planner takes TAB_FAC; scores and counters equal the TAB_HYB walk."""
    from nrw import mesh, tables
    csr = mesh.build_fluence_csr(0.001, 1.0 / 63.0, 200, 0.15, 1)
    src = mesh.source_values(csr.meta["mesh"]["meshpoints"], mesh.source_default)
    nc, _ = tables.classify_source(src, csr.n_rows)
    tabs = engine.upload(csr, node_class=nc)
    assert tabs.factored and tabs.struct.fac_period == 126 and tabs.struct.fac_cells == 1260
    starts = np.arange(0, csr.n_live, 7)
    out = {}
    for mode in ("fac", "hyb"):
        monkeypatch.setenv("NRW_TAB", mode)
        res = engine.walk(tabs, starts, 40, 30000, 3, score=True, class_value=[0, 1, 0, 0])
        assert ("TAB_FAC" if mode == "fac" else "TAB_HYB") in res.kernel, res.kernel
        out[mode] = (res.score.cpu().numpy(), res.counters.cpu().numpy()[:5])
    monkeypatch.delenv("NRW_TAB")
    res = engine.walk(tabs, starts, 40, 30000, 3, score=True, class_value=[0, 1, 0, 0])
    assert "TAB_FAC" in res.kernel                                   # the planner's own choice
    assert np.array_equal(out["fac"][0], out["hyb"][0]) and np.array_equal(out["fac"][1], out["hyb"][1])
    assert np.array_equal(res.score.cpu().numpy(), out["hyb"][0])


def test_very_long_time_horizons_stay_with_the_general_kernel(engine, c1_csr):
    """The trimmed kernel keeps 32-bit running sums per lane; the planner only picks it when a 32-id
    range cannot wrap them (32 * nt * max weight < 2^31)."""
    from nrw import mesh
    from oracle import philox_port
    _, nc, _ = _classes(c1_csr, mesh.source_default)
    tabs = engine.upload(c1_csr, node_class=nc)
    starts, counts = [435, 14], [500, 300]
    res = engine.walk(tabs, starts, counts, 1 << 23, 3, score=True, class_value=[0, 255, 0, 0])
    assert "walk_philox<" in res.kernel, res.kernel
    port = philox_port.walk_port(c1_csr, starts, counts, 1 << 23, 3, alias=tabs.alias_host, class_value=[0, 255, 0, 0])
    _cmp(res, port, hist=False)
    res = engine.walk(tabs, starts, counts, 1 << 18, 3, score=True, class_value=[0, 255, 0, 0])
    assert "walk_score" in res.kernel, res.kernel
