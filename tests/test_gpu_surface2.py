"""Round-2 surfaces on the GPU: the likely-path tally of the time-resolved counts (nrw_walk.count_path), the
Loihi-faithful tree sampler (loihirw/rw_core.py:11-156 as a device table), and the absorbed /
exit counts that --use_sinks returns next to rwSol (simple_flux_miniapp.py:189-201)."""
import numpy as np
import pytest

from conftest import C1

pytestmark = pytest.mark.gpu


def _counts(res):
    return res.count_tn.cpu().numpy().astype(np.int64)


def _ref_counts(port):
    ref = port["count_tn"].copy()
    ref[:, 0, :] = 0
    return ref


@pytest.mark.parametrize("per_start", [False, True])
def test_likely_path_tally_equals_the_plain_tally(engine, c1_csr, per_start, monkeypatch):
    """Global time-resolved counts with and without the leave-time histogram, against the numpy port: ragged starts
    (an empty one, the sink itself, a repeated node), a window shorter than the walk (t_cap < nt: overflow), and
    a window longer than any walk.  With the path the trimmed count walk runs (walk_score.cuh, CNT), under
    NRW_K1=general the general kernel's path tally, without a path the general kernel's hot-bin cache."""
    from oracle import philox_port
    tabs = engine.upload(c1_csr)
    starts = np.array([435, 0, 900, 14, 14, 899, 450, 464])
    counts = np.array([40000, 0, 17, 3000, 333, 64, 5000, 1200])
    for nt, t_cap in ((64, 64), (64, 40), (300, 300)):
        port = philox_port.walk_port(c1_csr, starts, counts, nt, 9, alias=tabs.alias_host, t_cap=t_cap,
                                     count_per_start=per_start)
        for path, general, want in ((True, False, "walk_count<"), (True, True, "32>"), (False, False, "32>")):
            if general:
                monkeypatch.setenv("NRW_K1", "general")
            else:
                monkeypatch.delenv("NRW_K1", raising=False)
            res = engine.walk(tabs, starts, counts, nt, 9, count_tn=True, t_cap=t_cap, count_per_start=per_start,
                              count_path=path)
            assert want in res.kernel, res.kernel           # "32>": TL_CNT_G, the bins do not fit shared memory
            c = res.counters_dict()
            for k in ("steps", "exits", "survivors", "walkers"):
                assert c[k] == port["counters"][k], (k, path, general)
            if t_cap < nt:
                assert c["t_overflow"] > 0
            assert np.array_equal(_counts(res), _ref_counts(port)), (nt, t_cap, path, general)
        monkeypatch.delenv("NRW_K1", raising=False)


def test_likely_path_tally_on_a_general_graph(engine):
    """Random sparse graph with absorbing rows, edgeless rows and ties between successors."""
    from nrw import mesh
    from oracle import philox_port
    rng = np.random.default_rng(5)
    n = 70000                                   # 70 000 x 40 bins > shared memory: global tally
    q_rows, q_cols, q_vals = [], [], []
    indptr = [0]
    for r in range(n):
        k = 0 if r % 97 == 0 else int(rng.integers(1, 5))
        cols = np.sort(rng.choice(n, size=k, replace=False))
        vals = rng.integers(1, 4, size=k).astype(np.float64)
        q_cols.append(cols)
        q_vals.append(vals / vals.sum() if k else vals)
        indptr.append(indptr[-1] + k)
    csr = mesh.TransitionCSR(n_rows=n, n_live=n, indptr=np.array(indptr, np.int64),
                             cols=np.concatenate(q_cols).astype(np.int32), vals=np.concatenate(q_vals))
    tabs = engine.upload(csr)
    starts, counts = [1, 2, 97, 4000], [30000, 500, 10, 7000]
    port = philox_port.walk_port(csr, starts, counts, 40, 3, alias=tabs.alias_host, t_cap=40)
    res = engine.walk(tabs, starts, counts, 40, 3, count_tn=True, t_cap=40, count_path=True)
    assert np.array_equal(_counts(res), _ref_counts(port))


def test_tree_sampler_equals_the_port_and_the_quantised_circuit(engine, c1_csr):
    """sampler='tree': the walk draws through rw_core's binary tree of 8-bit Bernoulli weights.  Integer tallies equal
    the numpy port on the same streams; the one-step distribution equals loihirw.rw_core.tree_distribution(
    quantise=True) within the Monte-Carlo error -- and differs from the exact row by the quantisation."""
    from loihirw import rw_core
    from nrw import mesh, tables
    from oracle import philox_port
    src = mesh.source_values(c1_csr.meta["mesh"]["meshpoints"], mesh.source_default)
    nc, _ = tables.classify_source(src, c1_csr.n_rows)
    tabs = engine.upload(c1_csr, node_class=nc)
    starts = np.arange(0, 900, 13)
    res = engine.walk(tabs, starts, 300, 64, 4, sampler="tree", score=True, class_value=[0, 1, 0, 0], node_hist=True)
    assert "TAB_TREE" in res.kernel
    port = philox_port.walk_port(c1_csr, starts, 300, 64, 4, sampler="tree", tree=tabs.tree_host, node_class=nc,
                                 class_value=[0, 1, 0, 0])
    assert np.array_equal(res.score.cpu().numpy(), port["score"])
    assert np.array_equal(res.node_hist.cpu().numpy(), port["node_hist"])
    assert res.counters_dict()["steps"] == port["counters"]["steps"]
    # one step from a stay row: empirical distribution vs the quantised circuit
    r, M = 435, 4_000_000
    one = engine.walk(tabs, [r], [M], 2, 8, sampler="tree", count_tn=True, t_cap=2)
    lo, hi = c1_csr.indptr[r], c1_csr.indptr[r + 1]
    emp = _counts(one)[0, 1][c1_csr.cols[lo:hi]] / M
    want = rw_core.tree_distribution(list(c1_csr.vals[lo:hi]), quantise=True)[:hi - lo]
    assert abs(emp.sum() - 1.0) < 1e-12
    assert (np.abs(emp - want) < 5.0 * np.sqrt(np.maximum(want, 1e-9) / M) + 1e-9).all()
    assert np.abs(want - c1_csr.vals[lo:hi]).max() > 1e-3          # the 8-bit weights are NOT the exact row


def test_use_sinks_returns_absorbed_counts_next_to_rwsol(tmp_path):
    """simple_flux --use_sinks (pc): rwSol unchanged; sinks[i, t] = walkers absorbed at step t; alive + absorbed is
    conserved step by step.  Both RNG modes."""
    import simple_flux_miniapp as app
    for rng, M in (("philox", 300000), ("replay", 500)):
        a = vars(app.build_parser().parse_args(["-r", "pc", "--rng", rng, "-M", str(M)]))
        a.update(matrix_filename=str(tmp_path / "m.csv"), results_filename=str(tmp_path / "r.csv"))
        np.random.seed(4)
        plain = app.run(a)
        a["use_sinks"] = True
        np.random.seed(4)
        sol = app.run(a)
        assert np.array_equal(sol, plain)
        sinks = app.last_sink_counts
        assert sinks.shape == sol.shape and sinks[:, 0].sum() == 0
        saved = np.genfromtxt(str(tmp_path / "r_sinks.csv"), delimiter=",")
        assert np.array_equal(saved, sinks)
        # start state 0 with ip = 5, in = 3: alive[t] = n+[t] + n-[t]; rwSol alone cannot give it, so re-walk counts
        csr = app.mesh.csr_from_dense(app.mesh.build_simple_flux_dense(0.005, 8.0, 2.0), n_live=2)
        np.random.seed(4)
        td, _, _ = app._counts_pc_fused(csr, M, a["neural_timesteps"], sol.shape[1], rng, "alias")
        alive = td[:, 0, :] + td[:, 1, :]
        assert np.array_equal(alive[:, 1:] + td[:, 2, 1:], alive[:, :-1])
        assert np.array_equal(td[:, 2, :sinks.shape[1]], sinks)
