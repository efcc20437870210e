"""GPU parity, bit-exact mode: the replay kernel (K2) + assembly (K3) + expansion (K5) against the
golden vectors produced by the reference itself (tests/golden, oracle/make_golden.py) and against
the oracle restatement.  Tolerance: none -- integer counts, stream position and float64 rwSol must be
identical to the last bit."""
import os

import numpy as np
import pytest

from conftest import C1

pytestmark = pytest.mark.gpu


def _events_to_dense(t, k, c, n_rows, nt):
    td = np.zeros((n_rows, nt))
    td[k, t] = c
    return td


@pytest.fixture(scope="module")
def c1_csv(tmp_path_factory, c1_csr):
    from nrw import mesh
    p = str(tmp_path_factory.mktemp("mat") / "curmat.csv")
    mesh.save_matrix_csv(p, c1_csr)
    return p


def test_run_pc_replay_matches_reference_tempdata(c1_csv, golden):
    import fluence_mini_app as app
    g = golden["fluence_c1_tempdata"]
    for ci, (start, seed, nt, M, steps) in enumerate(g["meta"]):
        np.random.seed(int(seed))
        td = app.run_pc(matrix_filename=c1_csv, neuraltime=int(nt), initial_walkers=[[int(start), int(M)]],
                        meshSize=900, runtime_args={"rng": "replay"})
        nxt = np.random.uniform()
        ref = _events_to_dense(g["t_%d" % ci], g["k_%d" % ci], g["c_%d" % ci], 901, int(nt))
        assert td.dtype == np.float64 and td.shape == (901, int(nt))
        assert np.array_equal(td, ref), "case %d start %d" % (ci, start)
        assert nxt == g["next_u_%d" % ci][0], "stream position, case %d" % ci
        assert int(td[:900, :int(nt) - 1].sum()) == steps


def test_run_pc_replay_chained_calls_share_the_stream(c1_csv, golden):
    """the stream is seeded once and continues across run_pc calls."""
    import fluence_mini_app as app
    g = golden["fluence_c1_tempdata"]
    np.random.seed(99)
    a = app.run_pc(matrix_filename=c1_csv, neuraltime=64, initial_walkers=[[14, 100]], meshSize=900,
                   runtime_args={"rng": "replay"})
    b = app.run_pc(matrix_filename=c1_csv, neuraltime=64, initial_walkers=[[435, 100]], meshSize=900,
                   runtime_args={"rng": "replay"})
    assert np.array_equal(a, _events_to_dense(g["chain_a_t"], g["chain_a_k"], g["chain_a_c"], 901, 64))
    assert np.array_equal(b, _events_to_dense(g["chain_b_t"], g["chain_b_k"], g["chain_b_c"], 901, 64))


def _fluence_args(tmp_path, **over):
    import fluence_mini_app as app
    a = vars(app.build_parser().parse_args(["-r", "pc"]))
    a.update(matrix_filename=str(tmp_path / "curmat.csv"), results_filename=str(tmp_path / "rwSol.csv"))
    a.update(over)
    return a


def test_run_replay_rwsol_bit_identical_c1(tmp_path, golden):
    """BASELINE config C1: full run(args), M=100, nt=64, seed 1 == the reference's rwSol.csv."""
    import fluence_mini_app as app
    g = golden["fluence_c1_rwsol"]
    args = _fluence_args(tmp_path, neural_timesteps=64, num_walks=100, rand_seed=1, rng="replay")
    np.random.seed(1)
    sol = app.run(args)
    nxt = np.random.uniform()
    assert np.array_equal(sol, g["rwSol"])
    assert nxt == g["next_u"][0]
    saved = np.genfromtxt(args["results_filename"], delimiter=",")
    assert np.array_equal(saved, g["rwSol"])
    # the matrix file is byte-compatible with what the reference's back ends parse
    from nrw import mesh
    q = np.genfromtxt(args["matrix_filename"], delimiter=",")
    gm = golden["fluence_c1_matrix"]
    assert np.array_equal(q[np.nonzero(q)], gm["vals"])


def test_run_replay_rwsol_bit_identical_c2_paper_source(tmp_path, golden):
    """BASELINE config C2 (M=1000, Nature-Electronics source) at nt=64, seed 1."""
    import fluence_mini_app as app
    g = golden["fluence_c2_rwsol"]
    args = _fluence_args(tmp_path, neural_timesteps=64, num_walks=1000, rand_seed=1, rng="replay",
                         source="paper")
    np.random.seed(1)
    sol = app.run(args)
    assert np.array_equal(sol, g["rwSol"])
    assert np.random.uniform() == g["next_u"][0]


def test_replay_with_absorption_weight_matches_oracle(tmp_path):
    """SigA > 0 only reweights the sum (fluence_mini_app.py:434, Q13): K3 must reproduce the
    oracle's sequential fp64 sum with exp(-SigA*dt*t) bit for bit."""
    import fluence_mini_app as app
    from oracle import restated
    args = _fluence_args(tmp_path, neural_timesteps=40, num_walks=30, SigA=0.7, rng="replay")
    np.random.seed(3)
    sol = app.run(args)
    np.random.seed(3)
    ref, _ = restated.fluence_run(dict(num_walks=30, timestep=0.01, anglestep=1.0 / 15.0, velocity=200,
                                       SigS=0.15, SigA=0.7, interval_length=1, neural_timesteps=40))
    assert np.array_equal(sol, ref)


def test_replay_small_stream_chunks_resume(engine, c1_csr):
    """The kernel hands back when the uploaded stream runs dry and resumes mid-start: results must
    not depend on the chunking."""
    tabs = engine.upload(c1_csr, alias=False)
    starts, M, nt = [435, 14, 0, 450], 200, 64
    np.random.seed(11)
    a = engine.replay(tabs, starts, M, nt)
    pos_a = np.random.uniform()
    np.random.seed(11)
    b = engine.replay(tabs, starts, M, nt, chunk=97)
    pos_b = np.random.uniform()
    assert a.draws_used == b.draws_used and a.n_events == b.n_events and pos_a == pos_b
    for x, y in ((a.ev_t, b.ev_t), (a.ev_node, b.ev_node), (a.ev_count, b.ev_count)):
        assert np.array_equal(x[:a.n_events].cpu().numpy(), y[:b.n_events].cpu().numpy())
    assert np.array_equal(a.ev_offsets.cpu().numpy(), b.ev_offsets.cpu().numpy())


def test_simple_flux_replay_matches_reference(tmp_path, golden):
    import simple_flux_miniapp as app
    g = golden["simple_flux"]
    for seed in (1, 2):
        a = vars(app.build_parser().parse_args(["-r", "pc", "--rng", "replay"]))
        a.update(matrix_filename=str(tmp_path / "m.csv"), results_filename=str(tmp_path / "r.csv"),
                 rand_seed=seed)
        np.random.seed(seed)
        sol = app.run(a)
        nxt = np.random.uniform()
        assert np.array_equal(sol, g["rwSol_seed%d" % seed])
        assert nxt == g["next_u_seed%d" % seed][0]
        assert np.array_equal(np.genfromtxt(a["matrix_filename"], delimiter=","), g["myQ"])
        # run_pc itself: tempdata[2, nt] for both start states on one stream
        np.random.seed(seed)
        for i in (0, 1):
            td = app.run_pc(matrix_filename=a["matrix_filename"], neuraltime=500000,
                            initial_walkers=[[i, 500]], meshSize=2, runtime_args={"rng": "replay"})
            ref = g["td_seed%d_start%d" % (seed, i)]
            assert td.shape == (2, 500000)
            assert np.array_equal(td[:, :ref.shape[1]], ref)
            assert td[:, ref.shape[1]:].sum() == 0
    # BASELINE config C3 parity shape (nt = 1002, --use_sinks is a no-op for pc, Q14)
    a = vars(app.build_parser().parse_args(["-r", "pc", "--rng", "replay", "--use_sinks", "-nt", "1002"]))
    a.update(matrix_filename=str(tmp_path / "m.csv"), results_filename=str(tmp_path / "r3.csv"), rand_seed=5)
    np.random.seed(5)
    sol = app.run(a)
    assert np.array_equal(sol, g["rwSol_c3_seed5"])
    assert np.random.uniform() == g["next_u_c3_seed5"][0]


def test_build_cum_kernel_matches_numpy_cumsum(engine, c1_csr):
    """K4 vs np.cumsum(myQ, axis=1) at the non-zero columns (fluence_mini_app.py:224)."""
    from nrw import tables
    tabs = engine.upload(c1_csr, alias=False)
    dev = tabs.cum.cpu().numpy()
    dense_cum = np.cumsum(c1_csr.to_dense(), axis=1)
    rows = np.repeat(np.arange(c1_csr.n_rows), c1_csr.row_lengths())
    assert np.array_equal(dev, dense_cum[rows, c1_csr.cols])
    assert np.array_equal(dev, tables.row_cum(c1_csr))


def test_expand_counts_matches_numpy(engine):
    import torch
    rng = np.random.default_rng(0)
    t_cap, n_rows, nt = 37, 70, 101
    cnt = rng.integers(0, 5, size=(t_cap, n_rows)).astype(np.int32)
    d = torch.from_numpy(cnt).to(engine.device)
    out = engine.expand_counts(d, t_cap, n_rows, nt, 3, 42).cpu().numpy()
    ref = np.zeros((n_rows, nt))
    ref[:, 1:t_cap] = cnt[1:].T
    ref[3, 0] = 42
    assert np.array_equal(out, ref)
