"""GPU parity of the SURVEY 8(f) "next" surfaces: loihirw.rad_test -> probes -> rw_analysis,
densityrw.DensityRW / run_miniapp on arbitrary sparse graphs, the run_loihi / run_fugu adapters of
the miniapps, and the refined (C4) mesh.  Integer results are compared bit for bit with run_pc in
replay mode (itself pinned to the reference's golden tempdata in test_gpu_replay.py), with the
oracle on general graphs and with the numpy port of the device sampler; statistical checks state
their tolerance inline."""
import numpy as np
import pytest

from conftest import C1

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c1_csv(tmp_path_factory, c1_csr):
    from nrw import mesh
    p = str(tmp_path_factory.mktemp("mat") / "curmat.csv")
    mesh.save_matrix_csv(p, c1_csr)
    return p


def _graph_csv(tmp_path, seed=0, n=40):
    """Random sparse graph: ragged rows, one probability-1 self-loop sink, one row without edges in
    the middle of the node range, one single-edge row."""
    rng = np.random.default_rng(seed)
    q = np.zeros((n, n))
    for i in range(n):
        k = int(rng.integers(1, 9))
        cols = rng.choice(n, size=k, replace=False)
        q[i, cols] = rng.dirichlet(np.ones(k))
    q[7] = 0; q[7, 7] = 1.0            # sink with self loop
    q[19] = 0                          # no outgoing edges at all
    q[23] = 0; q[23, 5] = 1.0          # deterministic hop
    p = str(tmp_path / "graph.csv")
    np.savetxt(p, q, delimiter=",")
    return p, q


# ---------------------------------------------------------------- loihirw
def test_rad_test_split_probes_decode_to_run_pc_tempdata(c1_csv):
    import fluence_mini_app as app
    from loihirw.rw_analysis import process_outputs_split
    from loihirw.rw_model import rad_test
    nt, M, start = 64, 100, 435
    np.random.seed(21)
    td = app.run_pc(matrix_filename=c1_csv, neuraltime=nt, initial_walkers=[[start, M]], meshSize=900,
                    runtime_args={"rng": "replay"})
    pos = np.random.uniform()
    np.random.seed(21)
    tprobe, eprobe, track_probes, core_neuroncount, max_coreid = rad_test(
        c1_csv, [start], nt, num_walkers=M, core_density=1.1, input_times=[1], track=6, size=(900, 1),
        rng="replay")
    assert np.random.uniform() == pos                    # same stream position as run_pc
    assert tprobe == [] and eprobe == [] and max_coreid == 83 and core_neuroncount.sum() == 901
    count, num_bins = process_outputs_split(track_probes, 0, 1, 900)
    assert num_bins == nt and count.shape == (900, nt + 1)
    assert np.array_equal(count[:, :nt], td[:900])
    # the sink is mesh point 900 of the probe list; its raster decodes to the exits per step
    csink, _ = process_outputs_split(track_probes, 900, 1, 1)
    assert np.array_equal(csink[0, :nt], td[900])
    # 'First empty number' = first tick of the first step at which no walker is left on the mesh
    period = int(np.nonzero(track_probes[0][0][1].data[:][0])[0][0]) + 1
    first_dead = int(np.nonzero(td[:900].sum(axis=0) == 0)[0][0])
    assert int(np.nonzero(track_probes[0][1][1].data[:][2])[0][0]) == first_dead * period


def test_rad_test_tracks_and_multi_location(c1_csv):
    from loihirw.rw_analysis import count_steps, process_outputs, process_outputs_ends
    from loihirw.rw_model import rad_test
    nt, M = 48, 60
    np.random.seed(5)
    flat = rad_test(c1_csv, [435, 14], nt, num_walkers=M, track=2, size=(900, 1), rng="replay")[2]
    np.random.seed(5)
    ends = rad_test(c1_csv, [435, 14], nt, num_walkers=M, track=1, size=(900, 1), rng="replay")[2]
    np.random.seed(5)
    assert rad_test(c1_csv, [435, 14], nt, num_walkers=M, track=0, size=(900, 1), rng="replay")[2] == []
    count, nb = process_outputs(flat, 3, 5, 901)
    assert nb == nt and count[435, 0] == M and count[14, 0] == M and count[:, 0].sum() == 2 * M
    total, steps = process_outputs_ends(ends, 4, 5, 901, 1)
    assert steps[0, 0] == nt and count_steps(ends, 4, 5, 901, 1)[0, 0] == nt
    assert np.array_equal(total[:, 0], count[:, :nt].sum(axis=1))
    # walker conservation: everyone who left a live node is somewhere the next step
    alive = count[:900, :nt].sum(axis=0)
    assert np.array_equal(alive[1:] + np.cumsum(count[900, :nt])[1:], np.full(nt - 1, 2.0 * M))
    # two input times = twice the walkers (define_inputs, rw_mesh.py:24-26)
    np.random.seed(5)
    tp = rad_test(c1_csv, [435], nt, num_walkers=M, input_times=[1, 7], track=2, size=(900, 1), rng="philox")[2]
    c2, _ = process_outputs(tp, 3, 5, 901)
    assert c2[435, 0] == 2 * M


def test_run_loihi_and_run_fugu_adapters_equal_run_pc(c1_csv):
    import fluence_mini_app as app
    nt, M = 64, 100
    for start in (0, 16, 435, 899):
        np.random.seed(8)
        td = app.run_pc(matrix_filename=c1_csv, neuraltime=nt, initial_walkers=[[start, M]], meshSize=900,
                        runtime_args={"rng": "replay"})
        np.random.seed(8)
        cl = app.run_functions["loihi"](matrix_filename=c1_csv, neuraltime=nt, initial_walkers=[[start, M]],
                                        meshSize=900, runtime_args={"rng": "replay"})
        np.random.seed(8)
        cf = app.run_functions["fugu"](matrix_filename=c1_csv, neuraltime=nt, initial_walkers=[[start, M]],
                                       meshSize=900, runtime_args={"rng": "replay", "fugu_backend": "snn",
                                                                   "use_sinks": False})
        assert cl.shape == (900, nt + 1) and np.array_equal(cl[:, :nt], td[:900])
        last = int(np.nonzero(td[:900].sum(axis=0))[0][-1])     # run_fugu drops the last occupied step
        assert cf.shape == (900, last) and np.array_equal(cf, td[:900, :last])


def test_fluence_run_through_the_loihi_dispatch_entry_equals_pc(tmp_path):
    """run(args) with run_mode 'loihi' takes the reference's generic per-start-node loop
    (fluence_mini_app.py:397-438) over rad_test + process_outputs_split; same seed => same rwSol."""
    import fluence_mini_app as app
    a = vars(app.build_parser().parse_args(["-r", "pc", "-M", "20", "-nt", "64", "--rng", "replay"]))
    a.update(matrix_filename=str(tmp_path / "m.csv"), results_filename=str(tmp_path / "r.csv"))
    np.random.seed(2)
    pc = app.run(dict(a))
    a["run_mode"] = "loihi"
    np.random.seed(2)
    lo = app.run(dict(a))
    assert np.array_equal(pc, lo)


# ---------------------------------------------------------------- densityrw
def test_density_rw_ring_graph_multi_source_matches_device_port(engine):
    from densityrw.DensityRW import DensityRW, build_test_graph
    from nrw import dropin, tables
    from oracle import philox_port
    walk = DensityRW(50, build_test_graph(25), init_walkers={(10,): 300, (13,): 500})
    np.random.seed(77)
    seed = dropin.philox_seed_from_global_stream()
    np.random.seed(77)
    counts = walk.run(rng="philox")
    assert counts.shape == (25, 50) and counts[10, 0] == 300 and counts[13, 0] == 500
    assert np.array_equal(counts.sum(axis=0), np.full(50, 800.0))        # a ring loses nobody
    at = tables.build_alias(walk.csr)
    port = philox_port.walk_port(walk.csr, [10, 13], [300, 500], 50, seed, alias=at, t_cap=50)
    ref = port["count_tn"][0].T.astype(np.float64)
    ref[10, 0], ref[13, 0] = 300, 500
    assert np.array_equal(counts, ref)
    # one step on a ring moves every walker to a neighbour of its start
    assert counts[[9, 11], 1].sum() == 300 and counts[[12, 14], 1].sum() == 500


def test_density_rw_general_graph_replay_matches_oracle(tmp_path):
    from densityrw.DensityRW import DensityRW, load_transitions
    from nrw import tables
    from oracle import restated
    path, q = _graph_csv(tmp_path)
    tr = load_transitions(path, remove_sink_connections=True, verbose=0)
    assert tr[(7,)].neighbors == [] and tr[(19,)].neighbors == [] and tr[(23,)].neighbors == [((5,), 1.0)]
    nt = 30
    walk = DensityRW(nt, tr, init_walkers={(3,): 400})
    np.random.seed(13)
    counts = walk.run(rng="replay")
    pos = np.random.uniform()
    csr = walk.csr
    cum = tables.row_cum(csr)
    np.random.seed(13)
    ref, used = restated.run_pc_counts(csr.indptr, csr.cols.astype(np.int64), cum, nt, [[3, 400]], csr.n_rows)
    assert np.random.uniform() == pos
    assert np.array_equal(counts, ref[:csr.n_rows])
    # walkers that reach the sink (7) or the edgeless node (19) are counted once and stop
    alive_nodes = np.ones(40, dtype=bool); alive_nodes[[7, 19]] = False
    absorbed = np.cumsum(counts[7] + counts[19])
    assert np.array_equal(counts[alive_nodes].sum(axis=0)[1:] + absorbed[1:], np.full(nt - 1, 400.0))
    # two sources in replay mode = the two single-source walks, in list order, on one stream
    np.random.seed(13)
    both = DensityRW(nt, tr, init_walkers={(3,): 400, (30,): 250}).run(rng="replay")
    np.random.seed(13)
    a, _ = restated.run_pc_counts(csr.indptr, csr.cols.astype(np.int64), cum, nt, [[3, 400]], csr.n_rows)
    b, _ = restated.run_pc_counts(csr.indptr, csr.cols.astype(np.int64), cum, nt, [[30, 250]], csr.n_rows)
    assert np.array_equal(both, (a + b)[:csr.n_rows])


def test_run_miniapp_spike_table_and_sink_tracking(tmp_path):
    from densityrw.rw_miniapp_fugu import run_miniapp
    from nrw import dropin
    path, q = _graph_csv(tmp_path, seed=4)
    nt = 25
    np.random.seed(31)
    table = run_miniapp(path, nt, [(3, 200), (11, 100)], False, 'snn', rng="philox")
    assert list(table.columns) == ["time", "neuron_number", "node"]
    td = dropin.spike_table_to_tempdata(table, 40)
    assert td[3, 0] == 200 and td[11, 0] == 100
    assert not td[7].any() and not td[19].any()                 # sinks are not reported without use_sinks
    np.random.seed(31)
    kept = run_miniapp(path, nt, [(3, 200), (11, 100)], True, 'snn', rng="philox")
    td2 = dropin.spike_table_to_tempdata(kept, 40)
    assert td2[7].any()                                         # the self-loop sink now holds its walkers
    assert (np.diff(td2[7]) >= 0).all()                         # ... forever: its count never decreases
    with pytest.raises(ValueError, match="Unsupported backend"):
        run_miniapp(path, nt, [(3, 1)], False, 'loihi')


# ---------------------------------------------------------------- BASELINE config C4 (refined mesh)
def test_refined_mesh_c4_through_the_densityrw_surface(engine, tmp_path):
    """dt=0.001, da=1/63: 158 760 nodes, nnz 1.9e7 -- the mesh the reference cannot build (201 GB
    dense).  CSR .npz -> load_transitions (lazy) -> DensityRW -> GPU walk; statistical check of the
    occupancy against the exact Markov-chain propagation (oracle/expectation.py): |z| < 5 per
    (node, step) cell with expectation >= 50 walkers."""
    from densityrw.DensityRW import DensityRW, load_transitions
    from nrw import mesh
    from oracle import expectation
    csr = mesh.build_fluence_csr(0.001, 1.0 / 63.0, 200, 0.15, 1)
    assert csr.n_rows == 158761 and csr.n_live == 158760
    path = str(tmp_path / "c4.npz")
    mesh.save_matrix_npz(path, csr)
    tr = load_transitions(path, remove_sink_connections=True, verbose=0)
    assert len(tr) == 158761 and len(tr[(80000,)].neighbors) in (1, 126)
    start, M, nt = 79443, 2000000, 40
    walk = DensityRW(nt, tr, init_walkers={(start,): M})
    np.random.seed(9)
    counts = walk.run(rng="philox")
    assert counts.shape == (158761, nt) and counts[start, 0] == M
    start_dist = np.zeros(csr.n_rows); start_dist[start] = 1.0
    occ = expectation.occupancy_expectation(csr.indptr, csr.cols, csr.vals, csr.n_rows, csr.n_live,
                                            start_dist, nt).T
    exp = occ * M
    big = (exp >= 50) & (occ < 1.0)                 # the start cell at t = 0 is certain, not random
    z = (counts[big] - exp[big]) / np.sqrt(exp[big] * (1 - occ[big]))
    assert big.sum() > 1000 and np.abs(z).max() < 5.0, np.abs(z).max()
    assert abs(z.std() - 1.0) < 0.1
