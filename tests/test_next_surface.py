"""SURVEY section 8(f) "next" rows on CPU: the NxSDK-free rw_analysis decoders, the probe writer,
densityrw.load_transitions / Transition / CSR conversion, the spike table of run_miniapp and the
run_fugu decode -- against fixtures produced by the UNMODIFIED reference
(oracle/make_golden_next.py -> tests/golden/next_surface.npz) and, when /root/reference is mounted,
against the live reference functions."""
import numpy as np
import pytest

from oracle import ref_import

have_ref = pytest.mark.skipif(not ref_import.reference_available(), reason="reference not mounted")


class P:
    def __init__(self, data):
        self.data = data


@pytest.fixture(scope="module")
def g(golden):
    return golden["next_surface"]


def _split_probes(b, spikes):
    z = np.zeros_like(b)
    return [[["Supervisor Global Spike", P(np.stack([b, b]))], ["Supervisor Fire Spike", P(np.stack([b, z, z]))]],
            [["Mesh Node: %d" % j, P(np.stack([b, spikes[j], z]))] for j in range(len(spikes))]]


def _flat_probes(b, spikes):
    z = np.zeros_like(b)
    sup = [P(np.stack([b, b]))]
    flat = [sup, sup, sup]
    for j in range(len(spikes)):
        flat.append([P(np.stack([b, spikes[j], z]))])
        flat.extend([sup] * 4)
    return flat


# ---------------------------------------------------------------- rw_analysis decoders vs reference
@pytest.mark.parametrize("case", [0, 1, 2])
def test_process_outputs_split_and_flat_match_reference(g, case):
    from loihirw import rw_analysis as A
    b, spikes = g["an%d_boundary" % case], g["an%d_spikes" % case]
    c6, nb6 = A.process_outputs_split(_split_probes(b, spikes), 0, 1, len(spikes))
    assert np.array_equal(c6, g["an%d_split_count" % case]) and nb6 == g["an%d_split_bins" % case][0]
    c2, nb2 = A.process_outputs(_flat_probes(b, spikes), 3, 5, len(spikes))
    assert np.array_equal(c2, g["an%d_flat_count" % case]) and nb2 == g["an%d_flat_bins" % case][0]
    # the decoded counts are the walker counts the rasters were drawn from
    assert np.array_equal(c6[:, :nb6], g["an%d_counts" % case])


def test_process_outputs_ends_and_count_steps_match_reference(g):
    from loihirw import probes, rw_analysis as A
    tp = probes.ends_track_probes(g["ends_totals"], int(g["ends_n_steps"][0]))
    count, num_steps = A.process_outputs_ends(tp, 4, 5, len(g["ends_totals"]), 1)
    assert np.array_equal(count, g["ends_count"]) and np.array_equal(num_steps, g["ends_num_steps"])
    assert np.array_equal(A.count_steps(tp, 4, 5, len(g["ends_totals"]), 1), g["ends_count_steps"])
    assert np.array_equal(count[:, 0], g["ends_totals"]) and num_steps[0, 0] == g["ends_n_steps"][0]


# ---------------------------------------------------------------- probe writer round trips
def _random_counts(seed, n_nodes=9, n_steps=13, peak=17):
    rng = np.random.default_rng(seed)
    c = rng.integers(0, peak + 1, size=(n_nodes, n_steps))
    c[rng.random(c.shape) < 0.5] = 0
    c[:, -3:] = 0                                   # the walk has died out before the end
    return c


@pytest.mark.parametrize("mode", ["unary", "count"])
def test_probe_writer_round_trips_through_our_decoders(mode):
    from loihirw import probes, rw_analysis as A
    c = _random_counts(3)
    tp, lay = probes.split_track_probes(c, live_rows=8, spike_mode=mode)
    count, nb = A.process_outputs_split(tp, 0, 1, c.shape[0])
    assert nb == c.shape[1] and count.shape == (c.shape[0], nb + 1)
    assert np.array_equal(count[:, :nb], c) and not count[:, nb].any()
    flat, _ = probes.flat_track_probes(c, spike_mode=mode)
    count2, nb2 = A.process_outputs(flat, 3, 5, c.shape[0])
    assert nb2 == nb and np.array_equal(count2[:, :nb], c)
    # 'First empty number' (fluence_mini_app.py:169-172): first tick of the first step without walkers
    first_dead = np.nonzero(c[:8].sum(axis=0) == 0)[0][0]
    nz = np.nonzero(tp[0][1][1].data[:][2])[0]
    assert nz[0] == first_dead * lay.period
    if mode == "unary":                             # one spike per walker, never two on a tick
        assert max(np.asarray(tp[1][j][1].data[:][1]).max() for j in range(c.shape[0])) == 1


@have_ref
@pytest.mark.parametrize("mode", ["unary", "count"])
def test_probe_writer_is_read_back_by_the_unmodified_reference_decoders(mode):
    from loihirw import probes
    A = ref_import.rw_analysis()
    c = _random_counts(5)
    tp, _ = probes.split_track_probes(c, spike_mode=mode)
    count, nb = A.process_outputs_split(tp, 0, 1, c.shape[0])
    assert nb == c.shape[1] and np.array_equal(count[:, :nb], c)
    flat, _ = probes.flat_track_probes(c, spike_mode=mode)
    count2, nb2 = A.process_outputs(flat, 3, 5, c.shape[0])
    assert np.array_equal(count2[:, :nb2], c)
    ends = probes.ends_track_probes(c.sum(axis=1), c.shape[1])
    ce, ns = A.process_outputs_ends(ends, 4, 5, c.shape[0], 1)
    assert np.array_equal(ce[:, 0], c.sum(axis=1)) and ns[0, 0] == c.shape[1]


def test_core_assignment_follows_the_reference_placement_arithmetic():
    from loihirw.rw_mesh import core_assignment
    core_id, max_coreid, chips = core_assignment(901, core_density=1.0, core_start=0)   # 10 per core
    assert core_id[0] == 1 and core_id[9] == 1 and core_id[10] == 2 and core_id[900] == 91
    assert max_coreid == 92 and chips == 1
    _, max_coreid, chips = core_assignment(1270, core_density=1.0)     # rolls over core 128
    assert max_coreid == 128 and chips == 2
    core_id, _, _ = core_assignment(901, core_density=1.1)   # run_loihi's default: 1.1 * 10 == 11.0
    assert core_id[10] == 1 and core_id[11] == 2 and core_id[900] == 82
    core_id, max_coreid, chips = core_assignment(901, core_density=0.77)   # 7.7 per core: `==` never fires
    assert (core_id == 1).all() and max_coreid == 2 and chips == 1


# ---------------------------------------------------------------- densityrw surface
def _flatten(tr):
    src, dst, prob = [], [], []
    for k in sorted(tr.keys()):
        for dest, p in tr[k].neighbors:
            src.append(k[0]); dst.append(dest[0]); prob.append(p)
    return np.array(src), np.array(dst), np.array(prob)


@pytest.mark.parametrize("remove", [True, False])
def test_load_transitions_matches_reference(g, tmp_path, remove):
    from densityrw.DensityRW import load_transitions, transitions_to_csr
    from nrw import mesh
    path = str(tmp_path / "graph.csv")
    np.savetxt(path, g["lt_matrix"], delimiter=",")
    tr = load_transitions(path, remove_sink_connections=remove, verbose=0)
    tag = "lt_rm%d" % int(remove)
    src, dst, prob = _flatten(tr)
    assert len(tr) == g[tag + "_n"][0]
    assert np.array_equal(src, g[tag + "_src"]) and np.array_equal(dst, g[tag + "_dst"])
    assert np.array_equal(prob, g[tag + "_prob"])
    assert all(tr[k].location == k for k in tr)
    # CSR conversion keeps every edge bit for bit; the lazy CSR-backed view is the same mapping
    csr = transitions_to_csr(tr)
    assert csr.n_rows == 7 and csr.nnz == len(src)
    assert np.array_equal(csr.cols, dst) and np.array_equal(csr.vals, prob)
    npz = str(tmp_path / "graph.npz")
    mesh.save_matrix_npz(npz, mesh.csr_from_dense(g["lt_matrix"], n_live=7))
    lazy = load_transitions(npz, remove_sink_connections=remove, verbose=0)
    s2, d2, p2 = _flatten(lazy)
    assert len(lazy) == 7 and (2,) in lazy and (7,) not in lazy
    assert np.array_equal(s2, src) and np.array_equal(d2, dst) and np.array_equal(p2, prob)
    csr2 = transitions_to_csr(lazy)
    assert np.array_equal(csr2.indptr, csr.indptr) and np.array_equal(csr2.cols, csr.cols)
    assert np.array_equal(csr2.vals, csr.vals)


def test_load_transitions_rejects_unknown_files():
    from densityrw.DensityRW import load_transitions
    with pytest.raises(ValueError, match="Matrix file incompatible"):
        load_transitions("graph.txt")


@have_ref
def test_ring_test_graph_matches_reference():
    from densityrw.DensityRW import build_test_graph
    ref = ref_import.density_rw().build_test_graph(11)
    mine = build_test_graph(11)
    assert sorted(ref) == sorted(mine)
    for k in ref:
        assert ref[k].location == mine[k].location and ref[k].neighbors == mine[k].neighbors


def test_ring_graph_csr_rows_are_stochastic():
    from densityrw.DensityRW import build_test_graph, transitions_to_csr
    csr = transitions_to_csr(build_test_graph(20))
    assert csr.n_rows == 20 and csr.nnz == 40 and csr.n_live == 20
    assert np.array_equal(np.add.reduceat(csr.vals, csr.indptr[:-1]), np.ones(20))
    assert csr.cols[csr.indptr[0]:csr.indptr[1]].tolist() == [1, 19]


def test_spike_table_decodes_like_the_reference_run_fugu(g):
    """counts -> run_miniapp's spike table -> the run_fugu post-processing; the golden tempdata is
    what the reference's own run_fugu (fluence_mini_app.py:66-107) returned for that table."""
    from densityrw.rw_miniapp_fugu import counts_to_spike_table
    from nrw import dropin
    counts = g["fugu_counts"]
    meshSize = counts.shape[0] - 1
    table = counts_to_spike_table(counts, include_rows=np.arange(meshSize + 1) < meshSize)
    assert list(table.columns) == ["time", "neuron_number", "node"]
    assert len(table) == counts[:meshSize].sum() and table["node"].iloc[0] == "(2,)"
    td = dropin.spike_table_to_tempdata(table, meshSize)
    assert td.dtype == np.float64 and np.array_equal(td, g["fugu_tempdata"])
    # the reference sizes tempdata by the index of the last occupied step, which drops that step
    last = np.nonzero(counts[:meshSize].sum(axis=0))[0][-1]
    assert td.shape == (meshSize, last) and np.array_equal(td, counts[:meshSize, :last])


# ---------------------------------------------------------------- rw_core probability tree
def test_get_prob_circuit_values_matches_reference(g):
    from loihirw import rw_core
    n = int(g["pc_n"][0])
    assert n >= 8
    for i in range(n):
        prob = list(g["pc%d_prob" % i])
        conn, relay, weight = rw_core.get_probCircuitValues(prob)
        assert np.array_equal(conn, g["pc%d_conn" % i]), i
        assert np.array_equal(relay, g["pc%d_relay" % i]), i
        assert np.array_equal(weight, g["pc%d_weight" % i]), i
    with pytest.raises(IndexError):                      # the reference's own failure on this fan-out
        rw_core.get_probCircuitValues([0.0, 0.6, 0.4])


def test_probability_tree_realises_the_row_distribution():
    from loihirw import rw_core
    rng = np.random.default_rng(2)
    for n in (2, 3, 5, 8, 30):
        p = rng.dirichlet(np.ones(n))
        d = rw_core.tree_distribution(list(p))
        assert np.allclose(d[:n], p, atol=1e-15) and not d[n:].any()
        dq = rw_core.tree_distribution(list(p), quantise=True)       # 8-bit weights: |dp| <= depth/128
        assert abs(dq.sum() - 1) < 1e-12 and np.abs(dq[:n] - p).max() <= np.ceil(np.log2(n)) / 128
