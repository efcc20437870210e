"""CPU tests of the product's host logic (no GPU, no compute calls into libnrw):
mesh builder bit-exactness, matrix files, sampling tables, index maps, CLI surface, sharding."""
import os
import re

import numpy as np
import pytest

from conftest import C1, ROOT
from nrw import dist, mesh, tables
from oracle import philox_port, ref_import, restated as R


# ---------------------------------------------------------------- builder
def test_pairwise_rowsum_emulation_matches_numpy():
    rng = np.random.default_rng(0)
    for n in (1, 5, 8, 9, 100, 128, 129, 901, 1000, 10001, 158761):
        for _ in range(12):
            W = int(rng.integers(1, min(n, 130) + 1))
            c0 = int(rng.integers(0, n - W + 1))
            block = rng.random((4, W))
            dense = np.zeros((4, n))
            dense[:, c0:c0 + W] = block
            assert np.array_equal(dense.sum(axis=1), mesh._pairwise_rowsum_window(n, c0, block))


def test_csr_builder_is_bit_identical_to_reference_matrix(c1_csr, golden):
    g = golden["fluence_c1_matrix"]
    assert c1_csr.n_rows == 901 and c1_csr.n_live == 900 and c1_csr.nnz == 13950
    assert np.array_equal(c1_csr.indptr, g["indptr"]) and np.array_equal(c1_csr.cols, g["cols"])
    assert np.array_equal(c1_csr.vals.view(np.uint64), g["vals"].view(np.uint64))


@pytest.mark.parametrize("dt,da", [(0.01, 0.02), (0.0025, 1.0 / 16.0), (0.005, 1.0 / 15.0)])
def test_csr_builder_matches_oracle_dense_on_other_meshes(dt, da):
    csr = mesh.build_fluence_csr(dt, da, 200, 0.15, 1)
    q, m = R.fluence_matrix_dense(dt, da, 200, 0.15, 1)
    assert np.array_equal(csr.to_dense(), q)
    assert csr.n_live == m["meshSize"]


def test_refined_mesh_shape():
    """BASELINE config C4: 1260 x 126 nodes -- the reference cannot even allocate this matrix."""
    csr = mesh.build_fluence_csr(0.001, 1.0 / 63.0, 200, 0.15, 1)
    m = csr.meta["mesh"]
    assert (m["NS"] - 1, m["ND"] - 1, csr.n_live) == (1260, 126, 158760)
    assert np.all(np.abs(np.add.reduceat(csr.vals, csr.indptr[:-2]) - 1) < 1e-12)


def test_bad_physics_raises_like_the_reference():
    with pytest.raises(ValueError, match="too large"):
        mesh.build_fluence_csr(0.05, 1.0 / 15.0, 200, 0.15, 1)       # fluence_mini_app.py:285-291
    with pytest.raises(ValueError, match="too large"):
        mesh.build_simple_flux_dense(0.05, 8.0, 2.0)                 # simple_flux_miniapp.py:232-238
    with pytest.raises(IndexError):
        mesh.build_fluence_csr(0.01, 0.07, 200, 0.15, 1)             # steps miss cell centres (Q15)


def test_simple_flux_matrix(golden):
    q = mesh.build_simple_flux_dense(0.005, 8.0, 2.0)
    assert np.array_equal(q, golden["simple_flux"]["myQ"])
    csr = mesh.csr_from_dense(q, n_live=2)
    assert csr.row_lengths().tolist() == [3, 3, 0]


def test_matrix_csv_is_byte_identical_to_savetxt_and_round_trips(tmp_path, c1_csr):
    a, b = str(tmp_path / "a.csv"), str(tmp_path / "b.csv")
    mesh.save_matrix_csv(a, c1_csr)
    np.savetxt(b, c1_csr.to_dense(), delimiter=",")
    assert open(a, "rb").read() == open(b, "rb").read()
    back = mesh.load_matrix(a, n_live=900)
    assert np.array_equal(back.vals.view(np.uint64), c1_csr.vals.view(np.uint64))
    assert np.array_equal(back.cols, c1_csr.cols) and np.array_equal(back.indptr, c1_csr.indptr)
    assert mesh.load_matrix(a, n_live=900) is back                   # cached, not re-parsed
    z = str(tmp_path / "m.npz")
    mesh.save_matrix_npz(z, c1_csr)
    assert np.array_equal(mesh.load_matrix(z).vals, c1_csr.vals)


# ---------------------------------------------------------------- tables
def test_row_cum_equals_dense_cumsum(c1_csr):
    cum = tables.row_cum(c1_csr)
    dense = np.cumsum(c1_csr.to_dense(), axis=1)
    rows = np.repeat(np.arange(c1_csr.n_rows), c1_csr.row_lengths())
    assert np.array_equal(cum, dense[rows, c1_csr.cols])
    assert (cum[c1_csr.indptr[1:-1] - 1] < 1).sum() == 416           # 416 rows end below 1


@pytest.mark.parametrize("layout", ["compact", "wide"])
def test_alias_tables_reproduce_row_probabilities(c1_csr, layout):
    at = tables.build_alias(c1_csr, layout=layout)
    assert at.n_patterns == 31 and at.n_alias == 901                 # 30 stay patterns + the exit row
    for r in list(range(0, 900, 37)) + [899]:
        d = tables.alias_row_distribution(c1_csr, at, r)
        lo, hi = c1_csr.indptr[r], c1_csr.indptr[r + 1]
        p = c1_csr.vals[lo:hi] / c1_csr.vals[lo:hi].sum()
        assert set(d) == set(c1_csr.cols[lo:hi].tolist())
        assert max(abs(d[int(c)] - x) for c, x in zip(c1_csr.cols[lo:hi], p)) < 2e-10   # 2^-32 coin
    assert tables.alias_row_distribution(c1_csr, at, 900) == {}      # sink: never stepped


def test_alias_on_arbitrary_sparse_graph():
    """Ragged rows, an absorbing row in the middle, a zero row, a row with one huge and many tiny
    probabilities -- the densityrw-style generic graph."""
    rng = np.random.default_rng(3)
    n = 40
    q = np.zeros((n, n))
    for r in range(n - 1):
        k = int(rng.integers(1, 9))
        c = rng.choice(n, size=k, replace=False)
        w = rng.random(k) ** 4
        q[r, c] = w / w.sum()
    q[7] = 0
    q[11] = 0
    q[11, 3] = 1.0
    csr = mesh.csr_from_dense(q, n_live=n - 1)
    for layout in ("compact", "wide"):
        at = tables.build_alias(csr, layout=layout)
        for r in range(n):
            d = tables.alias_row_distribution(csr, at, r)
            if r >= n - 1 or r == 7:
                assert d == {}
                continue
            ref = {int(c): float(x) for c, x in zip(np.nonzero(q[r])[0], q[r][np.nonzero(q[r])])}
            assert set(d) <= set(ref)
            assert max(abs(d.get(c, 0.0) - x) for c, x in ref.items()) < 2e-9


def test_compact_layout_limits():
    q = np.zeros((3, 70000))
    q = np.zeros((70001, 70001))
    q[0, 0] = 0.5
    q[0, 70000] = 0.5
    csr = mesh.csr_from_dense(q)
    assert tables.build_alias(csr).layout == "wide"                  # relative column > 65535
    with pytest.raises(ValueError):
        tables.build_alias(csr, layout="compact")


def test_classify_source(c1_csr):
    m = c1_csr.meta["mesh"]
    for src_fn, val in ((mesh.source_default, 3.0), (mesh.source_paper, 0.015)):
        src = mesh.source_values(m["meshpoints"], src_fn)
        nc, cv = tables.classify_source(src, c1_csr.n_rows)
        assert cv.tolist() == [0.0, val, 0.0, 0.0] and nc[900] == 0
        assert np.array_equal(nc[:900] == 1, src != 0)
    with pytest.raises(ValueError):
        tables.classify_source(np.arange(10.0), 11)


# ---------------------------------------------------------------- index maps / sources
def test_scatter_to_rwsol_reproduces_the_reference_index_quirk():
    from nrw.fluence import scatter_to_rwsol
    NS, ND = 31, 31
    per = np.arange(900, dtype=np.float64) + 1
    ref = np.zeros((30, 30))
    for i in range(900):
        p, a = R.result_index(i, ND)
        ref[p, a] = per[i]
    got = scatter_to_rwsol(per, NS, ND)
    assert np.array_equal(got, ref)
    assert got[0, 29] == 31 and got[29, 29] == 0
    assert mesh.result_index(0, ND) == R.result_index(0, ND) == (0, 29)


def test_source_terms():
    for x, y in ((0.1, 0.0), (0.3, -0.2), (0.0, 0.05859375), (0.0, 0.0586), (-0.5, 0.3)):
        assert mesh.source_default(x, y) == R.source_term_default(x, y)
        assert mesh.source_paper(x, y) == R.source_term_paper(x, y)


def test_exp_time_weights_are_scalar_exp():
    from nrw.dropin import exp_time_weights
    w = exp_time_weights(0.7, 0.01, 300)
    assert all(w[j] == np.exp(-0.7 * 0.01 * j) for j in range(300))
    assert np.array_equal(exp_time_weights(0, 0.01, 5), np.ones(5))


# ---------------------------------------------------------------- CLI surface
def _ref_cli(path):
    """(flags, default-repr, help) triples parsed from the reference's add_argument lines."""
    src = open(path).read()
    out = {}
    for m in re.finditer(r"parser\.add_argument\((.*)\)\n", src):
        body = m.group(1)
        flags = tuple(re.findall(r"'(-{1,2}[A-Za-z_]+)'", body.split("help")[0].split("default")[0]))
        h = re.search(r"help=(['\"])(.*)\1", body)
        out[flags] = h.group(2) if h else None
    return out


@pytest.mark.skipif(not ref_import.reference_available(), reason="reference not mounted")
@pytest.mark.parametrize("name", ["fluence_mini_app", "simple_flux_miniapp"])
def test_cli_matches_reference_flags_defaults_and_help(name):
    import importlib
    app = importlib.import_module(name)
    ours = {flags: kw for flags, kw in app.CLI}
    ref = _ref_cli(os.path.join(ref_import.REFERENCE_ROOT, name + ".py"))
    assert len(ref) >= 18
    for flags, help_ in ref.items():
        assert flags in ours, flags
        assert ours[flags].get("help") == help_, flags
    refargs = (ref_import.fluence_default_args() if name == "fluence_mini_app"
               else ref_import.simple_flux_default_args())
    got = vars(app.build_parser().parse_args([]))
    for k, v in refargs.items():
        if k == "run_mode":
            assert got[k] == "fugu"                                   # the reference's default mode
        else:
            assert got[k] == v, k
    added = set(got) - set(refargs)
    assert added <= {"rng", "source", "sampler"}                      # additive flags only


def test_dispatch_table_keeps_the_reference_modes():
    import fluence_mini_app as f
    import simple_flux_miniapp as s
    for app in (f, s):
        assert set(app.run_functions) == {"fugu", "loihi", "spinnaker", "pc"}
        assert app.run_functions["pc"] is app.run_pc
        with pytest.raises(NotImplementedError, match="-r pc"):
            app.run_functions["spinnaker"](matrix_filename="m", neuraltime=1, initial_walkers=[[0, 1]], meshSize=1)
        # 'fugu' / 'loihi' are the reference's adapters over this build's densityrw / loihirw surfaces
        from nrw import dropin
        assert app.run_functions["fugu"] is dropin.fugu_adapter
        assert app.run_functions["loihi"] is dropin.loihi_adapter


def test_pc_mode_fails_loudly_without_a_gpu(tmp_path, c1_csr):
    """No CPU fallback: on a machine without CUDA the pc back end raises, it does not emulate."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import fluence_mini_app as app
    from nrw.engine import NrwError
    p = str(tmp_path / "m.csv")
    mesh.save_matrix_csv(p, c1_csr)
    with pytest.raises(NrwError, match="no CPU fallback"):
        app.run_pc(matrix_filename=p, neuraltime=8, initial_walkers=[[0, 5]], meshSize=900)


# ---------------------------------------------------------------- sharding / device-sampler port
def test_shard_range_partitions_exactly():
    for total in (0, 1, 7, 100, 12500, 10 ** 9 + 7):
        for world in (1, 2, 3, 8):
            for align in (1, 32):
                spans = [dist.shard_range(total, r, world, align) for r in range(world)]
                assert spans[0][0] == 0 and spans[-1][1] == total
                assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
                assert all(lo <= hi for lo, hi in spans)


def test_philox_known_answers():
    """Random123 known-answer vectors for Philox2x32-10."""
    kat = [((0x00000000, 0x00000000), 0x00000000, (0xff1dae59, 0x6cd10df2)),
           ((0xffffffff, 0xffffffff), 0xffffffff, (0x2c3f628b, 0xab4fd7ad)),
           ((0x243f6a88, 0x85a308d3), 0x13198a2e, (0xdd7ce038, 0xf62a4c12))]
    for (c0, c1), key, (e0, e1) in kat:
        r0, r1 = philox_port.philox2x32_10(np.array([c0]), np.array([c1]), key)
        assert (int(r0[0]), int(r1[0])) == (e0, e1)


def test_port_sharding_invariance_and_statistics(c1_csr):
    """The device-sampler port itself: slices with start_id_base sum to the unsharded tallies, and
    its mean score agrees with the reference chain's expectation."""
    from oracle import expectation
    src = mesh.source_values(c1_csr.meta["mesh"]["meshpoints"], mesh.source_default)
    nc, cv = tables.classify_source(src, c1_csr.n_rows)
    at = tables.build_alias(c1_csr, node_class=nc)
    starts, M, nt, seed = np.arange(0, 900, 9), 400, 64, 3
    full = philox_port.walk_port(c1_csr, starts, M, nt, seed, alias=at, class_value=[0, 1, 0, 0],
                                 start_id_base=starts * M)
    acc = np.zeros_like(full["score"])
    for rank in range(2):
        lo, hi = dist.shard_range(M, rank, 2)
        part = philox_port.walk_port(c1_csr, starts, hi - lo, nt, seed, alias=at, class_value=[0, 1, 0, 0],
                                     start_id_base=starts * M + lo)
        acc += part["score"]
    assert np.array_equal(acc, full["score"])
    mean, var = expectation.fluence_moments(c1_csr.indptr, c1_csr.cols, c1_csr.vals, 901, 900, src, nt,
                                            0.0, C1["dt"], C1["v"])
    est = full["score"] * 3.0 * C1["v"] * C1["dt"] / M
    sig = np.sqrt(var[starts] / M)
    z = np.abs(est - mean[starts]) / np.where(sig > 0, sig, 1)
    assert z.max() < 4.5 and (est[sig == 0] == mean[starts][sig == 0]).all()


def test_factorise_rows_is_exact_or_absent():
    """tables.factorise_rows: the factored descriptors reproduce the table bit for bit on fluence
    meshes (any source classes), and graphs without that structure are refused."""
    from nrw import mesh, tables
    for dt, da in ((0.01, 1.0 / 15.0), (0.005, 1.0 / 15.0), (0.01, 0.2), (0.005, 0.04)):
        csr = mesh.build_fluence_csr(dt, da, 200, 0.15, 1)
        for nc in (None, (np.arange(csr.n_rows) * 7 % 4).astype(np.uint8)):
            at = tables.build_alias(csr, node_class=nc)
            fr = tables.factorise_rows(at, csr.n_live)
            assert fr is not None and fr.period * fr.cells == csr.n_live
            assert np.array_equal(fr.synthesise(csr.n_live), at.rowinfo[:csr.n_live])
            assert np.array_equal(fr.tail, at.rowinfo[csr.n_live:])
            k = np.arange(csr.n_live, dtype=np.uint64)
            assert np.array_equal((k * np.uint64(fr.magic)) >> np.uint64(fr.shift), k // np.uint64(fr.period))
    q = mesh.build_simple_flux_dense(0.005, 8.0, 2.0)
    c = mesh.csr_from_dense(q, n_live=2)
    assert tables.factorise_rows(tables.build_alias(c), c.n_live) is None
    rng = np.random.default_rng(0)
    q = rng.random((40, 40)) * (rng.random((40, 40)) < 0.2)
    q[:, 0] += 1e-3
    c = mesh.csr_from_dense(q / q.sum(axis=1, keepdims=True), n_live=39)
    assert tables.factorise_rows(tables.build_alias(c), c.n_live) is None
    # one perturbed descriptor breaks the structure: refused, never approximated
    csr = mesh.build_fluence_csr(0.01, 1.0 / 15.0, 200, 0.15, 1)
    at = tables.build_alias(csr)
    at.rowinfo[100, 1] += 30
    assert tables.factorise_rows(at, csr.n_live) is None


def test_likely_paths_follow_the_most_probable_successor(c1_csr):
    from nrw import tables
    path = tables.likely_paths(c1_csr, [435, 900, 14], 12)
    assert path.shape == (3, 12) and path[0, 0] == 435 and (path[1, 1:] == -1).all() and path[1, 0] == 900
    for s in range(3):
        for t in range(1, 12):
            prev = path[s, t - 1]
            if prev < 0 or prev >= c1_csr.n_live:
                assert path[s, t] == -1
                continue
            lo, hi = c1_csr.indptr[prev], c1_csr.indptr[prev + 1]
            assert path[s, t] == c1_csr.cols[lo + np.argmax(c1_csr.vals[lo:hi])]


def test_tree_sampler_tables_realise_the_quantised_probability_circuit(c1_csr):
    """nrw.tables.build_tree_tables (the device table of sampler='tree') against loihirw.rw_core: walking the heap of
    8-bit thresholds with exact 7-bit uniforms gives tree_distribution(quantise=True) for every row pattern."""
    from loihirw import rw_core
    from nrw import tables
    thr, ptr = tables.build_tree_tables(c1_csr)
    lens = c1_csr.row_lengths()
    assert ptr[-1] == len(thr) and (np.diff(ptr)[lens <= 1] == 0).all()
    for r in (0, 14, 15, 435, 449, 885):
        n = int(lens[r])
        if n <= 1:
            continue
        lo, hi = c1_csr.indptr[r], c1_csr.indptr[r + 1]
        n2 = 1 << int(np.ceil(np.log2(n)))
        heap = thr[ptr[r]:ptr[r + 1]].astype(np.float64) / 128.0
        assert len(heap) == n2 - 1
        prob = np.ones(n2)                                   # exact distribution of the heap walk
        for leaf in range(n2):
            node = leaf + n2 - 1
            while node > 0:
                parent = (node - 1) // 2
                prob[leaf] *= heap[parent] if node == 2 * parent + 1 else 1.0 - heap[parent]
                node = parent
        want = rw_core.tree_distribution(list(c1_csr.vals[lo:hi]), quantise=True)
        assert np.allclose(prob, want, atol=1e-15) and prob[n:].sum() == 0


def test_launch_descriptors_build_without_a_device(c1_csr):
    """The host half of Engine.walk / prepare_walk (start list, id bases, tally buffers, likely paths, tree tables,
    the ctypes structs) on CPU tensors: everything up to the library call."""
    import ctypes as C
    import torch
    from nrw import engine as E, tables
    eng = E.Engine.__new__(E.Engine)
    eng.torch, eng.device, eng.lib, eng.handle = torch, torch.device("cpu"), E.load_library(), None
    tabs = E.DeviceTables(c1_csr, None, None, None, None, None, None, None, None, E._Tables())
    w, res, keep = eng._walk_struct(tabs, [435, 14], [5, 7], 40, 3, None, "alias", True, True, [0, 1, 0, 0],
                                    np.ones(40), True, True, 40, True, start_id_base=[0, 100], count_path=True)
    assert w.n_starts == 2 and w.walker_hi == 12 and w.t_cap == 40 and w.count_per_start == 1
    assert res.count_tn.shape == (2, 40, 901) and res.score.numel() == 2 and res.node_hist.numel() == 901
    path = keep["path"].numpy()
    assert path.shape == (2, 40) and path[0, 0] == 435 and np.array_equal(path, tables.likely_paths(c1_csr, [435, 14], 40))
    assert w.count_path == keep["path"].data_ptr() and not w.reduce
    eng._ensure_tree(tabs)
    assert tabs.struct.tree == tabs.tree[0].data_ptr() and tabs.tree_host[1][-1] == len(tabs.tree_host[0])
    with pytest.raises(E.NrwError):                      # the library refuses a NULL engine: no launch happened
        eng.prepare_walk(tabs, [435], 5, 40, 3, score=True, class_value=[0, 1, 0, 0])


def test_structured_rows_reproduce_every_alias_draw(c1_csr):
    """nrw.tables.structure_rows: target = base + (coin <= T ? slot : same) must be the alias draw's target for EVERY
    (slot, coin) of every row -- checked here exhaustively over the slots and at the coin values around the
    thresholds -- on the default mesh, the bench mesh and a mesh with factored descriptors; tables without the
    closed form (a general graph, a tampered entry) are refused."""
    from nrw import mesh, tables
    for csr in (c1_csr, mesh.build_fluence_csr(0.01, 0.02, 200, 0.15, 1), mesh.build_fluence_csr(0.001, 1.0 / 31.0, 200, 0.15, 1)):
        nc = (np.arange(csr.n_rows) % 3).astype(np.uint8)
        at = tables.build_alias(csr, node_class=nc)
        ri_s, T, n_draw = tables.structure_rows(at)
        w0, w1 = at.rowinfo[:, 0].astype(np.int64), at.rowinfo[:, 1].astype(np.int64)
        assert np.array_equal(ri_s[:, 1], at.rowinfo[:, 1]) and np.array_equal(ri_s[:, 0] >> 15, at.rowinfo[:, 0] >> 15)
        n, off, same = (w0 >> 15) & 0x3FF, w0 & 0x7FFF, ri_s[:, 0].astype(np.int64) & 0x7FFF
        th = at.alias[:, 0].astype(np.int64)
        own, ali = (at.alias[:, 1] & 0xFFFF).astype(np.int64), (at.alias[:, 1] >> 16).astype(np.int64)
        rows = np.repeat(np.arange(csr.n_rows), n)
        slot = np.arange(len(rows)) - np.repeat(np.cumsum(n) - n, n)
        e = off[rows] + slot
        for coin in (0, T - 1, T, T + 1, 0xFFFFFFFF):
            alias_t = np.where(coin <= th[e], own[e], ali[e])
            str_t = np.where(coin <= T, slot, same[rows])
            assert np.array_equal(alias_t, str_t), coin
        fr = tables.factorise_rows(at, csr.n_live, rowinfo=ri_s)
        assert fr is not None and np.array_equal(fr.synthesise(csr.n_live), ri_s[:csr.n_live])
        # the one length of the rows a walker draws from (nrw_tables.str_n): every row that is neither absorbing nor
        # a row whose single successor is absorbing (`term` bit)
        alive_on = (n > 0) & (((w0 >> 26) & 1) == 0)
        assert alive_on.any() and (n[alive_on] == n_draw).all() and n_draw == csr.meta["mesh"]["ND"] - 1
    # no closed form: a general graph, and a fluence table with one alias target changed
    q = np.random.default_rng(1).dirichlet(np.ones(6), size=6)
    assert tables.structure_rows(tables.build_alias(mesh.csr_from_dense(q, n_live=5))) is None
    at = tables.build_alias(c1_csr)
    k = int(at.rowinfo[435, 0] & 0x7FFF) + 3
    at.alias[k, 1] = (at.alias[k, 1] & 0xFFFF) | (((at.alias[k, 1] >> 16) ^ 1) << 16)
    assert tables.structure_rows(at) is None
    at = tables.build_alias(c1_csr)
    at.alias[k, 0] -= 1                                     # a second threshold
    assert tables.structure_rows(at) is None
    # rows of two lengths that a walker can draw from: a mesh whose structure holds entry by entry but whose walk
    # cannot multiply by one length is refused too
    at = tables.build_alias(c1_csr)
    r = int(np.nonzero(((at.rowinfo[:, 0] >> 15) & 0x3FF) > 1)[0][0])
    at.rowinfo[r, 0] -= np.uint32(1 << 15)                  # one entry fewer (its last slot is never drawn)
    sr = tables.structure_rows(at)
    assert sr is None
    # two stay rows of one cell trade places: every entry still has the closed form, but a walker sent to offset i no
    # longer lands on the row whose own direction is i -- the walk carries that offset as the direction, so: refused
    at = tables.build_alias(c1_csr)
    stay = np.nonzero(((at.rowinfo[:, 0] >> 15) & 0x3FF) > 1)[0]
    a = int(stay[np.nonzero(np.diff(stay) == 1)[0][0]])
    at.rowinfo[[a, a + 1]] = at.rowinfo[[a + 1, a]]
    assert tables.structure_rows(at) is None


def test_result_file_writer_is_np_savetxt_byte_for_byte(tmp_path):
    """nrw.mesh.save_array_csv writes the result files of run() (fluence_mini_app.py:446, simple_flux_miniapp.py:358):
    the bytes of np.savetxt(path, a, delimiter=','), for 1-D and 2-D arrays, signed zeros, denormals, inf and nan."""
    from nrw import mesh
    rng = np.random.default_rng(1)
    big = (rng.random((300, 37)) - 0.5) * 10.0 ** rng.integers(-300, 300, size=(300, 37))    # through libnrw.so
    big[5, 5], big[6, 0], big[7, 36], big[8, 1], big[9, 9] = np.nan, np.inf, -np.inf, -0.0, 5e-324
    big[10, 3] = -np.nan
    cases = [rng.random(17), rng.random((5, 7)) - 0.5, rng.random((100, 100)) * 4, np.zeros((3, 1)), big, big[:, 0],
             rng.random(5000),
             np.arange(6).reshape(2, 3), np.zeros((0, 4)),
             np.array([[0.0, -0.0, np.inf, -np.inf, np.nan, 1e-320, 1.7976931348623157e308]])]
    for k, a in enumerate(cases):
        p1, p2 = str(tmp_path / ("a%d.csv" % k)), str(tmp_path / ("b%d.csv" % k))
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")                     # np.savetxt warns about the empty array
            np.savetxt(p1, a, delimiter=",")
            mesh.save_array_csv(p2, a)
        assert open(p1, "rb").read() == open(p2, "rb").read(), (k, np.shape(a))


def test_format_csv_entry_argument_checks():
    """nrw_format_csv_e18 (host only): too small an output buffer is NRW_ENOSPC with both sizes in the message, bad
    arguments are NRW_EINVAL, one thread and many threads give the same bytes."""
    import ctypes as C
    from nrw.engine import load_library
    lib = load_library()
    a = np.ascontiguousarray(np.random.default_rng(3).standard_normal((64, 40)))
    out = C.create_string_buffer(a.size * 28)
    n = C.c_int64(0)
    assert lib.nrw_format_csv_e18(C.c_void_p(a.ctypes.data), 64, 40, out, 100, C.byref(n), 4) == -3
    assert "100" in lib.nrw_last_error().decode()
    assert lib.nrw_format_csv_e18(None, 64, 40, out, a.size * 28, C.byref(n), 4) == -1
    assert lib.nrw_format_csv_e18(C.c_void_p(a.ctypes.data), 64, 0, out, a.size * 28, C.byref(n), 4) == -1
    texts = []
    for threads in (1, 3, 64):
        assert lib.nrw_format_csv_e18(C.c_void_p(a.ctypes.data), 64, 40, out, a.size * 28, C.byref(n), threads) == 0
        texts.append(out.raw[:n.value])
    assert texts[0] == texts[1] == texts[2]
    want = "".join(",".join("%.18e" % v for v in row) + "\n" for row in a).encode()
    assert texts[0] == want


def test_capture_digest_ignores_comments_and_white_space():
    """tools/ncu_summary.csrc_sha256 hashes the CODE of the CUDA sources (bench.py refuses a capture of other kernel
    code, not of a re-worded comment): strip_comments drops // and /* */ comments and collapses white space, and
    leaves string and character literals alone."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from ncu_summary import strip_comments
    src = 'int a = 1; // c1\n/* block\n comment */ const char* s = "// kept /* too */";\tchar c = \'"\';   // tail\nasm("x;" :: "r"(a));'
    assert strip_comments(src) == 'int a = 1; const char* s = "// kept /* too */"; char c = \'"\'; asm("x;" :: "r"(a));'
    assert strip_comments("a  =\n\n b ;") == strip_comments("a = b ; // same code")
