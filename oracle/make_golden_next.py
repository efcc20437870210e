"""Generate tests/golden/next_surface.npz from the UNMODIFIED reference -- TEST INFRASTRUCTURE.

Pins the "next" rows of SURVEY section 8(f): the reference's own spike-raster decoders
(loihirw/rw_analysis.py), its transition loader (densityrw/DensityRW.py:177-205), its probability
tree (loihirw/rw_core.py:11-156) and the spike-table post-processing of fluence_mini_app.run_fugu
(:72-107), each run here on small seeded inputs.  Run in the build container only:

    python -m oracle.make_golden_next
"""
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "neural-random-walker_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

from oracle import ref_import  # noqa: E402


class P:
    """probe stand-in: .data[neuron][time]"""
    def __init__(self, data):
        self.data = data


def random_rasters(rng, n_nodes, n_bins, max_count):
    """Boundary raster with irregular bin lengths + one-spike-per-walker rasters."""
    lens = rng.integers(max_count + 1, max_count + 6, size=n_bins)
    ends = np.cumsum(lens) - 1 + np.arange(n_bins)          # boundary ticks, strictly increasing
    T = int(ends[-1]) + 4
    b = np.zeros(T, dtype=np.uint8)
    b[ends] = 1
    starts = np.concatenate([[0], ends[:-1] + 1])
    counts = rng.integers(0, max_count + 1, size=(n_nodes, n_bins))
    counts[rng.random((n_nodes, n_bins)) < 0.4] = 0
    spikes = np.zeros((n_nodes, T), dtype=np.uint8)
    for j in range(n_nodes):
        for i in range(n_bins):
            width = ends[i] - starts[i]
            pos = rng.choice(width, size=counts[j, i], replace=False) + starts[i]
            spikes[j, pos] = 1
    return b, spikes, counts


def main():
    out = {}
    rng = np.random.default_rng(20240611)
    A = ref_import.rw_analysis()

    # ---- rw_analysis decoders ------------------------------------------------------------
    for case, (n_nodes, n_bins, mx) in enumerate([(5, 7, 6), (12, 20, 3), (3, 1, 9)]):
        b, spikes, counts = random_rasters(rng, n_nodes, n_bins, mx)
        T = len(b)
        z = np.zeros(T, dtype=np.uint8)
        mesh_data = [np.stack([b, spikes[j], z]) for j in range(n_nodes)]
        # track 6 (split)
        tp = [[["Supervisor Global Spike", P(np.stack([b, b]))], ["Supervisor Fire Spike", P(np.stack([b, z, z]))]],
              [["Mesh Node: %d" % j, P(mesh_data[j])] for j in range(n_nodes)]]
        c6, nb6 = A.process_outputs_split(tp, 0, 1, n_nodes)
        # track 2/5 (flat, start 3, stride 5; entries are probe lists)
        sup = [P(np.stack([b, b]))]
        flat = [sup, sup, sup]
        for j in range(n_nodes):
            flat.append([P(mesh_data[j])])
            flat.extend([sup] * 4)
        c2, nb2 = A.process_outputs(flat, 3, 5, n_nodes)
        out.update({"an%d_boundary" % case: b, "an%d_spikes" % case: spikes, "an%d_counts" % case: counts,
                    "an%d_split_count" % case: c6, "an%d_split_bins" % case: np.array([nb6]),
                    "an%d_flat_count" % case: c2, "an%d_flat_bins" % case: np.array([nb2])})
    # track 1/3 (end-state base-1000 counters)
    totals = np.array([0, 7, 999, 1000, 123456, 2999999])
    dig = lambda v: [[int(v) % 1000], [(int(v) // 1000) % 1000], [int(v) // 1000000]]
    n_steps = 77
    ends = [[], [P(dig(128 * n_steps))], [], []]
    for v in totals:
        ends.append([P(dig(v))])
        ends.extend([[]] * 4)
    ce, ns = A.process_outputs_ends(ends, 4, 5, len(totals), 1)
    cs = A.count_steps(ends, 4, 5, len(totals), 1)
    out.update(ends_totals=totals, ends_n_steps=np.array([n_steps]), ends_count=ce, ends_num_steps=ns,
               ends_count_steps=cs)

    # ---- densityrw.load_transitions ------------------------------------------------------
    D = ref_import.density_rw()
    q = np.zeros((7, 7))
    q[0, [1, 2]] = [0.25, 0.75]
    q[1, [0, 1, 6]] = [0.5, 0.125, 0.375]
    q[2, 2] = 1.0                      # sink: probability-1 self loop
    q[3, [0, 4]] = [1.0 / 3.0, 2.0 / 3.0]
    q[4, 5] = 1.0                      # single edge that is NOT a self loop: kept
    # row 5: no edges at all
    q[6, 6] = 0.5                      # self loop with p != 1: kept
    q[6, 0] = 0.5
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "graph.csv")
        np.savetxt(path, q, delimiter=",")
        for flag in (True, False):
            tr = D.load_transitions(path, remove_sink_connections=flag, verbose=0)
            src, dst, prob = [], [], []
            for k in sorted(tr):
                assert tr[k].location == k
                for (dest, p) in tr[k].neighbors:
                    src.append(k[0]); dst.append(dest[0]); prob.append(p)
            tag = "lt_rm%d" % int(flag)
            out.update({tag + "_src": np.array(src), tag + "_dst": np.array(dst), tag + "_prob": np.array(prob),
                        tag + "_n": np.array([len(tr)])})
    out["lt_matrix"] = q

    # ---- rw_core.get_probCircuitValues ---------------------------------------------------
    C = ref_import.rw_core()
    probs = [[0.5, 0.25, 0.25], [0.1, 0.2, 0.3, 0.4], [0.2, 0.0, 0.3, 0.1, 0.4], [1.0],
             [0.05] * 3 + [0.85], list(rng.dirichlet(np.ones(7))), [0.3, 0.3], [0.0, 0.6, 0.4],
             list(rng.dirichlet(np.ones(16))), [0.25, 0.25, 0.25, 0.25, 0.0, 0.0]]
    kept = 0
    for pr in probs:
        try:
            conn, relay, weight = C.get_probCircuitValues(list(pr))
        except (IndexError, ValueError) as e:      # the reference fails on some fan-outs (zeros, n=1)
            print("get_probCircuitValues(%r) raises %s in the reference: not pinned" % (pr, type(e).__name__))
            continue
        out.update({"pc%d_prob" % kept: np.array(pr), "pc%d_conn" % kept: conn, "pc%d_relay" % kept: relay,
                    "pc%d_weight" % kept: weight})
        kept += 1
    out["pc_n"] = np.array([kept])

    # ---- run_fugu post-processing on the product's spike table ---------------------------
    from densityrw.rw_miniapp_fugu import counts_to_spike_table
    F = ref_import.fluence()
    meshSize = 6
    counts = np.zeros((meshSize + 1, 9), dtype=np.int64)
    counts[2, 0] = 11
    for t in range(1, 8):
        left = int(counts[:meshSize, t - 1].sum())
        if left == 0:
            break
        stay = rng.multinomial(left - left // 4, np.ones(meshSize) / meshSize)
        counts[:meshSize, t] = stay
        counts[meshSize, t] = left // 4
    table = counts_to_spike_table(counts, include_rows=np.arange(meshSize + 1) < meshSize)
    fake = types.ModuleType("densityrw.rw_miniapp_fugu")
    fake.run_miniapp = lambda *a, **k: table.copy()
    pkg = types.ModuleType("densityrw")
    saved = {k: sys.modules.get(k) for k in ("densityrw", "densityrw.rw_miniapp_fugu")}
    sys.modules["densityrw"], sys.modules["densityrw.rw_miniapp_fugu"] = pkg, fake
    try:
        td = F.run_fugu(matrix_filename="unused.csv", neuraltime=9, initial_walkers=[[2, 11]],
                        meshSize=meshSize, runtime_args={"fugu_backend": "snn", "use_sinks": False})
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    out.update(fugu_counts=counts, fugu_tempdata=td)

    dst = os.path.join(ROOT, "tests", "golden", "next_surface.npz")
    np.savez_compressed(dst, **out)
    print("wrote", dst, os.path.getsize(dst), "bytes;", len(out), "arrays")
    print("run_fugu tempdata shape", td.shape, "counts nonzero steps", int((counts.sum(0) > 0).sum()))


if __name__ == "__main__":
    main()
