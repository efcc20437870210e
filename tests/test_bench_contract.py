"""bench.py's output contract: the keys the driver reads.  The GPU arm cannot run here, so its latest
committed line (profiles/) is checked; the reference arm is the CPU port and runs here for real."""
import glob
import json
import os
import subprocess
import sys

from conftest import ROOT

BASE_KEYS = ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
             "scaling", "vs_baseline", "dtype", "data", "config")


def _last_line(path):
    return json.loads(open(path).read().strip().splitlines()[-1])


def test_committed_gpu_lines_carry_the_contract_keys():
    """Round-2 lines: N = 1 (full, with the side legs), N = 2 and N = 8 (the walk's fused exchange)."""
    paths = [os.path.join(ROOT, "profiles", n) for n in ("r02v_bench_n1_full.json", "r02w_bench_n2.json",
                                                         "r02w_bench_n8.json")]
    for path in paths:
        d = _last_line(path)
        for k in BASE_KEYS + ("roofline", "e2e", "gpu_launches", "clocks", "ms_p50", "ms_p95", "fluence_err",
                              "drawn_steps_per_s", "value_always_draw"):
            assert k in d, (path, k)
        assert d["metric"] == "walker-steps/s" and d["higher_is_better"] is True and d["scaling"] == "weak"
        assert d["vs_baseline"] is None                        # BASELINE.md holds no published number
        assert "workload" in d["config"] and "model" not in d["config"]
        r = d["roofline"]
        for k in ("bound", "achieved", "peak", "unit", "frac", "traffic", "mem"):
            assert k in r, k
        # the binding resource is instruction issue; no key claims HBM, and the fraction is physical
        assert r["bound"] == "issue" and r["unit"] == "warp-inst/s"
        if r["frac"] is not None:
            assert 0.0 < r["frac"] <= 1.0 and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
            # recomputable from the line's own fields: warp-instructions / (SMs x 4 schedulers x clock x time)
            again = r["warp_inst_per_launch"] / (r["kernel_ms"] * 1e-3) / (148 * 4 * d["clocks"]["sm_mhz"] * 1e6)
            assert abs(again - r["frac"]) < 0.03 * r["frac"]
            assert 0.5 < r["lane_efficiency"] <= 1.0
        assert r["mem"]["hbm_frac"] is None or r["mem"]["hbm_frac"] < 0.01
        e = d["e2e"]
        assert e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0 and 0 < e["value"] < d["value"]
        assert d["gpu_launches"] == d["steps"]                 # one kernel per step on any N
        assert not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
        f = d["fluence_err"]
        assert f["max_abs_z"] < 5.5 and f["frac_within_3sigma"] >= 0.995 and f["zero_variance_cells_exact"]
        assert f["rel_l2"] < 1.5 * f["expected_rel_l2"]
        if d["n_gpus"] > 1:
            assert d["multi_rank_parity"]["sharded_equals_unsharded"] is True
            assert "tail" in d["collective"]
    one = _last_line(paths[0])
    c = one["cpu_baseline"]
    for k in ("value", "unit", "cores", "kind", "sample"):
        assert k in c, k
    assert c["kind"] in ("port", "reference") and c["unit"] == one["unit"]
    assert one["roofline"]["frac"] is not None and "stale" not in one["roofline"]
    for leg in ("c2", "c3", "c4", "e2e_run_args", "resources"):
        assert leg in one, leg
    assert one["e2e_run_args"]["warm_s"] < 0.025               # run(args) on 1e9 walkers: VERDICT r01 item 3
    assert one["c3"]["conservation_ok"] and one["c4"]["densityrw"]["counts_total_ok"]


def test_reference_arm_runs_on_the_host_and_prints_one_json_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1                                     # stdout carries the one JSON line only
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["value"] > 0
    for k in BASE_KEYS + ("cpu_baseline", "e2e"):
        assert k in d, k
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["e2e"]["value"] == d["value"] and d["cpu_baseline"]["value"] == d["value"]
    # the reference itself where a copy is reachable (this container), the oracle's port of it elsewhere (GPU box)
    have_ref = os.path.isfile("/root/reference/fluence_mini_app.py")
    assert d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["kind"] == ("reference" if have_ref else "port")
    # the arm describes the workload IT ran (the reference's M, nt and RNG), not the GPU arm's
    assert "M=100" in d["config"]["workload"] and "MT19937" in d["config"]["workload"]
    assert d["config"]["workers"] == d["cpu_baseline"]["cores"]
