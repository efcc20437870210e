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
    lines = sorted(glob.glob(os.path.join(ROOT, "profiles", "r01t_bench_n1_full.json"))
                   + glob.glob(os.path.join(ROOT, "profiles", "r01s_bench_n2.json")))
    assert len(lines) == 2
    for path in lines:
        d = _last_line(path)
        for k in BASE_KEYS + ("roofline", "e2e", "gpu_launches", "clocks"):
            assert k in d, (path, k)
        assert d["metric"] == "walker-steps/s" and d["higher_is_better"] is True and d["scaling"] == "weak"
        assert d["vs_baseline"] is None                        # BASELINE.md holds no published number
        assert "workload" in d["config"] and "model" not in d["config"]
        r = d["roofline"]
        for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
            assert k in r, k
        assert r["bound"] in ("hbm", "tensor") and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
        e = d["e2e"]
        assert e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0 and 0 < e["value"] < d["value"]
        assert d["gpu_launches"] >= d["steps"]
        assert not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    one = _last_line(os.path.join(ROOT, "profiles", "r01t_bench_n1_full.json"))
    c = one["cpu_baseline"]
    for k in ("value", "unit", "cores", "kind", "sample"):
        assert k in c, k
    assert c["kind"] in ("port", "reference") and c["unit"] == one["unit"]


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
