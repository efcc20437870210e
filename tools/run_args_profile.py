#!/usr/bin/env python
"""Where the time of the plugin-level call goes: fluence_mini_app.run(args) on the C5 mesh with
-M 100000 --rng philox (run on a B200).  Prints one JSON object: process start-up costs (import torch,
CUDA context, engine), then run(args) cold and warm with the split the drop-in logs at DEBUG, and the
top of a cProfile of the cold call.  Usage: python tools/run_args_profile.py [M] [--profile]"""
import cProfile
import io
import json
import os
import pstats
import sys
import tempfile
import time

T0 = time.perf_counter()
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "neural-random-walker_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    M = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 100000
    out = {}
    t = time.perf_counter()
    import torch
    out["import_torch_s"] = time.perf_counter() - t
    t = time.perf_counter()
    torch.cuda.init()
    torch.zeros(1, device="cuda")
    torch.cuda.synchronize()
    out["cuda_context_s"] = time.perf_counter() - t
    t = time.perf_counter()
    import fluence_mini_app as app
    out["import_app_s"] = time.perf_counter() - t
    t = time.perf_counter()
    from nrw.engine import get_engine
    get_engine()
    out["engine_create_s"] = time.perf_counter() - t
    tmp = tempfile.mkdtemp()
    args = vars(app.build_parser().parse_args(
        ["-r", "pc", "-M", str(M), "-dt", "0.01", "-da", "0.02", "--rng", "philox", "--rand_seed", "1",
         "--matrix_filename", os.path.join(tmp, "curmat.csv"), "--results_filename", os.path.join(tmp, "rwSol.csv")]))
    import numpy as np
    np.random.seed(1)
    prof = cProfile.Profile() if "--profile" in sys.argv else None
    calls = []
    for i in range(4):
        t = time.perf_counter()
        if prof is not None and i == 0:
            prof.enable()
        sol = app.run(args)
        if prof is not None and i == 0:
            prof.disable()
        calls.append(time.perf_counter() - t)
    out["run_args_s"] = calls
    out["run_args_split_last"] = getattr(app, "last_run_split", None)
    out["rwSol_sum"] = float(sol.sum())
    out["process_s"] = time.perf_counter() - T0
    if prof is not None:
        s = io.StringIO()
        pstats.Stats(prof, stream=s).sort_stats("cumulative").print_stats(30)
        out["cprofile_cold"] = s.getvalue().splitlines()[-45:]
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
