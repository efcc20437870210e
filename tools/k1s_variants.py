#!/usr/bin/env python
"""A/B timing of the score kernel's launch variants on a mesh (run on a B200).
Usage: python tools/k1s_variants.py [c1|c5|c4] [walkers_per_node]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "neural-random-walker_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    import torch
    from nrw import mesh, tables
    from nrw.engine import Engine
    which = sys.argv[1] if len(sys.argv) > 1 else "c1"
    par = dict(c1=(0.01, 1.0 / 15.0), c5=(0.01, 0.02), c4=(0.001, 1.0 / 63.0))[which]
    M = int(sys.argv[2]) if len(sys.argv) > 2 else dict(c1=150000, c5=12500, c4=200)[which]
    eng = Engine(0)
    csr = mesh.build_fluence_csr(par[0], par[1], 200, 0.15, 1)
    src = mesh.source_values(csr.meta["mesh"]["meshpoints"], mesh.source_default)
    nc, _ = tables.classify_source(src, csr.n_rows)
    tabs = eng.upload(csr, node_class=nc)
    starts = np.arange(csr.n_live)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=eng.device)
    knobs = ("NRW_CPS", "NRW_DET", "NRW_K1", "NRW_RANGE_TAIL", "NRW_TAIL_RANGES", "NRW_RANGES_PER_WARP", "NRW_BLOCK")
    configs = [("default", {}), ("cps2", {"NRW_CPS": "2"}), ("single_range_size", {"NRW_TAIL_RANGES": "0"}),
               ("always_draw", {"NRW_DET": "0"}), ("general", {"NRW_K1": "general"}),
               ("block768", {"NRW_BLOCK": "768"}), ("block512", {"NRW_BLOCK": "512"}), ("block256", {"NRW_BLOCK": "256"})]
    if len(sys.argv) > 3 and sys.argv[3] == "sweep":        # guided range sizes: smallest range x tail share
        configs = [("single_range_size", {"NRW_TAIL_RANGES": "0"}), ("general_control", {"NRW_K1": "general"})]
        for rt, tr in ((2048, 2), (1024, 2), (512, 2), (256, 2), (128, 2), (64, 2), (1024, 1), (1024, 3), (512, 1),
                       (512, 3), (1024, 4)):
            configs.append(("min%d_x%d" % (rt, tr), {"NRW_RANGE_TAIL": str(rt), "NRW_TAIL_RANGES": str(tr)}))
        for rpw in (3, 4, 6, 12, 16, 24):
            for rt, tr in ((1024, 2), (512, 2), (1024, 3)):
                configs.append(("rpw%d_min%d_x%d" % (rpw, rt, tr), {"NRW_RANGES_PER_WARP": str(rpw), "NRW_RANGE_TAIL": str(rt),
                                                                     "NRW_TAIL_RANGES": str(tr)}))
    only = os.environ.get("NRW_VARIANTS")                   # e.g. "default,always_draw": a subset by name
    if only:
        configs = [c for c in configs if c[0] in only.split(",")]
    for tag, env in configs:
        # the library reads its switches once per engine: one engine per variant (device tables are shared)
        sw = {k: None for k in knobs}
        sw.update(env)
        ev_eng = Engine(0, switches=sw)
        launch, res = ev_eng.prepare_walk(tabs, starts, M, 30000, seed=1, score=True, class_value=[0, 1, 0, 0])
        for _ in range(3):
            launch()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(10)]
        for a, b in ev:
            flush.fill_(1)
            launch.zero()
            a.record()
            launch.launch_kernel()
            b.record()
        torch.cuda.synchronize()
        ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
        steps = int(res.counters.cpu()[0].item())
        print(json.dumps({"lib": os.path.basename(os.environ.get("NRW_LIB", "libnrw.so")), "mesh": which, "M": M, "variant": tag, "kernel": res.kernel, "geometry": res.launch_geometry,
                          "kernel_ms": ms, "walker_steps": steps, "walker_steps_per_s": steps / (ms * 1e-3),
                          "score_sum": int(res.score.sum().item()),
                          "score_crc": int((res.score.cpu().numpy().astype(np.uint64) * (np.arange(len(starts), dtype=np.uint64) * np.uint64(2654435761) + np.uint64(1))).sum() & np.uint64(0xFFFFFFFFFFFF))}), flush=True)


if __name__ == "__main__":
    main()
