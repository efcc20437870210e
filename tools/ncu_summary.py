#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU): key raw metrics per launch and the hot SASS lines.
Usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [--sass] [--min-share 0.002]
       python tools/ncu_summary.py prof.ncu-rep --traffic-json profiles/walk_traffic.json --bench-line bench.json
           --source "profiles/<summary>.txt (...)" --round 2
The second form regenerates profiles/walk_traffic.json -- the per-launch facts of the dominant kernel that
bench.py's roofline object is built from (warp-instructions, lanes per instruction, DRAM bytes, pipe use) -- from
the capture, and takes the walker-steps per launch from the bench line of the SAME command, so that bench.py can
refuse a capture of another kernel or workload."""
import csv
import io
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'launch__registers_per_thread', 'launch__shared_mem_per_block_dynamic',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio',
        'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_issued.avg.pct_of_peak_sustained_active',
        'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_cbu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
        'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_red.sum',
        'lts__t_sectors.sum', 'sm__cycles_elapsed.avg', 'smsp__cycles_active.avg',
        'smsp__cycles_active.max', 'smsp__cycles_active.min', 'sm__cycles_active.max', 'sm__cycles_active.min']


def run(args):
    return subprocess.run(["ncu", "-i"] + args, capture_output=True, text=True).stdout


TAB_NAMES = {0: "TAB_SMEM", 1: "TAB_HYB", 2: "TAB_GLB", 3: "TAB_CDF", 4: "TAB_FAC", 5: "TAB_TREE"}


def plan_name(ncu_name):
    """ncu's demangled kernel name -> the planner's name (nrw_walk_philox_plan), e.g.
    walk_score_kernel<0, 1, 1, 1, 0> -> walk_score<TAB_SMEM,true,true,cps=1>."""
    import re
    m = re.search(r"walk_score_kernel<([^>]*)>", ncu_name)
    if m:
        args = [int(x) for x in re.findall(r"\d+", re.sub(r"\((?:int|bool)\)", "", m.group(1)))]
        tab, idw, det, cps, hist = args[:5]
        if len(args) > 5 and args[5]:
            return "walk_count<%s>" % TAB_NAMES[tab]
        strv = "_str" if len(args) > 6 and args[6] else ""
        return "%s%s<%s,%s,%s,cps=%d>" % ("walk_score_hist" if hist else "walk_score", strv, TAB_NAMES[tab],
                                        "true" if idw else "false", "true" if det else "false", cps)
    m = re.search(r"walk_philox_kernel<\(?(?:int\))?(\d+), \(?(?:int\))?(\d+)>", ncu_name)
    if m:
        return "walk_philox<%s,%d>" % (TAB_NAMES[int(m.group(1))], int(m.group(2)))
    return ncu_name


def strip_comments(src: str) -> str:
    """C++ source without its comments and with runs of white space collapsed (string literals kept as they are):
    what the digest below hashes, so that a comment or a re-wrapped line does not invalidate a capture."""
    import re
    pat = re.compile(r'"(?:\\.|[^"\\])*"|\'(?:\\.|[^\'\\])*\'|//[^\n]*|/\*.*?\*/', re.S)
    code = pat.sub(lambda m: m.group(0) if m.group(0)[0] in "\"'" else " ", src)
    return re.sub(r"\s+", " ", code).strip()


def csrc_sha256():
    """Digest of the CUDA sources the library is built from -- their CODE: comments and white space are left out --
    so that bench.py refuses a capture of other kernel code."""
    import glob
    import hashlib
    import os
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "neural-random-walker_b200", "nrw", "csrc")
    h = hashlib.sha256()
    for f in sorted(glob.glob(os.path.join(root, "*.cu")) + glob.glob(os.path.join(root, "*.cuh"))):
        h.update(os.path.basename(f).encode())
        h.update(strip_comments(open(f, "r", encoding="utf-8", errors="replace").read()).encode())
    return h.hexdigest()


def traffic_json(h, data, out_path, argv):
    import json
    r = data[0]
    val = lambda k: float(r[h.index(k)].replace(",", "")) if k in h and r[h.index(k)] not in ("", "n/a") else None
    unit_scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}
    def bytes_of(k, units):
        v = val(k)
        return None if v is None else v * unit_scale.get(units[h.index(k)], 1.0)
    units = argv["units"]
    facts = {"kernel": plan_name(r[h.index("Kernel Name")]),
             "dram_bytes_per_launch": (bytes_of("dram__bytes_read.sum", units) or 0.0) + (bytes_of("dram__bytes_write.sum", units) or 0.0),
             "round": argv.get("round"), "source": argv.get("source"), "csrc_sha256": csrc_sha256(),
             "ncu_kernel_us": (val("gpu__time_duration.sum") or 0.0) * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(
                 units[h.index("gpu__time_duration.sum")], 1.0),
             "warp_inst_per_launch": val("smsp__inst_executed.sum"),
             "threads_per_warp_inst": val("smsp__thread_inst_executed_per_inst_executed.ratio"),
             "issue_active_pct": val("smsp__issue_active.avg.pct_of_peak_sustained_active"),
             "alu_pipe_pct": val("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
             "fmaheavy_pipe_pct": val("sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed"),
             "smsp_active_share_of_elapsed": (val("smsp__cycles_active.avg") or 0.0) / max(val("sm__cycles_elapsed.avg") or 1.0, 1.0),
             "smem_bank_conflict_share": (val("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum") or 0.0)
                                         / max(val("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum") or 1.0, 1.0)}
    if argv.get("bench_line"):
        # the last JSON line of the log (under ncu the log ends with ncu's own ==PROF== lines)
        line = json.loads([l for l in open(argv["bench_line"]).read().splitlines() if l.startswith("{")][-1])
        facts["walker_steps_per_launch"] = int(line["roofline"]["walker_steps_per_launch"])
        assert line["roofline"]["kernel"] == facts["kernel"], (line["roofline"]["kernel"], facts["kernel"])
        facts["warp_inst_per_32_walker_steps"] = 32.0 * facts["warp_inst_per_launch"] / facts["walker_steps_per_launch"]
    json.dump(facts, open(out_path, "w"), indent=1)
    print("wrote", out_path, facts)


def main():
    rep = sys.argv[1]
    rows = list(csv.reader(io.StringIO(run([rep, "--page", "raw", "--csv"]))))
    h, u, data = rows[0], rows[1], rows[2:]
    if "--traffic-json" in sys.argv:
        opt = lambda k: sys.argv[sys.argv.index(k) + 1] if k in sys.argv else None
        traffic_json(h, data, opt("--traffic-json"), {"units": u, "bench_line": opt("--bench-line"), "source": opt("--source"),
                                                       "round": int(opt("--round") or 0)})
        return
    keys = KEYS + [k for k in h if k.startswith('smsp__average_warps_issue_stalled') and k.endswith('per_issue_active.ratio')]
    print("# kernel:", ' | '.join(r[h.index('Kernel Name')] for r in data))
    for k in keys:
        if k in h:
            i = h.index(k)
            vals = [r[i] for r in data]
            if all(v in ('0', '', 'n/a') for v in vals) and 'stalled' in k:
                continue
            print("%-86s %-14s %s" % (k, u[i], ' | '.join(vals)))
    if "--sass" in sys.argv:
        share = 0.002
        if "--min-share" in sys.argv:
            share = float(sys.argv[sys.argv.index("--min-share") + 1])
        rows = list(csv.reader(io.StringIO(run([rep, "--page", "source", "--csv", "--print-source", "sass"]))))
        blocks, cur = [], None
        for r in rows:
            if r and r[0] == 'Kernel Name':
                cur = []
                blocks.append((r[1], cur))
            elif cur is not None:
                cur.append(r)
        name, b = blocks[0]
        hh, d = b[0], b[1:]
        ia, isrc, ie, it, iss = (hh.index(x) for x in ('Address', 'Source', 'Instructions Executed',
                                                      'Thread Instructions Executed', '# Samples'))
        tot = sum(int(r[ie]) for r in d if r[ie].isdigit())
        print("# SASS of %s: lines with > %.2f%% of %d warp-instructions" % (name, share * 100, tot))
        print("# addr   warp-inst(M)  avg-threads  stall-samples  instruction")
        for r in d:
            e = int(r[ie]) if r[ie].isdigit() else 0
            if e > share * tot:
                print("%s %9.2fM %5.1f %6s  %s" % (r[ia][-5:], e / 1e6, int(r[it]) / max(e, 1), r[iss], r[isrc][:100]))


if __name__ == "__main__":
    main()
