#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU): key raw metrics per launch and the hot SASS lines.
Usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [--sass] [--min-share 0.002]"""
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


def main():
    rep = sys.argv[1]
    rows = list(csv.reader(io.StringIO(run([rep, "--page", "raw", "--csv"]))))
    h, u, data = rows[0], rows[1], rows[2:]
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
