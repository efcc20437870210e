#!/bin/bash
# One GPU-box pass: parity tests, kernel A/B, bench lines, ncu launch list and full capture.  Every leg has
# its own timeout and its own log under gpurun_out/, so a leg that fails or hangs does not take the others
# down.  Usage (from the repo root, on the box):  bash tools/gpu_round.sh [tag]
set -u
tag=${1:-r01}
out=gpurun_out
mkdir -p $out
leg() {  # leg <name> <timeout_s> <cmd...>
    local name=$1 t=$2; shift 2
    local t0=$(date +%s)
    timeout "$t" "$@" > "$out/${tag}_${name}.log" 2> "$out/${tag}_${name}.err"
    echo "$name rc=$? $(( $(date +%s) - t0 ))s" | tee -a "$out/${tag}_legs.txt"
}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader > "$out/${tag}_gpu.txt" 2>&1
leg pytest_gpu 540 python -m pytest tests -m gpu -q -p no:cacheprovider
leg variants_c5 150 python tools/k1s_variants.py c5
leg variants_c1 120 python tools/k1s_variants.py c1
leg bench_n1 420 python bench.py
leg bench_n1_score_hist 200 python bench.py --tally score+hist --no-extras
leg side_c4 200 python tools/side_benches.py c4
leg ncu_launches 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file "$out/${tag}_launches_bench_steps2.csv" python bench.py --steps 2 --warmup 3 --no-extras
leg ncu_walk_score_full 240 ncu --set full --clock-control none --import-source on -k regex:walk_score -s 4 -c 1 \
    -f -o "$out/${tag}_walk_score" python bench.py --steps 2 --warmup 3 --no-extras
leg smoke 300 python -c "import __graft_entry__ as g; g.smoke()"
cat "$out/${tag}_legs.txt"
tail -3 "$out/${tag}_pytest_gpu.log"
cat "$out/${tag}_variants_c5.log" "$out/${tag}_variants_c1.log" | cut -c1-260
cat "$out/${tag}_bench_n1.log" | cut -c1-600
