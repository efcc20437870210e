#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench line, ncu launch list and full capture.  Every leg has its own
# timeout and its own log under gpurun_out/, so a leg that fails or hangs does not take the others down.
# Usage (from the repo root, on the box):  bash tools/gpu_round.sh [tag] [legs...]
set -u
tag=${1:-r02}
shift || true
legs=${*:-"pytest smoke bench ncu_launches ncu_full"}
out=gpurun_out
mkdir -p $out
leg() {  # leg <name> <timeout_s> <cmd...>
    local name=$1 t=$2; shift 2
    local t0=$(date +%s)
    timeout "$t" "$@" > "$out/${tag}_${name}.log" 2> "$out/${tag}_${name}.err"
    echo "$name rc=$? $(( $(date +%s) - t0 ))s" | tee -a "$out/${tag}_legs.txt"
}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader > "$out/${tag}_gpu.txt" 2>&1
for l in $legs; do
case $l in
pytest) leg pytest_gpu 900 python -m pytest tests -m gpu -q -p no:cacheprovider -x --durations=8 ;;
smoke) leg smoke 300 python -c "import __graft_entry__ as g; g.smoke()" ;;
bench) leg bench_n1 600 python bench.py ;;
bench_strong) leg bench_n1_1e9 300 python bench.py --total-walkers-per-node 100000 --no-extras ;;
bench_hist) leg bench_n1_score_hist 200 python bench.py --tally score+hist --no-extras ;;
variants) leg variants_c5 150 python tools/k1s_variants.py c5; leg variants_c1 120 python tools/k1s_variants.py c1 ;;
ncu_launches) leg ncu_launches 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file "$out/${tag}_launches_bench_steps2.csv" python bench.py --steps 2 --warmup 3 --no-extras ;;
ncu_full) leg ncu_walk_score_full 240 ncu --set full --clock-control none --import-source on -k regex:walk_score -s 4 -c 1 \
    -f -o "$out/${tag}_walk_score" python bench.py --steps 2 --warmup 3 --no-extras ;;
ncu_c4d) leg ncu_c4d 240 ncu --set full --clock-control none --import-source on -k regex:walk_philox -s 2 -c 1 \
    -f -o "$out/${tag}_c4d_walk_philox" python tools/side_benches.py c4d ;;
esac
done
cat "$out/${tag}_legs.txt"
tail -5 "$out/${tag}_pytest_gpu.log" 2>/dev/null
tail -3 "$out/${tag}_smoke.log" "$out/${tag}_smoke.err" 2>/dev/null | cut -c1-400
cat "$out/${tag}_bench_n1.log" 2>/dev/null | cut -c1-1500
tail -5 "$out/${tag}_bench_n1.err" 2>/dev/null | cut -c1-600
