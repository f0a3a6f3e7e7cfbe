#!/usr/bin/env bash
# The reference harness's four configurations on 1, 2, 4 and 8 GPUs (BASELINE.json's metric:
# "bench_1-4 wall time"), plus a larger synthetic problem on the same four layouts.
#   scripts/bench_table.sh [gpu counts...]      -> gpurun_out/run_bench_table.jsonl
cd "$(dirname "$0")/.."
mkdir -p gpurun_out outputs
out=gpurun_out/run_bench_table.jsonl
: > "$out"
have=$(nvidia-smi -L 2>/dev/null | wc -l)
counts=${*:-1 2 4 8}
for g in $counts; do
  [ "$g" -le "$have" ] || continue
  for c in 1 2 3 4; do
    log=$(DMLP_NGPU=$g ./run_bench.sh "$c" 2>&1)
    b=$(echo "$log" | grep -oE 'Benchmark time: [0-9]+' | grep -oE '[0-9]+')
    e=$(echo "$log" | grep -oE '"time_ms": [0-9.]+' | grep -oE '[0-9.]+$')
    be=$(echo "$log" | grep -oE '"backend": "[^"]*"' | cut -d'"' -f4)
    r=$(echo "$log" | grep -oE 'Results: +[a-zA-Z]+' | awk '{print $2}')
    echo "{\"what\": \"run_bench\", \"config\": $c, \"gpus\": $g, \"baseline_ms\": ${b:-null}, \"engine_ms\": ${e:-null}, \"backend\": \"$be\", \"results\": \"$r\"}" | tee -a "$out"
  done
  # a problem large enough for a B200 (no text: generated in memory), on the four layouts
  S="--synth 1000000 32768 256 -100 100 8 16 10 7"
  for grid in "${g}x1" "1x${g}" $([ "$g" -ge 4 ] && echo "2x$((g / 2))"); do
    res=$(DMLP_NGPU=$g DMLP_GRID=$grid DMLP_BACKEND=gpu ./engine $S --json 2> gpurun_out/tmp_table.err | sort | md5sum | cut -c1-12)
    e=$(grep -oE '"time_ms": [0-9.]+' gpurun_out/tmp_table.err | grep -oE '[0-9.]+$')
    be=$(grep -oE '"backend": "[^"]*"' gpurun_out/tmp_table.err | cut -d'"' -f4)
    echo "{\"what\": \"synth 1000000 x 32768 x 256\", \"grid\": \"$grid\", \"gpus\": $g, \"engine_ms\": ${e:-null}, \"backend\": \"$be\", \"md5\": \"$res\"}" | tee -a "$out"
    [ "$g" -eq 1 ] && break
  done
done
