#!/bin/bash
# Re-rank exact pass: chunk width / ring depth variants (DMLP_RERANK_VARIANT; run under gpurun).
#   0: 32 columns x 2 stages   1: 16 x 2   2: 16 x 3   3: 32 x 3   4: 16 x 4
mkdir -p gpurun_out
: > gpurun_out/sweep_rerank.log
summ='import sys,json; d=json.loads(sys.stdin.read()); print(d["shape"], "flagged",d["flagged"], {k: round(v,2) for k,v in d["ms"].items()})'
for v in 0 1 2 3 4; do
  echo -n "variant=$v " | tee -a gpurun_out/sweep_rerank.log
  DMLP_RERANK_VARIANT=$v timeout 120 python scripts/phase_times.py --reps 2 | python -c "$summ" 2>&1 | tee -a gpurun_out/sweep_rerank.log
done
for v in 1 2 4; do
  DMLP_RERANK_VARIANT=$v timeout 200 python -m pytest tests/test_gpu_kernels.py -x -q -k "knn_local or rerank or model" > gpurun_out/rerank_tests_$v.log 2>&1
  echo "variant=$v tests rc=$? $(tail -1 gpurun_out/rerank_tests_$v.log)" | tee -a gpurun_out/sweep_rerank.log
done
