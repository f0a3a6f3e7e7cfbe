#!/bin/bash
# Re-rank exact pass: query rows staged once per warp (DMLP_RERANK_DEDUP=1) vs once per pair.
mkdir -p gpurun_out
: > gpurun_out/sweep_rerank.log
DMLP_RERANK_DEDUP=1 timeout 200 python -m pytest tests/test_gpu_kernels.py -x -q > gpurun_out/rerank_tests.log 2>&1
echo "dedup tests rc=$? $(tail -1 gpurun_out/rerank_tests.log)" | tee -a gpurun_out/sweep_rerank.log
summ='import sys,json; d=json.loads(sys.stdin.read()); print(d["shape"], "flagged",d["flagged"], {k: round(v,2) for k,v in d["ms"].items()})'
for v in 1 0; do
  echo -n "dedup=$v " | tee -a gpurun_out/sweep_rerank.log
  DMLP_RERANK_DEDUP=$v timeout 120 python scripts/phase_times.py --reps 3 | python -c "$summ" 2>&1 | tee -a gpurun_out/sweep_rerank.log
done
