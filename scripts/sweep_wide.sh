#!/bin/bash
# A/B of the wide (256 x 512 per pair) filter kernel and of the default one: tests, wait profile,
# phase times on the headline, mid and short-row shapes (run under gpurun).
mkdir -p gpurun_out
: > gpurun_out/sweep_wide.log
for w in 0 1; do
  DMLP_TOPK_WIDE=$w timeout 420 python -m pytest tests/test_gpu_kernels.py -x -q > gpurun_out/wide_tests_$w.log 2>&1
  echo "wide=$w tests rc=$? $(tail -1 gpurun_out/wide_tests_$w.log)" | tee -a gpurun_out/sweep_wide.log
done
for w in 0 1; do DMLP_TOPK_WIDE=$w timeout 120 python scripts/topk_prof.py; done 2>&1 | tee -a gpurun_out/sweep_wide.log
summ='import sys,json; d=json.loads(sys.stdin.read()); p=d["plan"]; print(d["shape"], "grid",p["grid"],"gm",p["group_m"],"chunks",p["chunks"], round(d["dist_topk_tflops"],1),"TF flagged",d["flagged"], {k: round(v,2) for k,v in d["ms"].items()})'
for w in 0 1; do
  echo "wide=$w" | tee -a gpurun_out/sweep_wide.log
  DMLP_TOPK_WIDE=$w timeout 120 python scripts/phase_times.py --reps 3 | python -c "$summ" 2>&1 | tee -a gpurun_out/sweep_wide.log
  DMLP_TOPK_WIDE=$w timeout 120 python scripts/phase_times.py --reps 3 --q 16384 --n 262144 --a 1024 | python -c "$summ" 2>&1 | tee -a gpurun_out/sweep_wide.log
  DMLP_TOPK_WIDE=$w timeout 120 python scripts/phase_times.py --reps 3 --q 10000 --n 1000000 --a 64 | python -c "$summ" 2>&1 | tee -a gpurun_out/sweep_wide.log
done
