#!/bin/bash
mkdir -p gpurun_out
for cfg in "37 2" "74 1" "18 4" "24 3" "12 6" "32 2" "64 1" "36 2"; do
  set -- $cfg
  echo -n "gmc=$1 chunks=$2: "
  DMLP_TOPK_GMC=$1 DMLP_TOPK_CHUNKS=$2 timeout 120 python scripts/phase_times.py --reps 2 | python -c "import sys,json; d=json.loads(sys.stdin.read()); p=d['plan']; print('grid',p['grid'],'gm',p['group_m'],'chunks',p['chunks'], round(d['ms']['dist_topk'],2),'ms', round(d['dist_topk_tflops'],1),'TF flagged',d['flagged'])"
done 2>&1 | tee gpurun_out/sweep_topk.log
