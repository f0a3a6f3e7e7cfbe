#!/bin/bash
mkdir -p gpurun_out
for cfg in "128 37 2" "64 37 2" "32 37 2" "16 37 2" "8 37 2" "32 18 4" "16 18 4" "16 12 6"; do
  set -- $cfg
  echo -n "sync_kb=$1 gmc=$2 chunks=$3: "
  DMLP_TOPK_SYNC=$1 DMLP_TOPK_GMC=$2 DMLP_TOPK_CHUNKS=$3 timeout 120 python scripts/phase_times.py --reps 2 | python -c "import sys,json; d=json.loads(sys.stdin.read()); p=d['plan']; print('grid',p['grid'],'gm',p['group_m'],'chunks',p['chunks'], round(d['ms']['dist_topk'],2),'ms', round(d['dist_topk_tflops'],1),'TF flagged',d['flagged'])"
done 2>&1 | tee gpurun_out/sweep_topk.log
