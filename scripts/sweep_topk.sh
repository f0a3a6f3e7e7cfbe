#!/bin/bash
# Sweep the work-decomposition knobs of dist_topk on the headline shape (run under gpurun).
#   DMLP_TOPK_GMC    cluster rows (CTA pairs) per chunk and wave
#   DMLP_TOPK_CHUNKS column ranges per query
#   DMLP_TOPK_SYNC   re-alignment period in data tiles (0 = off)
#   DMLP_TOPK_WIDE   256 x 512 tile per pair
mkdir -p gpurun_out
for cfg in "37 2 1 1" "24 3 1 1" "18 4 1 1" "12 6 1 1" "9 8 1 1" "18 4 2 1" "18 4 0 1" "9 8 0 1" "18 4 1 0"; do
  set -- $cfg
  echo -n "gmc=$1 chunks=$2 sync=$3 wide=$4: "
  DMLP_TOPK_GMC=$1 DMLP_TOPK_CHUNKS=$2 DMLP_TOPK_SYNC=$3 DMLP_TOPK_WIDE=$4 timeout 120 python scripts/phase_times.py --reps 2 | python -c "import sys,json; d=json.loads(sys.stdin.read()); p=d['plan']; print('grid',p['grid'],'gm',p['group_m'],'chunks',p['chunks'], round(d['ms']['dist_topk'],2),'ms', round(d['dist_topk_tflops'],1),'TF rerank', round(d['ms']['rerank'],2), 'flagged',d['flagged'])"
done 2>&1 | tee gpurun_out/sweep_topk.log
