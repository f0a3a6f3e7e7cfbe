#!/bin/bash
# End-to-end step (pinned host -> device -> host) against the tiling of the streamed schedule.
#   DMLP_E2E_XPARTS  data parts      DMLP_E2E_QROWS  query rows per chunk
mkdir -p gpurun_out
: > gpurun_out/sweep_e2e.log
for cfg in "4 18944" "8 9472" "8 18944" "4 9472" "2 18944"; do
  set -- $cfg
  echo -n "xparts=$1 qrows=$2: " | tee -a gpurun_out/sweep_e2e.log
  DMLP_E2E_XPARTS=$1 DMLP_E2E_QROWS=$2 timeout 200 python bench.py --steps 5 --warmup 3 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('device', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['ms_per_step'],2), 'ms  flagged', d['flagged_queries'])" 2>&1 | tee -a gpurun_out/sweep_e2e.log
done
