#!/usr/bin/env bash
# Retry a gpurun call while the pod answers "busy" (exit 3: nothing charged).
#   scripts/gpu_try.sh [--gpus N] <timeout_s> '<command>'
gp=()
if [ "$1" = "--gpus" ]; then gp=(--gpus "$2"); shift 2; fi
t=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "${gp[@]}" --timeout "$t" -- "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 45
done
exit 3
