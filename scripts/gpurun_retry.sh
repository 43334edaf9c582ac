#!/usr/bin/env bash
# Development helper: run a gpurun command, retrying while the pool answers "busy" (nothing is charged for those).
# usage: scripts/gpurun_retry.sh [--gpus N] TIMEOUT 'command'
GP=()
if [ "$1" = "--gpus" ]; then GP=(--gpus "$2"); shift 2; fi
T="$1"; shift
for i in $(seq 1 40); do
  out=$(/usr/local/graft/bin/gpurun "${GP[@]}" --timeout "$T" -- "$@" 2>&1)
  if echo "$out" | grep -q "status=transient"; then sleep 120; continue; fi
  echo "$out"; exit 0
done
echo "gave up: pool busy"; exit 3
