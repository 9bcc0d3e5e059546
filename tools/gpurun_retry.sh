#!/bin/bash
# local helper: retry a gpurun call until the pod has a slot (exit code 3 / "transient" = nothing charged)
# usage: tools/gpurun_retry.sh <timeout_s> <gpus> '<command>'
t=$1; g=$2; shift 2
for i in $(seq 1 40); do
  if [ "$g" = "1" ]; then out=$(gpurun --timeout $t -- "$@" 2>&1); else out=$(gpurun --gpus $g --timeout $t -- "$@" 2>&1); fi
  if echo "$out" | grep -q "status=transient\|status=busy\|rc=None"; then sleep 90; continue; fi
  echo "$out"; exit 0
done
echo "$out"; exit 3
