#!/bin/bash
# usage: [GPURUN_OPTS="--gpus 2"] tools/gpu_retry.sh <timeout> <command...>   -- retries gpurun while the pod answers busy (nothing charged)
T=$1; shift
for i in $(seq 1 40); do
  out=$(gpurun $GPURUN_OPTS --timeout $T -- "$@" 2>&1)
  if echo "$out" | grep -q "status=transient\|retry in a few minutes\|no box\|status=busy"; then sleep 90; continue; fi
  echo "$out" | tail -120
  exit 0
done
echo "gave up: pod busy"
