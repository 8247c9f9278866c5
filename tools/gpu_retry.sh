#!/bin/bash
# usage: tools/gpu_retry.sh <timeout> <command...>   -- retries gpurun while the pod answers busy (nothing charged)
T=$1; shift
for i in $(seq 1 40); do
  out=$(gpurun --timeout $T -- "$@" 2>&1)
  if echo "$out" | grep -q "status=transient\|retry in a few minutes\|no box"; then sleep 90; continue; fi
  echo "$out" | tail -120
  exit 0
done
echo "gave up: pod busy"
