#!/bin/bash
# usage: tools/gpu_retry.sh <timeout_s> '<command>'   -- retries gpurun while the pod answers busy / transient (exit 3 or "transient")
T=$1; shift
for i in $(seq 1 30); do
  out=$(gpurun --timeout "$T" -- "$@" 2>&1); rc=$?
  if echo "$out" | grep -q "status=transient\|no box\|busy" && ! echo "$out" | grep -q "status=ok"; then
    if [ $rc -eq 3 ] || echo "$out" | grep -q "status=transient"; then sleep 90; continue; fi
  fi
  echo "$out"; exit $rc
done
echo "$out"; exit 3
