#!/bin/bash
# usage: tools/gpu_retry_n.sh <gpus> <timeout_s> '<command>'  -- like gpu_retry.sh with --gpus N
N=$1; T=$2; shift 2
for i in $(seq 1 12); do
  out=$(gpurun --gpus "$N" --timeout "$T" -- "$@" 2>&1); rc=$?
  if [ $rc -eq 3 ] || echo "$out" | grep -q "status=transient"; then sleep 120; continue; fi
  echo "$out"; exit $rc
done
echo "$out"; exit 3
