#!/bin/bash
# usage: tools/gpuN.sh <ngpus> <timeout-seconds> '<command>'   -- multi-GPU gpurun with retries while the pod is busy
N=$1; T=$2; shift 2
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --gpus "$N" --timeout "$T" -- "$@"
  rc=$?
  [ $rc -ne 3 ] && exit $rc
  sleep 120
done
exit 3
