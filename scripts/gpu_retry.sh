#!/bin/bash
# usage: scripts/gpu_retry.sh <timeout_s> [--gpus N] -- '<command>'   (retries while gpurun answers 3 = no box free)
T=$1; shift
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun --timeout "$T" "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
