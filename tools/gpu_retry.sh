#!/bin/bash
# Helper for the build sessions: retry a gpurun call while the pod answers "busy" (exit code 3).
# usage: tools/gpu_retry.sh <timeout-seconds> <gpus> '<command>'
T=$1; G=$2; shift 2
for i in $(seq 1 40); do
  if [ "$G" = "1" ]; then gpurun --timeout $T -- "$@"; else gpurun --gpus $G --timeout $T -- "$@"; fi
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
