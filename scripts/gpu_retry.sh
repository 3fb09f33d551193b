#!/bin/bash
# Retries a gpurun call while the pod answers "busy" (exit 3).  Usage: scripts/gpu_retry.sh <timeout> '<command>' [--gpus N]
T=$1; CMD=$2; shift 2
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" --timeout $T -- "$CMD"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 45
done
exit 3
