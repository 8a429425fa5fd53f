#!/bin/bash
# usage: gpu_retry.sh <gpurun-timeout> <script> [extra gpurun args]; retries while the pod answers "busy" (exit 3)
T=$1; S=$2; shift 2
for i in $(seq 1 60); do
  gpurun --timeout $T "$@" -- "bash $S"; rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 100
done
exit 3
