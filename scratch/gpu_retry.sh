#!/bin/bash
# retries a gpurun call while the pod answers "busy" (exit 3): scratch/gpu_retry.sh <timeout> <command...>
t=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$t" -- "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
