#!/bin/bash
# Usage: [GPUS=2] tools/gpurun_retry.sh <timeout-seconds> '<command>'  -- retries while the pod answers "busy" (rc 3).
T=$1; shift
for k in $(seq 1 40); do
  /usr/local/graft/bin/gpurun ${GPUS:+--gpus $GPUS} --timeout "$T" -- "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
