#!/bin/bash
# usage: tools/gpurun_retry.sh <timeout-seconds> '<command>'  -- retries while gpurun answers "busy" (exit 3), nothing is charged then
T=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$T" -- "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 60
done
exit 3
