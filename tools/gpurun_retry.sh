#!/bin/bash
# Local helper: retry a gpurun call while the pod answers "busy" (exit code 3), every 2 minutes, at most 12 times.
#   tools/gpurun_retry.sh <timeout seconds> '<command>'
for i in $(seq 1 12); do
  /usr/local/graft/bin/gpurun --timeout "$1" -- "$2"
  rc=$?
  [ $rc -ne 3 ] && exit $rc
  sleep 120
done
exit 3
