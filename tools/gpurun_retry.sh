#!/bin/bash
# tools/gpurun_retry.sh <log> <gpurun args...>: retries a gpurun call while the pod answers "busy" (exit code 3, nothing charged)
LOG=$1; shift
for attempt in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  rc=$?
  [ $rc -ne 3 ] && exit $rc
  sleep 120
done
exit 3
