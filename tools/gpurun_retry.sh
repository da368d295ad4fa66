#!/bin/bash
# usage: tools/gpurun_retry.sh <log> <gpurun args...>   - retries while the pod answers "no slot right now" (exit code 3, nothing charged)
log=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  echo "attempt $i rc=$rc $(date +%T)" >> "$log.attempts"
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
