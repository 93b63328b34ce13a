#!/bin/bash
# usage: tools/gpurun_retry.sh <log> [gpurun args...] -- '<command>'
# Retries while gpurun answers "no box / slot free right now" (exit code 3; nothing is charged for those).
LOG="$1"; shift
for attempt in $(seq 1 20); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" "$LOG"; then exit $rc; fi
  sleep 150
done
exit 3
