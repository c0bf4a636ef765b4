#!/bin/bash
# usage: tools/gpurun_retry.sh <log> <gpurun args...>   -- retries while gpurun answers 3 (no box / slot free right now)
LOG=$1; shift
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1; rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc $rc (attempt $i)" >> "$LOG"; exit $rc; fi
  sleep 120
done
echo "gave up" >> "$LOG"
