#!/bin/bash
# usage: tools/gpuretry.sh <timeout-seconds> '<command string>'
# Runs one gpurun call; exit code 3 (no box / slot free right now, nothing charged) is retried every two minutes.
for k in $(seq 1 20); do
  /usr/local/graft/bin/gpurun --timeout "$1" -- "$2"; rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
