#!/bin/bash
# usage: tools/gpuretry_n.sh <gpus> <timeout-seconds> '<command string>'   (gpurun --gpus N with the busy-retry of gpuretry.sh)
for k in $(seq 1 12); do
  /usr/local/graft/bin/gpurun --gpus "$1" --timeout "$2" -- "$3"; rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 150
done
exit 3
