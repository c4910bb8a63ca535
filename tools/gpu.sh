#!/bin/bash
# usage: [GPUS=N] tools/gpu.sh <timeout-seconds> '<command>'   -- gpurun with retries while the pod has no free slot (exit 3)
t=$1; shift
g=""; [ -n "$GPUS" ] && g="--gpus $GPUS"
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun $g --timeout "$t" -- "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 45
done
exit 3
