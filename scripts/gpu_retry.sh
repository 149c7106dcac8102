#!/bin/bash
# usage: scripts/gpu_retry.sh <gpus> <outfile> <timeout-seconds> <command string>
# gpurun with retries while the pod answers "busy / transient" (nothing is charged for those).
N=$1; OUT=$2; TO=$3; shift 3
for i in $(seq 1 30); do
  if [ "$N" = 1 ]; then /usr/local/graft/bin/gpurun --timeout $TO -- "$@" > $OUT 2>&1
  else /usr/local/graft/bin/gpurun --gpus $N --timeout $TO -- "$@" > $OUT 2>&1; fi
  if ! grep -q "status=transient\|status=busy\|rc=3\b" $OUT; then break; fi
  sleep 100
done
