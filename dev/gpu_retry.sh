#!/bin/bash
# usage: dev/gpu_retry.sh <timeout-seconds> <log> [--gpus N] -- <command>; retries while the pod answers "busy" (nothing charged)
T=$1; LOG=$2; shift 2
EXTRA=""
if [ "$1" = "--gpus" ]; then EXTRA="--gpus $2"; shift 2; fi
shift  # the --
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $T $EXTRA -- "$@" > $LOG 2>&1
  rc=$?
  if grep -q "status=transient\|nothing was charged" $LOG; then sleep 90; continue; fi
  break
done
echo "gpu_retry: done rc=$rc after $i tries" >> $LOG
