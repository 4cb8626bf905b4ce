#!/bin/bash
# gpurun with retries while the pod answers "busy" (exit 3: nothing charged).  Usage: tools/gpurun_retry.sh <timeout_s> [--gpus N] '<command>'
T=$1; shift
EXTRA=""
if [ "$1" == "--gpus" ]; then EXTRA="--gpus $2"; shift 2; fi
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun --timeout $T $EXTRA -- "$1"
  rc=$?
  if [ $rc -ne 3 ] && [ $rc -ne 2 ]; then exit $rc; fi
  sleep 90
done
exit 3
