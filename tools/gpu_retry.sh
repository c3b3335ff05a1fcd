#!/bin/bash
# usage: tools/gpu_retry.sh <timeout-seconds> '<command>' [logfile]
# Retries gpurun while the pod answers "no slot right now" (exit 3, nothing charged).
# Holds off while /tmp/repo.lock exists (the tree is being edited).
T=$1; CMD=$2; LOG=${3:-/tmp/gpurun_retry.log}
for i in $(seq 1 60); do
  while [ -e /tmp/repo.lock ]; do sleep 5; done
  /usr/local/graft/bin/gpurun ${GPUS:+--gpus $GPUS} --timeout "$T" -- "$CMD" > "$LOG" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 20
done
exit 3
