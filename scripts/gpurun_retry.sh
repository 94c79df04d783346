#!/bin/bash
# usage: [GPUS=N] scripts/gpurun_retry.sh <timeout-seconds> <log-file> '<command>'   (development helper)
# Retries while the pod answers "busy / transient" (exit 3, nothing charged); stops on any other outcome.
T=$1; LOG=$2; shift 2
G=""; [ -n "$GPUS" ] && G="--gpus $GPUS"
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun $G --timeout "$T" -- "$@" > "$LOG" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 60
done
exit 3
