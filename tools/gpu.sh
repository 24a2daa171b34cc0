#!/bin/bash
# usage: [GPUS=N] tools/gpu.sh <timeout-seconds> <logfile> <command...>   — retries while the pod answers busy (exit 3)
T=$1; LOG=$2; shift 2
G=""; [ -n "$GPUS" ] && G="--gpus $GPUS"
mkdir -p gpurun_out
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun $G --timeout $T -- "$@" > "$LOG" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc" >> "$LOG"; exit $rc; fi
  sleep 45
done
echo "gave up" >> "$LOG"
