#!/bin/bash
# usage: [GPUS=N] tools/gpurun_retry.sh <timeout seconds> <command string>   — retries while the pool answers "busy" (exit 3)
t=$1; shift
for i in $(seq 40); do
  if [ -n "$GPUS" ]; then /usr/local/graft/bin/gpurun --gpus $GPUS --timeout $t -- "$@"; else /usr/local/graft/bin/gpurun --timeout $t -- "$@"; fi
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
