#!/bin/bash
# Repeat a gpurun call while the pod answers "busy / transient" (exit code 3).  Usage: tools/gpuretry.sh <timeout-s> '<command>' [gpus]
# The last call's output is kept in gpurun_out/.gpuretry.log.
T=$1; CMD=$2; G=${3:-1}
mkdir -p gpurun_out
for i in $(seq 1 40); do
  if [ "$G" = 1 ]; then /usr/local/graft/bin/gpurun --timeout "$T" -- "$CMD" > gpurun_out/.gpuretry.log 2>&1
  else /usr/local/graft/bin/gpurun --gpus "$G" --timeout "$T" -- "$CMD" > gpurun_out/.gpuretry.log 2>&1; fi
  rc=$?
  if [ $rc -ne 3 ]; then break; fi
  sleep 60
done
tail -120 gpurun_out/.gpuretry.log
exit $rc
