#!/bin/bash
# usage: tools/gpurun_retry.sh <timeout_s> <log> <gpus> <command...>   retries while gpurun answers "busy" (exit 3)
t=$1; log=$2; g=$3; shift 3
for i in $(seq 1 40); do
  if [ "$g" = "1" ]; then /usr/local/graft/bin/gpurun --timeout $t -- "$@" > $log 2>&1; rc=$?
  else /usr/local/graft/bin/gpurun --gpus $g --timeout $t -- "$@" > $log 2>&1; rc=$?; fi
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
