#!/bin/bash
# usage: [GPUS=2] tools/grun.sh <tag> <timeout_s> <command...>   — one gpurun call with retries while the pod answers "busy" (rc 3)
tag=$1; to=$2; shift 2
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun ${GPUS:+--gpus $GPUS} --timeout "$to" -- "$@" > gpurun_out/${tag}_call.log 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
