#!/bin/bash
# gpurun with retries while the pod answers "busy / draining" (exit code 3: nothing charged)
# usage: [GPURUN_ARGS="--gpus 2"] tools/gpurun_retry.sh <timeout_s> <command...>
t=$1; shift
for try in $(seq 1 40); do
  /usr/local/graft/bin/gpurun ${GPURUN_ARGS:-} --timeout "$t" -- "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  echo "[retry $try] pod busy, sleeping 45 s"
  sleep 45
done
exit 3
