#!/bin/bash
# tools/gpurun_retry.sh <timeout-seconds> [--gpus N] -- '<command>': gpurun, retried while no box is free (exit 3).
T=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$T" "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  echo "[gpurun_retry] no box free (attempt $i), sleeping 90 s" >&2
  sleep 90
done
exit 3
