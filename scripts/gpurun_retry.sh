#!/bin/bash
# usage: scripts/gpurun_retry.sh <tries> <gpurun args...>   -- retries while the pod answers "busy" (rc 3)
tries=$1; shift
for i in $(seq 1 "$tries"); do
  /usr/local/graft/bin/gpurun "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
