#!/bin/bash
# usage: tools/gpurun_retry.sh [gpurun options] -- '<command>'   retries while the pod answers "busy" (exit 3), at most ~40 min
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun "$@"; rc=$?
  [ $rc -ne 3 ] && exit $rc
  sleep 100
done
exit 3
