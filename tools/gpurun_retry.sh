#!/bin/bash
# gpurun with retries while the pod answers "no box or slot free" (exit code 3; nothing is charged for those).
# usage: tools/gpurun_retry.sh [gpurun options] -- '<command>'
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 60
done
exit 3
