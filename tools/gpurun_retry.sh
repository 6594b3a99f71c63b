#!/bin/bash
# gpurun with retries on "busy / no slot" answers (exit 3 or status=transient; nothing is charged for those).
#   tools/gpurun_retry.sh <timeout_s> '<command>' [extra gpurun args...]      log -> stdout
t=$1; cmd=$2; shift 2
for attempt in 1 2 3 4 5 6 7 8 9 10; do
  out=$(/usr/local/graft/bin/gpurun "$@" --timeout "$t" -- "$cmd" 2>&1); rc=$?
  echo "$out"
  if [ $rc -ne 3 ] && ! echo "$out" | grep -q "status=transient"; then exit $rc; fi
  echo "[gpurun_retry] attempt $attempt answered busy; sleeping 150 s"
  sleep 150
done
exit 3
