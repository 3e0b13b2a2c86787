#!/bin/bash
# gpurun with retries on "busy / no slot" (exit code 3, nothing charged). Usage: scripts/gpurun_retry.sh [gpurun options] -- 'command'
for attempt in $(seq 1 12); do
  /usr/local/graft/bin/gpurun "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  echo "[gpurun_retry] attempt $attempt answered busy; sleeping 120 s" >&2
  sleep 120
done
exit 3
