#!/bin/bash
# Dev tool: run a gpurun call, retrying while the pod answers "busy" (exit code 3: nothing charged).
#   tools/gpurun_retry.sh [--gpus N] --timeout S -- '<command>'
for attempt in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  echo "[gpurun_retry] busy (attempt $attempt), retrying in 150 s" >&2
  sleep 150
done
exit 3
