#!/bin/bash
# Developer tool (build container): gpurun with retries while the pod answers "no box free" (exit 3).
# usage: tools/gpurun_retry.sh LOGFILE TIMEOUT_S 'command' [gpurun options]
LOG=$1; T=$2; CMD=$3; shift 3
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" --timeout "$T" -- "$CMD" > "$LOG" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc (try $i)" >> "$LOG"; exit $rc; fi
  sleep 60
done
exit 3
