#!/bin/bash
# usage: tools/gpurun_retry.sh LOGFILE [gpurun args...] -- 'command'
# Retries while the pod answers busy (exit 3), up to ~25 minutes.
LOG=$1; shift
for i in $(seq 1 12); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" "$LOG"; then break; fi
  sleep 120
done
echo "gpurun_retry finished rc=$rc" >> "$LOG"
