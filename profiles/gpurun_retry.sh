#!/bin/bash
# gpurun with retries while the pod answers "transient" (no slot free; nothing charged).  usage: gpurun_retry.sh LOG [gpurun args...]
LOG=$1; shift
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  grep -q "status=transient" "$LOG" || break
  sleep 60
done
tail -60 "$LOG"
