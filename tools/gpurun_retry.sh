#!/bin/bash
# tools/gpurun_retry.sh LOG [gpurun args...] : retries while the pod answers "busy / draining" (rc 3 / transient)
LOG=$1; shift
for i in 1 2 3 4 5 6 7 8 9 10 11 12; do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  if ! grep -q "status=transient\|nothing was charged" "$LOG"; then break; fi
  sleep 120
done
echo done >> "$LOG"
