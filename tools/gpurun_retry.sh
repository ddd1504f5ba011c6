#!/bin/bash
# usage: tools/gpurun_retry.sh [gpurun options] -- <command>   — retries while the pod answers "transient" (no slot free)
for attempt in $(seq 1 30); do
  out=$(/usr/local/graft/bin/gpurun "$@" 2>&1)
  if echo "$out" | grep -q "status=transient\|status=refused"; then
    echo "[retry $attempt] $(echo "$out" | grep -m1 "status=")" >&2
    sleep 100
    continue
  fi
  echo "$out"
  exit 0
done
echo "$out"
exit 3
