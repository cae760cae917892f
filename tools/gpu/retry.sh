#!/bin/bash
# usage: tools/gpu/retry.sh <log> <gpurun args...>   — re-submits while the pod answers "busy" (exit 3 / transient)
log=$1; shift
for i in $(seq 1 40); do
  gpurun "$@" > "$log" 2>&1
  rc=$?
  if grep -q "status=transient" "$log" || [ $rc -eq 3 ]; then sleep 90; continue; fi
  exit $rc
done
exit 3
