#!/bin/bash
# gpurun with retries while the pod answers busy (exit 3 / status=transient).  usage: tools/grun.sh LOG [gpurun args...]
log=$1; shift
for i in $(seq 1 ${RETRIES:-30}); do
  gpurun "$@" > "$log" 2>&1; rc=$?
  if grep -q "status=transient\|no box or slot" "$log" || [ $rc -eq 3 ]; then sleep ${WAIT:-90}; continue; fi
  break
done
echo "grun done rc=$rc" >> "$log"
