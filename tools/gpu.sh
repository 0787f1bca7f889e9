#!/bin/bash
# usage: tools/gpu.sh <logfile> <timeout_s> [--gpus N] -- '<command>'   (retries while the pod answers busy)
log=$1; shift; to=$1; shift
extra=()
while [ "$1" != "--" ]; do extra+=("$1"); shift; done; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$to" "${extra[@]}" -- "$1" > "$log" 2>&1
  rc=$?
  if grep -q "status=transient\|nothing was charged" "$log" || [ $rc -eq 3 ]; then sleep 45; continue; fi
  break
done
echo "gpu.sh done rc=$rc" >> "$log"
