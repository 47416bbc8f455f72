#!/bin/bash
# usage: scripts/gpu_retry.sh <tag> <timeout> '<command>' : retries gpurun while the pod answers busy (rc 3 / transient)
tag=$1; to=$2; cmd=$3
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$to" -- "$cmd" > gpurun_out/call_$tag.txt 2>&1
  if ! grep -q "status=transient\|nothing was charged" gpurun_out/call_$tag.txt; then break; fi
  sleep 120
done
echo done >> gpurun_out/call_$tag.txt
