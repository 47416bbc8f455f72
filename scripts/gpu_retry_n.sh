#!/bin/bash
# usage: scripts/gpu_retry_n.sh <tag> <ngpus> <timeout> '<command>' : like gpu_retry.sh for a multi-GPU box
tag=$1; n=$2; to=$3; cmd=$4
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --gpus "$n" --timeout "$to" -- "$cmd" > gpurun_out/call_$tag.txt 2>&1
  if ! grep -q "status=transient\|nothing was charged" gpurun_out/call_$tag.txt; then break; fi
  sleep 120
done
echo done >> gpurun_out/call_$tag.txt
