#!/bin/bash
# usage: tools/gpurun_retry.sh <timeout s> <gpus> '<command>'  - retries while the pod answers "busy" (nothing is charged then)
mkdir -p /root/repo/gpurun_out
for i in $(seq 1 40); do
  out=$(/usr/local/graft/bin/gpurun --timeout "$1" --gpus "$2" -- "$3" 2>&1)
  echo "$out" > /root/repo/gpurun_out/.last_retry_out.txt
  if echo "$out" | grep -q "status=transient"; then sleep 120; continue; fi
  break
done
echo "$out" | tail -60
