#!/bin/bash
# usage: tools/gpurun_retry.sh <timeout_s> <command...>   — retries while the pod answers "transient" (nothing charged)
T=$1; shift
for i in $(seq 1 12); do
  out=$(/usr/local/graft/bin/gpurun --timeout $T "$@" 2>&1)
  if echo "$out" | grep -q "status=transient"; then sleep 200; continue; fi
  echo "$out" | tail -60
  exit 0
done
echo "gave up: pod busy"
