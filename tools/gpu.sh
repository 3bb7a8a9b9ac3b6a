#!/bin/bash
# tools/gpu.sh TIMEOUT 'command' [extra gpurun args] -- gpurun with retries while the pod answers "busy"
t=$1; shift; cmd=$1; shift
for i in 1 2 3 4 5 6 7 8 9 10 11 12; do
  out=$(/usr/local/graft/bin/gpurun "$@" --timeout "$t" -- "$cmd" 2>&1)
  if echo "$out" | grep -q "status=transient"; then sleep 75; continue; fi
  echo "$out"; exit 0
done
echo "$out"; exit 3
