#!/bin/bash
# usage: tools/gpurun_retry.sh <out-file> <gpurun args...>   -- repeats a gpurun call while the pod answers busy / transient
out=$1; shift
for i in $(seq 1 12); do
  /usr/local/graft/bin/gpurun "$@" > "$out" 2>&1
  if grep -q "status=transient\|status=busy\|nothing was charged" "$out"; then sleep 150; else break; fi
done
tail -40 "$out" | cut -c1-240
