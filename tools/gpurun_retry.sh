#!/bin/bash
# dev helper: retry a gpurun call while the pod answers "transient" (no slot); usage: gpurun_retry.sh <timeout> <command>
t=$1; shift
for i in $(seq 1 40); do
  out=$(/usr/local/graft/bin/gpurun --timeout $t -- "$@" 2>&1)
  if echo "$out" | grep -q "status=transient"; then sleep 90; continue; fi
  echo "$out" | tail -60
  exit 0
done
echo "gave up: no GPU slot"
