#!/bin/bash
# usage: tools/gpurun_retry.sh [gpurun args...] -- 'cmd'   : retries while the pod answers busy (exit 3 / transient)
for i in $(seq 1 12); do
  out=$(gpurun "$@" 2>&1); rc=$?
  if echo "$out" | grep -q "status=transient"; then sleep 120; continue; fi
  echo "$out"; exit $rc
done
echo "$out"; exit 3
