#!/bin/bash
# gpurun with retries while the pod answers busy (exit 3 / "transient"): tools/gpurun_retry.sh [gpurun args] -- 'cmd'
for i in $(seq 1 12); do
  out=$(/usr/local/graft/bin/gpurun "$@" 2>&1); rc=$?
  if echo "$out" | grep -q "status=transient"; then sleep 150; continue; fi
  echo "$out"; exit $rc
done
echo "$out"; exit 3
