#!/bin/bash
# local helper (build container): retries a gpurun call while the pod answers "busy / draining" (nothing charged)
# usage: tools/gpurun_retry.sh [gpurun options] -- 'command'
for i in $(seq 1 40); do
  out=$(/usr/local/graft/bin/gpurun "$@" 2>&1); rc=$?
  if echo "$out" | grep -q "status=transient\|nothing was charged — retry\|no box or slot"; then
    echo "[retry $i] busy; sleeping 90 s"; sleep 90; continue
  fi
  echo "$out" | tail -60; exit $rc
done
echo "gave up after 40 tries"; exit 3
