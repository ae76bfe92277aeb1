#!/bin/bash
# usage: tools/gpurun_retry.sh [gpurun options] -- 'command' : re-submits while the pod answers "busy / draining" (nothing charged)
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun "$@" > /tmp/gpurun_retry.out 2>&1
  rc=$?
  if grep -q "status=transient" /tmp/gpurun_retry.out || [ $rc -eq 3 ]; then
    echo "[retry $i] busy" ; sleep 150 ; continue
  fi
  cat /tmp/gpurun_retry.out
  exit $rc
done
cat /tmp/gpurun_retry.out
exit 3
