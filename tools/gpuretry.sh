#!/bin/bash
# usage: tools/gpuretry.sh <log> <gpurun args...>   -- retries while the pod answers "transient"/busy (nothing charged)
log=$1; shift
for i in $(seq 1 30); do
  gpurun "$@" > $log 2>&1
  if grep -q "status=transient\|no box\|busy" $log && ! grep -q "status=ok" $log; then sleep 100; continue; fi
  break
done
