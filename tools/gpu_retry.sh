#!/bin/bash
# usage: tools/gpu_retry.sh TIMEOUT [--gpus N] -- 'command'  — gpurun, retried while the pod answers busy / transient
T=$1; shift
OUT=$(mktemp)
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout $T "$@" > $OUT 2>&1
  rc=$?
  if grep -q "status=transient\|status=busy" $OUT || [ $rc -eq 3 ]; then sleep 90; continue; fi
  break
done
grep -v "^\[gpurun\] sending" $OUT
rm -f $OUT
