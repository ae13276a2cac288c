#!/bin/bash
# Retry a gpurun call while the pod answers "transient" (nothing charged): tools/gpu_retry.sh <timeout> <command...>
T=$1; shift
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun --timeout $T -- "$@" > gpurun_out/.retry.log 2>&1
  rc=$?
  if grep -q "status=transient" gpurun_out/.retry.log || [ $rc -eq 3 ]; then sleep 45; continue; fi
  break
done
tail -40 gpurun_out/.retry.log
echo "gpurun rc=$rc after $i tries"
