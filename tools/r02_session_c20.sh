#!/bin/bash
OUT=gpurun_out/r02c20; mkdir -p $OUT
for t in 2 4 16; do
  echo "== KIMA_HOST_THREADS=$t, cache on"
  KIMA_HOST_THREADS=$t timeout 900 python tools/dev_host_repro.py CoRoT7 4096 60 5 1 2>&1 | cut -c1-200
done 2>&1 | tee $OUT/host_threads.txt
