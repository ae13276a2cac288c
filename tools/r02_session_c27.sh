#!/bin/bash
OUT=gpurun_out/r02c27; mkdir -p $OUT
echo "== fixed, cache off"; KIMA_HOST_THREADS=16 timeout 900 python tools/dev_host_repro.py CoRoT7 4096 60 24 0 2>&1 | cut -c1-300 | tee $OUT/off.txt
echo "== fixed, cache on"; KIMA_HOST_THREADS=16 timeout 900 python tools/dev_host_repro.py CoRoT7 4096 60 24 1 2>&1 | cut -c1-300 | tee $OUT/on.txt
