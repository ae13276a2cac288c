#!/bin/bash
OUT=gpurun_out/r02c23; mkdir -p $OUT
echo "== control: cache off"; KIMA_HOST_THREADS=16 timeout 1500 python tools/dev_host_repro.py CoRoT7 4096 60 20 0 2>&1 | cut -c1-700 | tee $OUT/off.txt
