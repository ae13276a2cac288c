#!/bin/bash
OUT=gpurun_out/r02c25; mkdir -p $OUT
echo "== trace"; KIMA_HOST_THREADS=16 timeout 1500 python tools/dev_host_repro.py CoRoT7 4096 60 22 0 2>&1 | cut -c1-400 | tee $OUT/trace.txt
