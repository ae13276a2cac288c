#!/bin/bash
OUT=gpurun_out/r02c21; mkdir -p $OUT
KIMA_HOST_THREADS=16 timeout 1200 python tools/dev_host_repro.py CoRoT7 4096 60 9 1 2>&1 | cut -c1-200 | tee $OUT/host16.txt
