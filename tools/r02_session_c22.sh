#!/bin/bash
OUT=gpurun_out/r02c22; mkdir -p $OUT
echo "== shared stream"; KIMA_HOST_THREADS=16 timeout 1200 python tools/dev_host_repro.py CoRoT7 4096 60 14 1 2>&1 | cut -c1-700 | tee $OUT/shared.txt
