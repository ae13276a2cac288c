#!/bin/bash
OUT=gpurun_out/r02c10; mkdir -p $OUT
timeout 800 python tools/dev_host_repro.py 51Peg 4096 200 5 2>&1 | tee $OUT/repro_51peg.txt | cut -c1-300
KIMA_HOST_THREADS=1 timeout 900 python tools/dev_host_repro.py CoRoT7 4096 40 4 1 2>&1 | tee $OUT/repro_corot_1thread.txt | cut -c1-300
