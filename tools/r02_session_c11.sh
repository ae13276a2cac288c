#!/bin/bash
OUT=gpurun_out/r02c11; mkdir -p $OUT
timeout 900 python tools/dev_cache_stress2.py 4096 400 2>&1 | tee $OUT/stress2.txt | cut -c1-1600
