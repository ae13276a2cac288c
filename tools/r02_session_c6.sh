#!/bin/bash
OUT=gpurun_out/r02c6; mkdir -p $OUT
timeout 1200 python tools/dev_cache_stress.py 177 4096 6000 2>&1 | tee $OUT/stress.txt | cut -c1-1500
