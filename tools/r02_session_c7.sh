#!/bin/bash
OUT=gpurun_out/r02c7; mkdir -p $OUT
timeout 1500 python tools/dev_cache_verify.py 60 5 2>&1 | tee $OUT/verify.txt | cut -c1-300
