#!/bin/bash
OUT=gpurun_out/r02c9; mkdir -p $OUT
timeout 1700 python tools/dev_cache_verify2.py 60 6 2 2>&1 | tee $OUT/verify2.txt | cut -c1-420
