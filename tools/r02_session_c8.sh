#!/bin/bash
OUT=gpurun_out/r02c8; mkdir -p $OUT
timeout 1700 python tools/dev_cache_sampler3.py 60 6 4 2>&1 | tee $OUT/repeat.txt | cut -c1-400
