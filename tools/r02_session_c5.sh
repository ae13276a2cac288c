#!/bin/bash
OUT=gpurun_out/r02c5; mkdir -p $OUT
timeout 900 python tools/dev_cache_sampler2.py 4096 100 60 2>&1 | tee $OUT/repro.txt
