#!/bin/bash
OUT=gpurun_out/r02c12; mkdir -p $OUT
timeout 1200 python tools/dev_cache_sampler4.py 30 6 2>&1 | tee $OUT/device_repro.txt | cut -c1-300
