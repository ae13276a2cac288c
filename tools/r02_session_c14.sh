#!/bin/bash
OUT=gpurun_out/r02c14; mkdir -p $OUT
KIMA_CUDA_LIB=variants/prof.so timeout 300 python tools/dev_small_timeline.py 2>&1 | grep "^{" | tee $OUT/timeline.jsonl
