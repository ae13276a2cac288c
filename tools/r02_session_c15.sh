#!/bin/bash
OUT=gpurun_out/r02c15; mkdir -p $OUT
for v in "$@"; do KIMA_CUDA_LIB=variants/$v.so timeout 200 python tools/dev_small_mix.py 2>&1 | grep "^{" | tee -a $OUT/mix.jsonl; done
