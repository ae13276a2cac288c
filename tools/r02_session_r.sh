#!/bin/bash
# r02 session r: gps_kernel schedule variants, one process per library
OUT=gpurun_out/r02r; mkdir -p $OUT
for v in "$@"; do
  KIMA_CUDA_LIB=variants/$v.so timeout 200 python tools/dev_small_variants.py >> $OUT/variants.jsonl 2>&1
done
cat $OUT/variants.jsonl
