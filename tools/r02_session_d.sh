#!/bin/bash
# r02 session d: gps_kernel schedule variants (isolated diag scheduler, U look-ahead, tile-wise start of block column 0)
OUT=gpurun_out/r02d; mkdir -p $OUT
for v in base iso look tile looktile all; do
  KIMA_CUDA_LIB=$PWD/variants/gps_$v.so timeout 120 python tools/dev_small_variants.py >> $OUT/variants.jsonl 2>&1
done
cat $OUT/variants.jsonl
KIMA_CUDA_LIB=$PWD/variants/gps_allprof.so timeout 120 python tools/dev_small_prof.py 177 1184 > $OUT/gps_prof_all_177.log 2>&1; cat $OUT/gps_prof_all_177.log
timeout 600 python -m pytest tests/test_predictive.py -m gpu -x -q > $OUT/pytest_predictive.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_predictive.log
tail -5 $OUT/pytest_predictive.log
