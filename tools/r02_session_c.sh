#!/bin/bash
# r02 session c: predictive standard deviation (kc_predict_std) parity + in-kernel phase profile of gps_kernel
OUT=gpurun_out/r02c; mkdir -p $OUT
timeout 600 python -m pytest tests/test_predictive.py -m gpu -x -q > $OUT/pytest_predictive.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_predictive.log
tail -15 $OUT/pytest_predictive.log
KIMA_CUDA_LIB=$PWD/variants/prof.so timeout 120 python tools/dev_small_prof.py 177 1184 > $OUT/gps_prof_177.log 2>&1; cat $OUT/gps_prof_177.log
KIMA_CUDA_LIB=$PWD/variants/prof.so timeout 120 python tools/dev_small_prof.py 100 1184 > $OUT/gps_prof_100.log 2>&1; cat $OUT/gps_prof_100.log
