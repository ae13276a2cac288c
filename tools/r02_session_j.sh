#!/bin/bash
# r02 session j: gps_kernel with look-ahead U by the idle workers
OUT=gpurun_out/r02j; mkdir -p $OUT
timeout 300 python tools/dev_small.py > $OUT/dev_small.jsonl 2>&1; cat $OUT/dev_small.jsonl
KIMA_CUDA_LIB=$PWD/variants/gps_ulookprof.so timeout 120 python tools/dev_small_prof.py 177 1184 > $OUT/gps_prof_177.log 2>&1; cat $OUT/gps_prof_177.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest.log; tail -3 $OUT/pytest.log
