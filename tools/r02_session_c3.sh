#!/bin/bash
OUT=gpurun_out/r02c3; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gp_factor_cache.py -m gpu -x -q > $OUT/pytest_cache.txt 2>&1; echo "rc=$?" >> $OUT/pytest_cache.txt; tail -25 $OUT/pytest_cache.txt
timeout 300 python tools/dev_cache.py 2>&1 | tee $OUT/dev_cache.jsonl | cut -c1-260
timeout 300 python tools/dev_small.py > $OUT/dev_small.jsonl 2>&1; cut -c1-120 $OUT/dev_small.jsonl
