#!/bin/bash
OUT=gpurun_out/r02c16; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gp_factor_cache.py tests/test_gpu_parity.py -m gpu -x -q > $OUT/pytest.txt 2>&1; echo "rc=$?" >> $OUT/pytest.txt; tail -4 $OUT/pytest.txt
timeout 300 python tools/dev_small.py > $OUT/dev_small.jsonl 2>&1; cut -c1-120 $OUT/dev_small.jsonl
timeout 300 python tools/dev_cache.py 2>&1 | tee $OUT/dev_cache.jsonl | cut -c1-250
