#!/bin/bash
# factor cache: its tests, then the parity tests of the GP paths and the sampler tests
OUT=gpurun_out/${1:-r02cache}; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gp_factor_cache.py -m gpu -x -q > $OUT/pytest_cache.txt 2>&1; echo "rc=$?" >> $OUT/pytest_cache.txt; tail -25 $OUT/pytest_cache.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $OUT/pytest_parity.txt 2>&1; echo "rc=$?" >> $OUT/pytest_parity.txt; tail -3 $OUT/pytest_parity.txt
timeout 300 python tools/dev_small.py > $OUT/dev_small.jsonl 2>&1; cut -c1-120 $OUT/dev_small.jsonl
