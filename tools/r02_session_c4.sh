#!/bin/bash
OUT=gpurun_out/r02c4; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gp_factor_cache.py -m gpu -x -q > $OUT/pytest_cache.txt 2>&1; echo "rc=$?" >> $OUT/pytest_cache.txt; tail -3 $OUT/pytest_cache.txt
timeout 1500 python -m pytest tests/test_device_proposals.py tests/test_host_cpp.py tests/test_sampler_logic.py -m gpu -x -q > $OUT/pytest_sampler.txt 2>&1; echo "rc=$?" >> $OUT/pytest_sampler.txt; tail -5 $OUT/pytest_sampler.txt
timeout 1500 python tools/dev_cache_sampler.py 2>&1 | tee $OUT/cache_sampler.jsonl | cut -c1-600
