#!/bin/bash
OUT=gpurun_out/r02final4; mkdir -p $OUT
timeout 150 python -m pytest tests/test_gp_factor_cache.py -m gpu -q -k "chunked" > $OUT/pytest_chunk.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_chunk.log; tail -15 $OUT/pytest_chunk.log
