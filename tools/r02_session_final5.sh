#!/bin/bash
# two ranks, host proposals, CoRoT7 (factor cache on by default on every rank): the single-process chain
OUT=gpurun_out/r02final5; mkdir -p $OUT
timeout 200 python -m pytest tests/test_host_cpp.py -m gpu -q -k "two_ranks and CoRoT7" > $OUT/pytest_2rank.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_2rank.log; tail -8 $OUT/pytest_2rank.log
