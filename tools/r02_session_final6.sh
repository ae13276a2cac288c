#!/bin/bash
# two ranks, device proposals: kill_lagging's first loop and the record-holder search sharded over the ranks
OUT=gpurun_out/r02final6; mkdir -p $OUT
KIMA_CHECK_SHARDED_KILL=1 timeout 200 python -m pytest tests/test_device_proposals.py -m gpu -q -k "two_ranks" > $OUT/pytest_2rank_dev.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_2rank_dev.log; tail -12 $OUT/pytest_2rank_dev.log
