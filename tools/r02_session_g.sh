#!/bin/bash
# r02 session g: celerite predictor (kc_predict with kernel=celerite) parity
OUT=gpurun_out/r02g; mkdir -p $OUT
timeout 600 python -m pytest tests/test_predictive.py -m gpu -x -q > $OUT/pytest_predictive.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_predictive.log
tail -25 $OUT/pytest_predictive.log
