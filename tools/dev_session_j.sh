#!/bin/bash
OUT=gpurun_out/r01j; mkdir -p $OUT
timeout 900 python -m pytest tests/test_sampler_gpu.py tests/test_device_proposals.py -m gpu -q -x -k "sorted or mcmc_on_device" > $OUT/pytest.log 2>&1; echo "rc=$?" >> $OUT/pytest.log
tail -3 $OUT/pytest.log
