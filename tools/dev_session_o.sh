#!/bin/bash
OUT=gpurun_out/r01o; mkdir -p $OUT
timeout 900 python -m pytest tests/test_sampler_gpu.py tests/test_device_proposals.py -m gpu -q -x > $OUT/pytest.log 2>&1; echo "rc=$?" >> $OUT/pytest.log
tail -3 $OUT/pytest.log
