#!/bin/bash
OUT=gpurun_out/r01m; mkdir -p $OUT
timeout 600 python -m pytest tests/test_device_proposals.py -m gpu -q -x -k "many_lanes_sorted or reproducible" > $OUT/pytest.log 2>&1; echo "rc=$?" >> $OUT/pytest.log
tail -3 $OUT/pytest.log
KIMA_PROFILE=1 timeout 300 python tools/dev_devprop_bench2.py 2>&1 | head -1 | cut -c1-1500 | tee $OUT/devprop.txt
