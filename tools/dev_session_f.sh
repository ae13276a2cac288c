#!/bin/bash
OUT=gpurun_out/r01f; mkdir -p $OUT
timeout 600 python -m pytest tests/test_device_proposals.py -m gpu -q -k "state_file or reproducible" > $OUT/pytest_state.log 2>&1; echo "rc=$?" >> $OUT/pytest_state.log
tail -4 $OUT/pytest_state.log
KIMA_PROFILE=1 timeout 300 python tools/dev_devprop_bench2.py > $OUT/devprop_e2e.jsonl 2> $OUT/devprop_e2e.err
cat $OUT/devprop_e2e.jsonl | cut -c1-1200
bash tools/dev_regcap.sh
