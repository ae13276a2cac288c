#!/bin/bash
# r02 session f: GPU suite with the new parity cases (tail lanes, five more chain-equality set-ups, CoRoT7 evidence)
OUT=gpurun_out/r02f; mkdir -p $OUT
timeout 1800 python -m pytest tests -m gpu -x -q -s > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
grep -E "CoRoT7 log Z|passed|failed|rc=" $OUT/pytest_gpu.log | tail -8
