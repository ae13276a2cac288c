#!/bin/bash
# the GPU suite alone (the rest of the closing pass is tools/r02_session_final2.sh)
OUT=gpurun_out/r02final3; mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log; tail -4 $OUT/pytest_gpu.log
