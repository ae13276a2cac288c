#!/bin/bash
OUT=gpurun_out/r02c24; mkdir -p $OUT
echo "== device loop"; timeout 400 python tools/dev_determinism.py 177 2048 240 2>&1 | tail -20 | tee $OUT/dev.txt
echo "== host loop"; timeout 200 python tools/dev_determinism.py 177 2048 90 host 2>&1 | tail -20 | tee $OUT/host.txt
