#!/bin/bash
OUT=gpurun_out/r02c13; mkdir -p $OUT
timeout 300 python tools/dev_small_cfg.py 2>&1 | tee $OUT/cfg.jsonl
