#!/bin/bash
OUT=gpurun_out/r02c2; mkdir -p $OUT
timeout 300 python tools/dev_cache.py 2>&1 | tee $OUT/dev_cache.jsonl
