#!/bin/bash
OUT=gpurun_out/r02x; mkdir -p $OUT
for v in old new; do echo "== $v"; timeout 60 tools/_build/chol_bench_$v; done 2>&1 | tee $OUT/chol_bench.txt
