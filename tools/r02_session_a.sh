#!/bin/bash
# r02 session a: the whole GPU suite at HEAD (with the no-alloc test last), the no-alloc test alone
# in a fresh process, and the default bench line as the round's starting point.
OUT=gpurun_out/r02a; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
lscpu | grep -i "model name\|^CPU(s)" > $OUT/lscpu.txt
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $OUT/pytest_noalloc_alone.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_noalloc_alone.log
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" >> $OUT/bench.err
tail -5 $OUT/pytest_gpu.log; tail -5 $OUT/pytest_noalloc_alone.log; tail -c 1500 $OUT/bench.json
