#!/bin/bash
# Developer timing: where the move-sorted kernels start to pay (lanes x CTA size).
OUT=gpurun_out/r01n; mkdir -p $OUT
for L in 2048 4096 8192; do
  for B in 0 64 128 256; do
    echo "block $B lanes $L: $(KC_SMP_BLOCK=$B timeout 120 python tools/dev_sampler_prof.py $L 50 3 2>&1 | tail -1 | cut -c1-230)" >> $OUT/sorted_small.txt
  done
done
cat $OUT/sorted_small.txt
