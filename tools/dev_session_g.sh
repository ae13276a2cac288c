#!/bin/bash
# Developer session: the move-sorted proposal kernels (KC_SMP_BLOCK) — parity tests forced through them, then timings.
OUT=gpurun_out/r01g; mkdir -p $OUT
for B in 64 256; do
KC_SMP_BLOCK=$B timeout 900 python -m pytest tests/test_sampler_gpu.py -m gpu -q -x > $OUT/pytest_sorted_$B.log 2>&1; echo "rc=$?" >> $OUT/pytest_sorted_$B.log
tail -3 $OUT/pytest_sorted_$B.log
done
for B in 0 64 128 256; do
  for L in 16384 65536 262144; do
    echo "block $B lanes $L: $(KC_SMP_BLOCK=$B timeout 120 python tools/dev_sampler_prof.py $L 50 3 2>&1 | tail -1)" >> $OUT/sorted.txt
  done
done
echo "block 128 lanes 16384 GP N=177 npmax 5: $(KC_SMP_BLOCK=128 timeout 200 python tools/dev_sampler_prof.py 16384 10 2 177 5 gp 2>&1 | tail -1)" >> $OUT/sorted.txt
echo "block 0 lanes 16384 GP N=177 npmax 5: $(KC_SMP_BLOCK=0 timeout 200 python tools/dev_sampler_prof.py 16384 10 2 177 5 gp 2>&1 | tail -1)" >> $OUT/sorted.txt
cut -c1-400 $OUT/sorted.txt
