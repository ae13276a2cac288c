#!/bin/bash
# r02 session o: in-kernel phase clock of the small-N kernel under full load (KC_GPS_PROFILE build)
OUT=gpurun_out/r02o; mkdir -p $OUT
for N in 177; do
  KIMA_CUDA_LIB=variants/prof.so timeout 120 python tools/dev_small_prof.py $N 8192 > $OUT/prof_$N.txt 2>&1
  cat $OUT/prof_$N.txt
done
