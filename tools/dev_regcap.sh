#!/bin/bash
# Developer experiment: the step kernels compiled with other register budgets
# (variants/ built here with -maxrregcount R -DKC_SMP_MIN_CTAS=n; r156 = __launch_bounds__(32, 1)).
OUT=gpurun_out/regcap; mkdir -p $OUT
for R in ${REGS:-128 156}; do
  cp variants/libkima_cuda_r$R.so kima_b200/libkima_cuda.so
  for L in 65536 262144; do
    echo "regs $R lanes $L: $(timeout 120 python tools/dev_sampler_prof.py $L 50 3 2>&1 | tail -1)" >> $OUT/regcap.txt
  done
done
cp variants/libkima_cuda_r128.so kima_b200/libkima_cuda.so
cat $OUT/regcap.txt
