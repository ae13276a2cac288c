# developer session for the small-N GP kernel: parity, timings, in-kernel phase profile
OUT=gpurun_out/${1:-small}
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "small or golden or gp" > $OUT/pytest.log 2>&1; tail -2 $OUT/pytest.log
timeout 200 python tools/dev_small.py 2>&1 | tee $OUT/small.log
if [ -f variants/prof.so ]; then
  cp kima_b200/libkima_cuda.so /tmp/o.so; cp variants/prof.so kima_b200/libkima_cuda.so
  timeout 100 python tools/dev_small_prof.py 177 1184 2>&1 | tee $OUT/p177.log
  cp /tmp/o.so kima_b200/libkima_cuda.so
fi
