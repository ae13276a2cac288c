#!/bin/bash
# Two-GPU session: the two-rank device-proposal tests, the multi-GPU tests, then bench.py on 2 ranks.
TAG=${1:-r01e}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_device_proposals.py tests/test_host_cpp.py tests/test_gpu_parity.py -m gpu -q -k "two_ranks or two_gpus or ranks or multi" > $OUT/pytest_2gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 \
   bench.py --gpus 2 --steps 10 --warmup 3 > $OUT/bench_2gpu.json 2> $OUT/bench_2gpu.err; echo "bench rc=$?" >> $OUT/bench_2gpu.err
tail -6 $OUT/pytest_2gpu.log; tail -c 1500 $OUT/bench_2gpu.json
