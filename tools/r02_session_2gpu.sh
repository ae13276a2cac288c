#!/bin/bash
# r02: two B200s of one box: the bench's multi-GPU path (torchrun, one rank per GPU) and the 2-GPU byte-parity tests
OUT=gpurun_out/r02_2gpu; mkdir -p $OUT
nvidia-smi --query-gpu=index,name --format=csv > $OUT/smi.txt 2>&1
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 > $OUT/bench_2gpu.json 2> $OUT/bench_2gpu.err; echo "bench rc=$?" >> $OUT/bench_2gpu.err
tail -5 $OUT/bench_2gpu.err; tail -c 2500 $OUT/bench_2gpu.json
timeout 900 python -m pytest tests/test_host_cpp.py tests/test_device_proposals.py -m gpu -q -k "two_ranks" > $OUT/pytest_2gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_2gpu.log; tail -4 $OUT/pytest_2gpu.log
