#!/bin/bash
# r02: four B200s of one box: the bench's multi-GPU path (torchrun, one rank per GPU)
OUT=gpurun_out/r02_4gpu; mkdir -p $OUT
nvidia-smi --query-gpu=index,name --format=csv > $OUT/smi.txt 2>&1
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 4 --steps 5 --warmup 3 > $OUT/bench_4gpu.json 2> $OUT/bench_4gpu.err; echo "bench rc=$?" >> $OUT/bench_4gpu.err
tail -3 $OUT/bench_4gpu.err; tail -c 1500 $OUT/bench_4gpu.json
