#!/bin/bash
# bench.py on N ranks of one box (usage: gpu_session_ngpu.sh <tag> <N>).
TAG=${1:-r01e}; N=${2:-4}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
   bench.py --gpus $N --steps 10 --warmup 3 > $OUT/bench_${N}gpu.json 2> $OUT/bench_${N}gpu.err; echo "bench rc=$?" >> $OUT/bench_${N}gpu.err
tail -c 900 $OUT/bench_${N}gpu.json
