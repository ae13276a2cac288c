#!/bin/bash
# One GPU-box session: parity tests, bench (both arms), ncu launch list + full captures.
# Usage: tools/gpu_session.sh <tag>
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
lscpu | head -20 > $OUT/lscpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" >> $OUT/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" >> $OUT/bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err
timeout 300 python tools/dev_small.py > $OUT/dev_small.jsonl 2>&1
timeout 300 python tools/dev_gp.py timing > $OUT/dev_gp.jsonl 2>&1
timeout 300 python tools/dev_predict.py > $OUT/dev_predict.jsonl 2>&1
timeout 60 tools/_build/fp64_latency > $OUT/fp64_latency.txt 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
   --log-file $OUT/launches.csv python bench.py --steps 2 --warmup 3 --skip-cpu-baseline --skip-sampler > $OUT/ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:keplerian -c 1 -s 3 \
   -o $OUT/kep_full -f python bench.py --steps 2 --warmup 3 --skip-cpu-baseline --skip-sampler --workloads kep > $OUT/ncu_kep.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:gpb_update -c 1 -s 40 \
   -o $OUT/gp_full -f python bench.py --steps 2 --warmup 3 --skip-cpu-baseline --skip-sampler --workloads gp > $OUT/ncu_gp.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gps_kernel -c 1 \
   -o $OUT/gps_full -f python tools/dev_small_prof.py 177 1184 > $OUT/ncu_gps.log 2>&1
ls -la $OUT
tail -3 $OUT/pytest_gpu.log; tail -8 $OUT/smoke.log; cat $OUT/bench.json
