#!/bin/bash
# One GPU-box session (round r01e): all GPU tests, smoke, bench (both arms), ncu launch list.
TAG=${1:-r01e}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" >> $OUT/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" >> $OUT/bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
   --log-file $OUT/launches_sampler.csv python tools/dev_sampler_prof.py 65536 20 1 > $OUT/ncu_list_sampler.log 2>&1
tail -5 $OUT/pytest_gpu.log; tail -8 $OUT/smoke.log; cat $OUT/bench.json | head -c 3000
timeout 300 python tools/dev_devprop_bench2.py > $OUT/devprop_e2e.jsonl 2> $OUT/devprop_e2e.err
cat $OUT/devprop_e2e.jsonl
for lanes in 4096 16384 65536 262144; do timeout 300 python tools/dev_sampler_prof.py $lanes 50 3; done > $OUT/dev_sampler_timings.jsonl 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:smp_sorted_propose -c 1 -s 20 \
   -o $OUT/smp_full -f python tools/dev_sampler_prof.py 65536 30 1 > $OUT/ncu_smp.log 2>&1
tail -2 $OUT/ncu_smp.log
