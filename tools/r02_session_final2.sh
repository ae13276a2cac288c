#!/bin/bash
# r02 closing pass at HEAD: GPU suite, smoke, both bench arms (the ncu passes of r02_session_final.sh stay valid: same work per launch)
OUT=gpurun_out/r02final2; mkdir -p $OUT
timeout 1800 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log; tail -3 $OUT/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" >> $OUT/smoke.log; tail -2 $OUT/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" >> $OUT/bench.err; tail -2 $OUT/bench.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?" >> $OUT/bench_ref.err
timeout 600 python tools/dev_cache_sampler.py 2>&1 | tee $OUT/cache_sampler.jsonl | cut -c1-500
