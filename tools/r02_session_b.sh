#!/bin/bash
# r02 session b: the new GP kernels (8-column panel Cholesky, TMA/mbarrier pipeline) — quick parity
# gate first (a hang must not eat the box), then timings, the GPU suite, the new bench line and ncu.
OUT=gpurun_out/r02b; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 300 python tools/dev_gp.py parity > $OUT/dev_gp_parity.jsonl 2>&1; rc=$?; echo "parity rc=$rc" >> $OUT/dev_gp_parity.jsonl
tail -20 $OUT/dev_gp_parity.jsonl
if [ $rc -ne 0 ]; then echo "GATE FAILED"; exit 1; fi
timeout 300 python tools/dev_small.py > $OUT/dev_small.jsonl 2>&1; echo "small rc=$?" >> $OUT/dev_small.jsonl
cat $OUT/dev_small.jsonl
timeout 300 python tools/dev_gp.py timing > $OUT/dev_gp_timing.jsonl 2>&1; echo "timing rc=$?" >> $OUT/dev_gp_timing.jsonl
cat $OUT/dev_gp_timing.jsonl
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
tail -5 $OUT/pytest_gpu.log
timeout 900 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" >> $OUT/bench.err
tail -3 $OUT/bench.err; head -c 3000 $OUT/bench.json
M=smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__ops_path_tensor_src_fp64.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
for k in kep gp CoRoT7; do
  timeout 600 ncu --metrics $M --clock-control none --csv --log-file $OUT/exec_$k.csv python bench.py --profile-one $k > $OUT/ncu_exec_$k.log 2>&1
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gpb_update -s 40 -c 1 -o $OUT/gp_update_full -f python bench.py --profile-one gp > $OUT/ncu_gp.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gpb_diag -s 40 -c 1 -o $OUT/gp_diag_full -f python bench.py --profile-one gp > $OUT/ncu_gpd.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gps_kernel -s 1 -c 1 -o $OUT/gps_full -f python bench.py --profile-one CoRoT7 > $OUT/ncu_gps.log 2>&1
ls -la $OUT
