#!/bin/bash
# r02 final session: the round's measurement pass at HEAD: GPU suite, smoke, both bench arms, ncu launch list,
# executed-FLOP pass and full captures of the three GP kernels
OUT=gpurun_out/r02final; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
lscpu | grep -i "model name\|^CPU(s)" > $OUT/lscpu.txt
timeout 1800 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log; tail -3 $OUT/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" >> $OUT/smoke.log; tail -3 $OUT/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" >> $OUT/bench.err; tail -2 $OUT/bench.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?" >> $OUT/bench_ref.err
M=smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__ops_path_tensor_src_fp64.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
for k in kep gp gp4096 CoRoT7 51Peg many_planets; do
  timeout 900 ncu --metrics $M --clock-control none --csv --log-file $OUT/exec_$k.csv python bench.py --profile-one $k > $OUT/ncu_exec_$k.log 2>&1
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gpb_update -s 40 -c 1 -o $OUT/gp_update_full -f python bench.py --profile-one gp > $OUT/ncu_gp.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gpb_diag -s 40 -c 1 -o $OUT/gp_diag_full -f python bench.py --profile-one gp > $OUT/ncu_gpd.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gps_kernel -s 1 -c 1 -o $OUT/gps_full -f python bench.py --profile-one CoRoT7 > $OUT/ncu_gps.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:keplerian_warp -s 1 -c 1 -o $OUT/kep_full -f python bench.py --profile-one kep > $OUT/ncu_kep.log 2>&1
ls -la $OUT | head -40
