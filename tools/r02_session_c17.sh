#!/bin/bash
OUT=gpurun_out/r02c17; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "lane_order or small_path" > $OUT/pytest.txt 2>&1; echo "rc=$?" >> $OUT/pytest.txt; tail -4 $OUT/pytest.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gps_kernel -s 1 -c 1 -o $OUT/gps_full -f python bench.py --profile-one CoRoT7 > $OUT/ncu_gps.log 2>&1; tail -2 $OUT/ncu_gps.log
