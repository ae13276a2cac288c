#!/bin/bash
OUT=gpurun_out/r02c18; mkdir -p $OUT
for k in 51Peg BL2009; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:keplerian_warp -s 1 -c 1 -o $OUT/kep_$k -f python bench.py --profile-one $k > $OUT/ncu_$k.log 2>&1; tail -1 $OUT/ncu_$k.log
done
