#!/bin/bash
# developer session for the device-side sampler: timings over lane counts + one ncu capture
OUT=gpurun_out/${1:-smp}
mkdir -p $OUT
for lanes in 4096 16384 65536 262144; do timeout 300 python tools/dev_sampler_prof.py $lanes 50 3; done 2>&1 | tee $OUT/timings.jsonl
timeout 600 ncu --set full --clock-control none --import-source on -k regex:smp_accept_propose -c 1 -s 20 \
   -o $OUT/smp_full -f python tools/dev_sampler_prof.py 65536 30 1 > $OUT/ncu_smp.log 2>&1
tail -3 $OUT/ncu_smp.log
