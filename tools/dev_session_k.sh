#!/bin/bash
# Developer check: the tabled kill_lagging_device against the previous build (host_old/), many replacements.
OUT=gpurun_out/r01k; mkdir -p $OUT
E2E=tests/golden/e2e
ls $E2E | head -3
for v in new old; do
  for ex in BL2009 51Peg; do
    d=/tmp/kl_${v}_$ex; rm -rf $d; mkdir -p $d
    cp $E2E/* $d/ 2>/dev/null
    printf "2\n50\n10\n5\n0\n1\n100\n600\n" > $d/OPTIONS
    exe=host/_build/kima_gpu_$ex; [ $v = old ] && exe=host_old/_build/kima_gpu_$ex
    (cd $d && KIMA_DEVICE_PROPOSALS=1 $GRAFT_REPO_ROOT/$exe -t 64 -s 43 > stdout.txt 2> stderr.txt; echo "$v $ex rc=$?")
  done
done
for ex in BL2009 51Peg; do
  for f in sample.txt sample_info.txt levels.txt sampler_state.txt stdout.txt; do
    cmp /tmp/kl_new_$ex/$f /tmp/kl_old_$ex/$f && echo "$ex $f identical"
  done
  grep -c "Replacing lagging" /tmp/kl_new_$ex/stdout.txt
done 2>&1 | tee $OUT/kill_compare.txt
KIMA_PROFILE=1 timeout 300 python tools/dev_devprop_bench2.py 2>&1 | head -1 | cut -c1-1500 | tee $OUT/devprop.txt
