#!/bin/bash
# developer probe: device proposals vs host proposals (reference chain) at the same lane count
set -e
R=$PWD
for mode in dev host; do
  d=/tmp/dp_$mode; rm -rf $d; mkdir -p $d; cp tests/golden/e2e/BL2009_dataset2.kms.rv $d/
  printf "2\n1000\n500\n100\n0\n10\n100\n${SAVES:-4000}\n" > $d/OPTIONS
  ( cd $d; if [ $mode = dev ]; then export KIMA_DEVICE_PROPOSALS=1; fi; KIMA_PROFILE=1 $R/host/_build/kima_gpu_BL2009 -t ${LANES:-256} -s 43 > out.log 2> err.log; tail -3 err.log )
  mkdir -p gpurun_out/dp_$mode; cp $d/levels.txt $d/sample_info.txt $d/err.log gpurun_out/dp_$mode/; grep -n 'Creating level\|Done creating' $d/out.log | awk -F: '{print $1, $2}' > gpurun_out/dp_$mode/levels_created.txt; tail -5 $d/out.log
done
python - <<'PY'
import sys; sys.path.insert(0, 'tests')
from dnest_post import log_evidence
for m in ("dev", "host"):
    z, w, H = log_evidence("/tmp/dp_" + m)
    print(m, "logZ", z, "H", H)
PY
