#!/bin/bash
# developer probe: scatter of log Z over seeds, device proposals vs host proposals, same lanes
R=$PWD
for seed in ${SEEDS:-7 8 9}; do for mode in dev host; do
  d=/tmp/dq_${mode}_$seed; rm -rf $d; mkdir -p $d; cp tests/golden/e2e/BL2009_dataset2.kms.rv $d/
  printf "2\n1000\n500\n100\n0\n10\n100\n${SAVES:-2000}\n" > $d/OPTIONS
  ( cd $d; if [ $mode = dev ]; then export KIMA_DEVICE_PROPOSALS=1; fi; $R/host/_build/kima_gpu_BL2009 -t ${LANES:-256} -s $seed > out.log 2> err.log )
done; done
python - <<'PY'
import sys, glob, numpy as np; sys.path.insert(0, 'tests')
from dnest_post import log_evidence
for d in sorted(glob.glob("/tmp/dq_*")):
    z, w, H = log_evidence(d)
    lv = np.loadtxt(d + "/levels.txt", ndmin=2)
    tr = lv[:, 4]
    print(d, "logZ %.2f H %.1f levels %d tries min/max %.2e %.2e" % (z, H, len(lv), tr.min(), tr.max()))
PY
