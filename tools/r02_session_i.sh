#!/bin/bash
# r02 session i/l: parity gate (blocked path vs the C port), gps and blocked timings, GPU parity tests
OUT=gpurun_out/r02y; mkdir -p $OUT

timeout 300 python tools/dev_gp.py parity > $OUT/dev_gp_parity.jsonl 2>&1; rc=$?; echo "parity rc=$rc" >> $OUT/dev_gp_parity.jsonl
python - <<'PY'
import json
mx=0
for l in open("gpurun_out/r02y/dev_gp_parity.jsonl"):
    if l.startswith("{"):
        d=json.loads(l); mx=max(mx,d["max"])
        if d["max"]>1e-10 or not d["inf_match"]: print("BAD", d)
print("parity worst", mx)
PY
tail -1 $OUT/dev_gp_parity.jsonl
if [ $rc -ne 0 ]; then echo "GATE FAILED"; tail -20 $OUT/dev_gp_parity.jsonl; exit 1; fi
timeout 300 python tools/dev_small.py > $OUT/dev_small.jsonl 2>&1; cat $OUT/dev_small.jsonl
timeout 300 python tools/dev_gp.py timing > $OUT/dev_gp_timing.jsonl 2>&1; cat $OUT/dev_gp_timing.jsonl
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest.log; tail -3 $OUT/pytest.log
for N in 177; do
  KIMA_CUDA_LIB=variants/prof.so timeout 120 python tools/dev_small_prof.py $N 8192 > $OUT/prof_$N.txt 2>&1
  cat $OUT/prof_$N.txt
done
