#!/bin/bash
# r02 session e: look-ahead partial of the next diagonal block inside the update launch (blocked GP path),
# U look-ahead in gps_kernel: parity gate, timings, GPU suite
OUT=gpurun_out/r02e; mkdir -p $OUT
timeout 300 python tools/dev_gp.py parity > $OUT/dev_gp_parity.jsonl 2>&1; rc=$?; echo "parity rc=$rc" >> $OUT/dev_gp_parity.jsonl
python - <<'PY'
import json
mx=0
for l in open("gpurun_out/r02e/dev_gp_parity.jsonl"):
    if l.startswith("{"):
        d=json.loads(l); mx=max(mx,d["max"]); ok=d["inf_match"]
        if d["max"]>1e-10 or not ok: print("BAD", d)
print("parity worst", mx)
PY
tail -2 $OUT/dev_gp_parity.jsonl
if [ $rc -ne 0 ]; then echo "GATE FAILED"; exit 1; fi
timeout 300 python tools/dev_gp.py timing > $OUT/dev_gp_timing.jsonl 2>&1; cat $OUT/dev_gp_timing.jsonl
timeout 300 python tools/dev_small.py > $OUT/dev_small.jsonl 2>&1; cat $OUT/dev_small.jsonl
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
tail -5 $OUT/pytest_gpu.log
