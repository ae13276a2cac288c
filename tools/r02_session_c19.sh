#!/bin/bash
OUT=gpurun_out/r02c19; mkdir -p $OUT
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_host_cpp.py -m gpu -x -q > $OUT/pytest.txt 2>&1; echo "rc=$?" >> $OUT/pytest.txt; tail -4 $OUT/pytest.txt
timeout 900 python bench.py --steps 10 --warmup 3 --skip-sampler --skip-predictive --skip-cpu-baseline > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; tail -2 $OUT/bench.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02c19/bench.json"))
print("kep", d["value"], d["roofline"]["workloads"]["kep"]["frac"])
for s in d["shapes"]: print(s["name"], "%.4g"%s["value"], "%.3f"%s["roofline"]["frac"])
for k,v in d.get("np_mix",{}).items(): print("np_mix",k,"%.4g"%v["value"], "%.3f"%v["roofline"]["frac"])
PY
