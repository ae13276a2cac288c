#!/bin/bash
# Last check of a round: smoke and the default bench line at HEAD.
OUT=gpurun_out/r01p; mkdir -p $OUT
timeout 100 python __graft_entry__.py --smoke 2>&1 | tail -2
timeout 170 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r01p/bench.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], d["roofline"]["frac"], d["gp"]["value"], d["gpu_launches"], d["clocks"])
for r in d["sampler"]["device_proposals"]:
    print(r["example"], r["steps_per_s_run"], r.get("steps_per_s_loop"))
PY
