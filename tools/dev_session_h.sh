#!/bin/bash
# Developer session: where the one-off time before the sampling loop goes (three identical runs).
OUT=gpurun_out/r01h; mkdir -p $OUT
for i in 1 2 3; do
python - <<'PY' 2>&1 | tee -a $OUT/create.txt
import sys, json
sys.path.insert(0, "tools")
import sampler_bench
r = sampler_bench.run("gpu", "51Peg", 65536, 100, 20, env={"KIMA_DEVICE_PROPOSALS": "1"})
print(json.dumps({k: r.get(k) for k in ("steps_per_s_run", "steps_per_s_loop", "one_off", "run_s")}))
PY
done
