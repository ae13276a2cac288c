"""Developer timing of the end-to-end device-proposal runs with the host-side profile line."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sampler_bench
for name, lanes, saves in (("51Peg", 65536, 20), ("51Peg", 16384, 40), ("CoRoT7", 16384, 10)):
    r = sampler_bench.run("gpu", name, lanes, 100, saves, env={"KIMA_DEVICE_PROPOSALS": "1"}) or {}
    print(json.dumps({k: r.get(k) for k in ("example", "threads", "steps_per_s_run", "steps_per_s_loop", "loop_s", "one_off", "run_s", "wall_s", "mcmc_steps", "profile", "error")}), flush=True)
