"""Developer timing: MCMC steps/s of the sampling loop, host proposals vs device proposals
(KIMA_DEVICE_PROPOSALS=1), over lane counts."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sampler_bench

names = sys.argv[1].split(",") if len(sys.argv) > 1 else ["51Peg", "CoRoT7"]
lanes_list = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [4096, 16384, 65536]
for n in names:
    for lanes in lanes_list:
        for mode in ("host", "device"):
            if mode == "host" and lanes > 16384:
                continue
            env = {"KIMA_DEVICE_PROPOSALS": "1"} if mode == "device" else {}
            saves = 40 if mode == "device" else 20
            r = sampler_bench.run("gpu", n, lanes, 100, saves, env=env)
            r = r or {}
            print(json.dumps({"example": n, "lanes": lanes, "mode": mode, "steps_per_s_run": r.get("steps_per_s_run"),
                              "run_s": r.get("run_s"), "wall_s": r.get("wall_s"), "mcmc_steps": r.get("mcmc_steps"),
                              "error": r.get("error")}), flush=True)
