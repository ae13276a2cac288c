"""Developer check: is the host-proposal sampler reproducible run to run when the GPU is fast (51Peg, 4096 lanes)?"""
import filecmp, os, shutil, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from dev_cache_sampler import one, FILES
name, threads, ts, saves, reps = sys.argv[1], int(sys.argv[2]), 100, int(sys.argv[3]), int(sys.argv[4])
env = {"KIMA_CUDA_GP_CACHE": sys.argv[5]} if len(sys.argv) > 5 else {}
base = None
for i in range(reps):
    d, res = one(name, threads, ts, saves, env)
    if base is None:
        base = d; print("run 0", res); continue
    diff = [f for f in FILES if not filecmp.cmp(os.path.join(base, f), os.path.join(d, f), shallow=False)]
    print("run", i, "differs in", diff, res.get("steps_per_s_run"))
    shutil.rmtree(d, ignore_errors=True)
