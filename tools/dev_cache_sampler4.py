"""Developer check: CoRoT7 device-proposal runs, cache off once and on N times: files byte for byte."""
import filecmp, os, shutil, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from dev_cache_sampler import one, FILES
saves, reps = int(sys.argv[1]), int(sys.argv[2])
env = {"KIMA_DEVICE_PROPOSALS": "1"}
base, res0 = one("CoRoT7", 16384, 100, saves, dict(env, KIMA_CUDA_GP_CACHE="0"))
print("off", res0)
for i in range(reps):
    d, res = one("CoRoT7", 16384, 100, saves, dict(env, KIMA_CUDA_GP_CACHE="1"))
    diff = [f for f in FILES if not filecmp.cmp(os.path.join(base, f), os.path.join(d, f), shallow=False)]
    print("on", i, "differs in", diff, res.get("steps_per_s_run"))
    shutil.rmtree(d, ignore_errors=True)
