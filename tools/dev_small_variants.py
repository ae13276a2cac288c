"""Developer check on a GPU box: time gps_kernel builds given as KIMA_CUDA_LIB (one process per library)."""
import json, os, sys
sys.path.insert(0, ".")
import numpy as np
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params

tag = os.path.basename(os.environ.get("KIMA_CUDA_LIB", "product"))
for N, B in ((177, 8192), (100, 8192), (208, 4096), (64, 8192), (96, 8192), (128, 8192), (150, 8192)):
    flags = ModelFlags(GP=True, npmax=5)
    d = make_data(N)
    P, nps = make_params(flags, d, B, "max")
    with LikelihoodEngine(d, flags, B) as eng:
        res = eng.evaluate(P, nps)
        ms = []
        for _ in range(5):
            eng.flush_l2(); eng.log_likelihood_device(B); eng.sync(); ms.append(eng.last_kernel_ms())
        eng.set_option("gp_small", 0)
        ref = eng.evaluate(P, nps)
    rel = float(np.max(np.abs(res - ref) / np.abs(ref)))
    print(json.dumps(dict(lib=tag, N=N, B=B, ms=min(ms), evals_s=B / min(ms) * 1e3, rel_vs_blocked=rel)), flush=True)
