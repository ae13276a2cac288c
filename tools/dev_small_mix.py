"""Developer check: gps_kernel builds given as KIMA_CUDA_LIB, timed with Np ~ U{0..5} (SMs out of step) and with Np = 5 (in step)."""
import json, os, sys
sys.path.insert(0, ".")
import numpy as np
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params
tag = os.path.basename(os.environ.get("KIMA_CUDA_LIB", "product"))
for N, B in ((177, 16384), (100, 16384), (64, 16384), (208, 8192)):
    d = make_data(N)
    flags = ModelFlags(GP=True, npmax=5, hyperpriors=True)
    out = dict(lib=tag, N=N, B=B)
    for mode in ("mix", "max"):
        P, nps = make_params(flags, d, B, mode)
        with LikelihoodEngine(d, flags, B) as eng:
            eng.evaluate(P, nps)
            ms = []
            for _ in range(4):
                eng.flush_l2(); eng.log_likelihood_device(B); eng.sync(); ms.append(eng.last_kernel_ms())
        out[mode + "_ms"] = min(ms)
    print(json.dumps(out), flush=True)
