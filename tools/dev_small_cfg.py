"""Developer check on a GPU box: the small GP path at N=177 with / without hyperpriors, Np fixed / mixed."""
import json, sys
sys.path.insert(0, ".")
import numpy as np
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params
N, B = 177, 16384
d = make_data(N)
for hyper in (False, True):
    for mode in ("max", "mix"):
        flags = ModelFlags(GP=True, npmax=5, hyperpriors=hyper)
        P, nps = make_params(flags, d, B, mode)
        with LikelihoodEngine(d, flags, B) as eng:
            eng.evaluate(P, nps)
            ms = []
            for _ in range(4):
                eng.flush_l2(); eng.log_likelihood_device(B); eng.sync(); ms.append(eng.last_kernel_ms())
        print(json.dumps(dict(N=N, B=B, hyperpriors=hyper, np=mode, ms=min(ms), evals_s=B / min(ms) * 1e3)), flush=True)
