"""Developer check on a GPU box: the small GP path at N=177, hyperpriors on, as a function of the planet count."""
import json, sys
sys.path.insert(0, ".")
import numpy as np
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params
N, B = 177, 16384
d = make_data(N)
for hyper in (True, False):
    flags = ModelFlags(GP=True, npmax=5, hyperpriors=hyper)
    P, nps_mix = make_params(flags, d, B, "mix")
    with LikelihoodEngine(d, flags, B) as eng:
        for tag, nps in [("mix", nps_mix)] + [("all %d" % k, np.full(B, k, np.int32)) for k in (0, 1, 3, 5)] + [("sorted mix", np.sort(nps_mix))]:
            eng.evaluate(P, nps)
            ms = []
            for _ in range(4):
                eng.flush_l2(); eng.log_likelihood_device(B); eng.sync(); ms.append(eng.last_kernel_ms())
            print(json.dumps(dict(N=N, B=B, hyperpriors=hyper, np=tag, ms=min(ms), evals_s=B / min(ms) * 1e3)), flush=True)
