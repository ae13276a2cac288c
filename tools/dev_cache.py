"""Developer check on a GPU box: factor-cache miss / hit / uncached times of the small GP path."""
import json, sys
sys.path.insert(0, ".")
import numpy as np
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params

for N, B in ((177, 8192), (100, 8192), (64, 8192), (208, 4096)):
    flags = ModelFlags(GP=True, npmax=5, hyperpriors=(N == 177))
    d = make_data(N)
    P, nps = make_params(flags, d, B, "mix")
    Q = P.copy(); Q[0] *= 1.01
    with LikelihoodEngine(d, flags, B) as eng:
        eng.evaluate(P, nps)
        plain = []
        for _ in range(3):
            eng.flush_l2(); eng.evaluate(P, nps); plain.append(eng.last_kernel_ms())
        n = eng.enable_gp_cache(B)
        slots = np.arange(B, dtype=np.int32)
        R = P.copy()
        miss, hit = [], []
        for it in range(3):
            R[eng.layout["row_eta"] + 1] *= 1.001   # new C in every lane
            eng.set_slots(slots); eng.flush_l2(); eng.evaluate(R, nps); miss.append(eng.last_kernel_ms())
            R[eng.layout["row_vsys"]] += 0.1         # same C
            eng.set_slots(slots); eng.flush_l2(); eng.evaluate(R, nps); hit.append(eng.last_kernel_ms())
    print(json.dumps(dict(N=N, B=B, slots=n, plain_ms=min(plain), miss_ms=min(miss), hit_ms=min(hit),
                          hit_evals_s=B / min(hit) * 1e3, mix_625_ms=0.375 * min(miss) + 0.625 * min(hit))), flush=True)
