"""Developer check on a GPU box: small-N GP kernel vs the blocked path, timings."""
import sys, json
sys.path.insert(0, ".")
import numpy as np
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params

def timing(N, B, npmax=5, reps=5):
    flags = ModelFlags(GP=True, npmax=npmax)
    d = make_data(N)
    P, nps = make_params(flags, d, B, "max")
    out = {}
    with LikelihoodEngine(d, flags, B) as eng:
        for small in (1, 0):
            eng.set_option("gp_small", small)
            res = eng.evaluate(P, nps)
            ms = []
            for _ in range(reps):
                eng.flush_l2(); eng.log_likelihood_device(B); eng.sync(); ms.append(eng.last_kernel_ms())
            out[small] = (min(ms), res)
    F = N ** 3 / 3 + 37.0 * N * N + N * (330.0 * npmax + 45)
    rel = np.max(np.abs(out[1][1] - out[0][1]) / np.abs(out[0][1]))
    print(json.dumps(dict(N=N, B=B, small_ms=out[1][0], blocked_ms=out[0][0], small_evals_s=B / out[1][0] * 1e3,
                          blocked_evals_s=B / out[0][0] * 1e3, small_tflops=B / out[1][0] * 1e3 * F * 1e-12,
                          rel_small_vs_blocked=float(rel))), flush=True)

if __name__ == "__main__":
    for N, B in ((177, 8192), (64, 8192), (100, 8192), (208, 4096), (32, 8192)):
        timing(N, B)
