"""Developer check on a GPU box: kc_predict timings (posterior-predictive curves)."""
import sys, json, time
sys.path.insert(0, ".")
import numpy as np
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params

def timing(N, B, nt, npmax=5, reps=3):
    flags = ModelFlags(GP=True, npmax=npmax)
    d = make_data(N)
    P, nps = make_params(flags, d, B, "max")
    tg = np.linspace(d.t.min(), d.t.max(), nt)
    with LikelihoodEngine(d, flags, B) as eng:
        eng.predict(P, nps, tg)
        best = 1e9; kms = 1e9
        for _ in range(reps):
            t0 = time.perf_counter(); eng.predict(P, nps, tg); best = min(best, time.perf_counter() - t0)
            kms = min(kms, eng.last_kernel_ms())
        eng.set_option("gp_small", 0)
        eng.evaluate(P, nps)
        fms = 1e9
        for _ in range(reps):
            eng.log_likelihood_device(B); eng.sync(); fms = min(fms, eng.last_kernel_ms())
    print(json.dumps(dict(N=N, B=B, nt=nt, predict_wall_ms=best * 1e3, predict_kernels_ms=kms, factor_only_ms=fms,
                          curves_per_s=B / best)), flush=True)

if __name__ == "__main__":
    for N, B, nt in ((177, 4096, 1000), (1000, 1184, 1000), (1000, 1184, 5000)):
        timing(N, B, nt)
