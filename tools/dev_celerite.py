"""Developer check on a GPU box: celerite path timing (GPU vs the reference's CPU path)."""
import sys, json, os, time
sys.path.insert(0, ".")
import numpy as np
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params
from oracle.oracle import Oracle

for N, B in ((1000, 8192), (177, 8192), (1000, 65536)):
    flags = ModelFlags(GP=True, kernel=1, npmax=10 if N == 1000 else 5)
    d = make_data(N)
    P, nps = make_params(flags, d, B, "max")
    with LikelihoodEngine(d, flags, B) as eng:
        got = eng.evaluate(P, nps)
        ms = []
        for _ in range(5):
            eng.flush_l2(); eng.log_likelihood_device(B); eng.sync(); ms.append(eng.last_kernel_ms())
    ref = Oracle("reference")
    ref.set_rvdata(d); ref.set_flags(flags)
    nref = 2048
    t0 = time.perf_counter()
    r, _ = ref.eval_batch(P[:, :nref].copy(), nps[:nref], os.cpu_count())
    dt = time.perf_counter() - t0
    rel = np.max(np.abs(got[:nref] - r) / np.abs(r))
    print(json.dumps(dict(N=N, B=B, ms=min(ms), evals_s=B / min(ms) * 1e3, cpu_evals_s=nref / dt, cores=os.cpu_count(),
                          rel=float(rel))), flush=True)
