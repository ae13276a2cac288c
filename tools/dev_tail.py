"""Tail of the logL parity distribution on large batches, both Kepler modes."""
import sys, json
sys.path.insert(0, ".")
import numpy as np
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params
from oracle.oracle import Oracle
port = Oracle("port")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
for name, flags, N in [("N1000_np10", ModelFlags(npmax=10), 1000), ("N200_np5_hyper", ModelFlags(npmax=5, hyperpriors=True), 200)]:
    d = make_data(N)
    P, nps = make_params(flags, d, B, "max", seed=7)
    port.set_rvdata(d); port.set_flags(flags)
    ref, secs = port.eval_batch(P, nps, 16)
    for mode in (0, 1):
        eng = LikelihoodEngine(d, flags, B, kepler_mode=mode)
        got = eng.evaluate(P, nps)
        rel = np.abs(got - ref) / np.abs(ref)
        q = {str(k): float(np.quantile(rel, k)) for k in (0.5, 0.99, 0.999, 0.9999)}
        print(json.dumps(dict(case=name, mode=mode, B=B, max=float(rel.max()), n_gt_1e10=int((rel > 1e-10).sum()),
                              n_gt_1e11=int((rel > 1e-11).sum()), n_gt_1e12=int((rel > 1e-12).sum()), q=q,
                              absmax=float(np.abs(got - ref).max()), medabs_logL=float(np.median(np.abs(ref))))), flush=True)
        eng.close()
