"""Developer check: two cached contexts driven back to back from two host threads (their launch sequences overlap on
the GPU), all results kept and compared with an uncached context afterwards."""
import json, sys, threading
sys.path.insert(0, ".")
import numpy as np
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params

N, P_PART, ITERS = 177, int(sys.argv[1]) if len(sys.argv) > 1 else 4096, int(sys.argv[2]) if len(sys.argv) > 2 else 300
flags = ModelFlags(GP=True, npmax=5, hyperpriors=True)
data = make_data(N)
log = {11: [], 12: []}


def drive(seed):
    rng = np.random.default_rng(seed)
    state, nps_all = make_params(flags, data, P_PART, "mix", seed=seed)
    with LikelihoodEngine(data, flags, P_PART) as eng:
        L = eng.layout
        assert eng.enable_gp_cache(P_PART) == P_PART
        for it in range(ITERS):
            sel = rng.permutation(P_PART)[: rng.integers(P_PART // 4, P_PART // 2)]
            prop = state[:, sel].copy()
            n = max(L["npmax"], 1)
            prop[L["row_K"]:L["row_K"] + n] *= rng.uniform(0.9, 1.1, (n, sel.size))
            ch = rng.random(sel.size) < 0.375
            prop[L["row_eta"] + 1, ch] *= rng.uniform(0.95, 1.05, ch.sum())
            eng.set_slots(sel.astype(np.int32))
            got = eng.evaluate(prop, nps_all[sel])
            log[seed].append((prop, nps_all[sel].copy(), got))
            acc = rng.random(sel.size) < 0.35
            state[:, sel[acc]] = prop[:, acc]


ts = [threading.Thread(target=drive, args=(s,)) for s in (11, 12)]
for t in ts: t.start()
for t in ts: t.join()
bad = []
with LikelihoodEngine(data, flags, P_PART) as plain:
    for seed, items in log.items():
        for it, (prop, nps, got) in enumerate(items):
            ref = plain.evaluate(prop, nps)
            if not np.array_equal(got, ref):
                w = np.nonzero(got != ref)[0]
                bad.append(dict(seed=seed, it=it, n=int(w.size), B=int(got.size), first_lane=int(w[0]), got=float(got[w[0]]), ref=float(ref[w[0]])))
print(json.dumps(dict(particles=P_PART, iters=ITERS, bad_batches=len(bad), first=bad[:8])))
