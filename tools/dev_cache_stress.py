"""Developer check on a GPU box: random accept / reject histories through the factor cache, every batch compared
bit for bit with an uncached context; two cached contexts driven from two host threads (the host sampler's
two lane groups)."""
import json, sys, threading
sys.path.insert(0, ".")
import numpy as np
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params

N = int(sys.argv[1]) if len(sys.argv) > 1 else 177
P_PART = int(sys.argv[2]) if len(sys.argv) > 2 else 2048     # particles per context
ITERS = int(sys.argv[3]) if len(sys.argv) > 3 else 300
flags = ModelFlags(GP=True, npmax=5, hyperpriors=True)
data = make_data(N)
bad = []


def drive(seed):
    rng = np.random.default_rng(seed)
    state, nps_all = make_params(flags, data, P_PART, "mix", seed=seed)
    with LikelihoodEngine(data, flags, P_PART) as eng, LikelihoodEngine(data, flags, P_PART) as plain:
        L = eng.layout
        assert eng.enable_gp_cache(P_PART) == P_PART
        cur = plain.evaluate(state, nps_all)
        for it in range(ITERS):
            # the lanes of this step: a random subset of the particles, in random order
            sel = rng.permutation(P_PART)[: rng.integers(P_PART // 4, P_PART)]
            prop = state[:, sel].copy()
            kind = rng.random(sel.size)
            n = max(L["npmax"], 1)
            prop[L["row_K"]:L["row_K"] + n] *= rng.uniform(0.9, 1.1, (n, sel.size))
            prop[L["row_vsys"]] += rng.normal(0, 0.3, sel.size)
            ch = kind < 0.375                                   # these proposals change C
            which = rng.integers(0, 5, sel.size)
            for r in range(4):
                m = ch & (which == r)
                prop[L["row_eta"] + r, m] *= rng.uniform(0.95, 1.05, m.sum())
            m = ch & (which == 4)
            prop[L["row_jit"], m] *= rng.uniform(0.9, 1.1, m.sum())
            eng.set_slots(sel.astype(np.int32))
            got = eng.evaluate(prop, nps_all[sel])
            ref = plain.evaluate(prop, nps_all[sel])
            if not np.array_equal(got, ref):
                w = np.nonzero(got != ref)[0]
                bad.append(dict(seed=seed, it=it, n=int(w.size), lane=int(w[0]), slot=int(sel[w[0]]), got=float(got[w[0]]), ref=float(ref[w[0]]),
                                changed_C=bool(ch[w[0]])))
            # accept about a third; now and then copy a particle over another (kill_lagging_particles)
            acc = rng.random(sel.size) < 0.35
            state[:, sel[acc]] = prop[:, acc]
            if it % 7 == 0:
                a, b = rng.integers(0, P_PART, 2)
                state[:, a] = state[:, b]
                nps_all[a] = nps_all[b]


ts = [threading.Thread(target=drive, args=(s,)) for s in (11, 12)]
for t in ts: t.start()
for t in ts: t.join()
print(json.dumps(dict(N=N, particles=P_PART, iters=ITERS, mismatches=len(bad), first=bad[:5])))
