"""Golden case for the ill-conditioned tail of the Keplerian likelihood (soft spot 1b).

The reference turns the eccentric anomaly into the true anomaly through acos
(/root/reference/src/RVmodel.cpp:1467); near f = 0, pi its result carries an error of
about eps / |sin f|, which K |sin(f + w)| passes on to mu and (y - mu) / var to logL.  Over
65 536 prior draws one lane in 65 536 differed from the GPU by more than 1e-10 relative
(profiles/parity_tail_r01.jsonl: case N200_np5_hyper, both Kepler modes — it is the
reference's own conditioning, not the GPU's).  This script draws the same batch, computes for
every lane the reference-derived bound

    bound_b = sum_i |r_i| / var_i * sum_p K_p |sin(f_ip + w_p)| * 4 eps / max(|sin f_ip|, 1e-9)

from the reference's own f, mu and variances, and commits the 48 lanes with the largest
bound / |logL| together with their reference logL (libkima_ref.so) and the bound.
Run in the build container:   python tests/golden/make_golden_tail.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from kima_b200.flags import ModelFlags  # noqa: E402
from kima_b200.synthetic import make_data, make_params  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def conditioning_bound(ref, flags, d, P, nps):
    """The bound above for every lane of the block P (vectorised over lanes and epochs)."""
    L = ref.layout()
    B = P.shape[1]
    N = d.N()
    t = np.tile(d.t, B)
    acc = np.zeros(B * N)
    for p in range(flags.npmax):
        per = P[L["row_P"] + p]
        if flags.hyperpriors:
            per = np.exp(per)
        K, phi, ecc, w = (P[L[r] + p] for r in ("row_K", "row_phi", "row_ecc", "row_w"))
        Tp = d.M0_epoch - (per * phi) / (2.0 * np.pi)
        f = ref.true_anomaly(t, np.repeat(per, N), np.repeat(ecc, N), np.repeat(Tp, N))
        live = np.repeat(p < nps, N)
        df = 4e-16 / np.maximum(np.abs(np.sin(f)), 1e-9)
        acc += np.where(live, np.repeat(K, N) * np.abs(np.sin(f + np.repeat(w, N))) * df, 0.0)
    out = np.zeros(B)
    jit = P[L["row_jit"]]
    for b in range(B):
        mu = ref.mu(P, b, int(nps[b]))
        var = d.sig ** 2 + jit[b] ** 2
        out[b] = np.sum(np.abs(d.y - mu) / var * acc[b * N:(b + 1) * N])
    return out


def main():
    ref = Oracle("reference")
    flags = ModelFlags(npmax=5, hyperpriors=True)
    d = make_data(200)
    B = 65536
    P, nps = make_params(flags, d, B, "max", seed=7)   # the batch of tools/dev_tail.py
    ref.set_rvdata(d)
    ref.set_flags(flags)
    logL, _ = ref.eval_batch(P, nps, os.cpu_count() or 8)
    bound = conditioning_bound(ref, flags, d, P, nps)
    score = bound / np.abs(logL)
    pick = np.argsort(-score)[:48]
    print("lanes with bound > 1e-10 |logL|:", int((score > 1e-10).sum()), " > 1e-11:", int((score > 1e-11).sum()),
          " worst %.3g" % score.max())
    np.savez_compressed(os.path.join(HERE, "tail_kep_N200_np5_hyper.npz"), params=P[:, pick], nps=nps[pick],
                        logL=logL[pick], bound=bound[pick], lanes=pick,
                        flags=json.dumps({k: int(getattr(flags, k)) for k in flags.__dataclass_fields__}))


if __name__ == "__main__":
    main()
