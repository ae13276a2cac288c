"""Golden vectors for the posterior-predictive path, made by running
pykima's own keplerian() and GP.predict() (/root/reference/pykima/keplerian.py,
pykima/GP.py) in the build container:   python tests/golden/make_golden_predictive.py

pykima's modules are loaded by file path (importing the package pulls in matplotlib and
corner; GP.py wants `celerite` at import time, which is absent here and unused by the
standard-kernel classes, so an empty stand-in module is registered for the import only).
The GP object is given the C++ likelihood's variances sig^2 + jitter^2 as `yerr` and
white_noise 0, so that K + diag(yerr) is calculate_C's matrix (src/RVmodel.cpp:186-216).
"""
import importlib.util
import json
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from kima_b200.flags import ModelFlags, layout_from_flags  # noqa: E402
from kima_b200.synthetic import make_data, make_params  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
PK = "/root/reference/pykima"


def load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def main():
    stub = types.ModuleType("celerite")
    stub.GP = object
    stub.terms = types.SimpleNamespace(Term=object)
    sys.modules["celerite"] = stub
    kep = load("pk_keplerian", os.path.join(PK, "keplerian.py"))
    gpm = load("pk_GP", os.path.join(PK, "GP.py"))

    cases = [("predictive_gp_N60", ModelFlags(GP=True, npmax=2), 60, 6, 150),
             ("predictive_gp_N300_trend", ModelFlags(GP=True, npmax=3, trend=True, degree=2), 300, 4, 700),
             ("predictive_kep_N120_hyper", ModelFlags(npmax=4, hyperpriors=True), 120, 8, 333)]
    for name, flags, N, B, nt in cases:
        d = make_data(N)
        P, nps = make_params(flags, d, B, "mix", seed=7)
        L = layout_from_flags(flags, d.number_instruments, d.number_indicators)
        # keep the eccentricities where pykima's Newton iteration converges to ~1e-10
        P[L["row_ecc"]:L["row_ecc"] + flags.npmax] = np.minimum(P[L["row_ecc"]:L["row_ecc"] + flags.npmax], 0.8)
        tg = np.linspace(d.t.min() - 30.0, d.t.max() + 30.0, nt)
        model_grid = np.zeros((B, nt))
        model_data = np.zeros((B, N))
        gp_grid = np.zeros((B, nt))
        gp_std = np.zeros((B, nt))
        for b in range(B):
            v = np.zeros(nt)
            vd = np.zeros(N)
            for j in range(int(nps[b])):
                per = P[L["row_P"] + j, b]
                if flags.hyperpriors:
                    per = np.exp(per)
                K, phi, ecc, w = (P[L[r] + j, b] for r in ("row_K", "row_phi", "row_ecc", "row_w"))
                t0 = d.M0_epoch - (per * phi) / (2.0 * np.pi)  # results.py:770
                v += kep.keplerian(tg, per, K, ecc, w, t0, 0.0)
                vd += kep.keplerian(d.t, per, K, ecc, w, t0, 0.0)
            vsys = P[L["row_vsys"], b]
            v += vsys
            vd += vsys
            if flags.trend:
                tp = np.r_[[P[L["row_trend"] + k, b] for k in range(flags.degree)][::-1], 0.0]
                v += np.polyval(tp, tg - d.get_t_middle())        # results.py:790-795
                vd += np.polyval(tp, d.t - d.get_t_middle())
            model_grid[b], model_data[b] = v, vd
            if flags.GP:
                eta = [P[L["row_eta"] + k, b] for k in range(4)]
                jit = P[L["row_jit"], b]
                g = gpm.GP(gpm.QPkernel(*eta), d.t, d.sig ** 2 + jit ** 2, white_noise=0.0)
                gp_grid[b] = g.predict(d.y - vd, tg, return_std=False)
                gp_std[b] = g.predict(d.y - vd, tg, return_std=True)[1]   # GP.py:113-131
        np.savez_compressed(os.path.join(HERE, name + ".npz"), t=d.t, y=d.y, sig=d.sig, params=P, nps=nps,
                            tgrid=tg, model_grid=model_grid, model_data=model_data, gp_grid=gp_grid, gp_std=gp_std,
                            flags=json.dumps({k: int(getattr(flags, k)) for k in flags.__dataclass_fields__}))
        print(name, "N=%d B=%d nt=%d  max|model|=%.3g max|gp|=%.3g" % (N, B, nt, np.abs(model_grid).max(), np.abs(gp_grid).max()))


def celerite_case():
    """kernel=celerite: the conditional mean of the reference's OWN solver (CholeskySolver::predict through
    oracle/_ref/libkima_ref.so, oracle/ref_harness.cpp: ref_celerite_predict) -- pykima's GP_celerite wraps the
    Python celerite package, which is not installed here; the C++ solver is the code the likelihood runs."""
    from oracle.oracle import Oracle
    ref = Oracle("reference")
    for name, flags, N, B, nt in (("predictive_celerite_N200", ModelFlags(GP=True, kernel=1, npmax=2), 200, 6, 450),
                                  ("predictive_celerite_N90_trend", ModelFlags(GP=True, kernel=1, npmax=1, trend=True, degree=1),
                                   90, 4, 77)):
        d = make_data(N)
        P, nps = make_params(flags, d, B, "mix", seed=9)
        ref.set_rvdata(d)
        ref.set_flags(flags)
        tg = np.linspace(d.t.min() - 25.0, d.t.max() + 25.0, nt)
        gp_grid = np.array([ref.celerite_predict(P, b, int(nps[b]), tg) for b in range(B)])
        gp_data = np.array([ref.celerite_predict(P, b, int(nps[b]), d.t) for b in range(B)])
        np.savez_compressed(os.path.join(HERE, name + ".npz"), t=d.t, y=d.y, sig=d.sig, params=P, nps=nps, tgrid=tg,
                            gp_grid=gp_grid, gp_data=gp_data,
                            flags=json.dumps({k: int(getattr(flags, k)) for k in flags.__dataclass_fields__}))
        print(name, "N=%d B=%d nt=%d  max|gp|=%.3g" % (N, B, nt, np.abs(gp_grid).max()))


if __name__ == "__main__":
    if "celerite" not in sys.argv[1:]:
        main()
    celerite_case()
