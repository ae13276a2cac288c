"""Generate the committed golden vectors from the UNMODIFIED reference.

Run in the build container only (needs /root/reference and
oracle/_ref/libkima_ref.so):   python tests/golden/make_golden.py

Every fixture holds inputs (data arrays as left behind by the reference's own
file parsers, flags, parameter block) and the outputs of the reference's
calculate_mu / calculate_C / log_likelihood / kepler::* on them.  The fixtures
travel to the GPU box; /root/reference does not.
"""
import json
import math
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from kima_b200.data import RVData  # noqa: E402
from kima_b200.flags import ModelFlags  # noqa: E402
from kima_b200.synthetic import make_data, make_params  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
EX = "/root/reference/examples"


def rvdata_from_ref(d: dict) -> RVData:
    multi = d["n_inst"] > 1
    return RVData.from_arrays(d["t"], d["y"], d["sig"], d["obsi"] if multi else None,
                              d["actind"] if d["n_ind"] else None, n_inst=d["n_inst"])


def flags_dict(f: ModelFlags) -> dict:
    return {k: int(getattr(f, k)) for k in f.__dataclass_fields__}


ONLY = set(sys.argv[1:])  # `make_golden.py name ...` regenerates just those fixtures


def save_case(ref: Oracle, name: str, flags: ModelFlags, data: RVData, B: int, np_mode: str,
              seed: int, with_cov: bool = False, load_info=None, edge=None):
    if ONLY and name not in ONLY:
        return
    ref.set_rvdata(data)
    ref.set_flags(flags)
    P, nps = make_params(flags, data, B, np_mode, seed=seed)
    if edge is not None:
        edge(P, nps, ref.layout())
    logL, _ = ref.eval_batch(P, nps)
    mu0, mu0_lo = ref.mu(P, 0, int(nps[0]), with_lo=True)
    out = dict(t=data.t, y=data.y, sig=data.sig, obsi=data.obsi, actind=data.actind,
               n_inst=data.number_instruments, params=P, nps=nps, logL=logL, mu0=mu0,
               mu0_lo=mu0_lo, flags=json.dumps(flags_dict(flags)),
               layout=json.dumps(ref.layout()), load_info=json.dumps(load_info or {}))
    if with_cov:
        out["C0"] = ref.cov(P, 0)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print("%-28s N=%4d B=%3d  logL[0]=%.6f  finite=%d" % (name, data.N(), B, logL[0],
                                                          int(np.isfinite(logL).sum())))


def kepler_vectors(ref):
    # ---- Kepler known answers -------------------------------------------------------
    rng = np.random.default_rng(12345)
    n = 20000
    M = np.concatenate([rng.uniform(0, 2 * np.pi, n), rng.uniform(-100, 1e5, n)])
    e = np.concatenate([rng.uniform(0, 1, n), rng.choice([0, 1e-12, 0.5, 0.9, 0.99, 1 - 1e-9], n)])
    # edge grid of SURVEY 8(d): M at multiples of pi, the listed eccentricities
    Mg, eg = np.meshgrid(np.array([0, 0.5, 1, 1.5, 2, 3, 7]) * np.pi,
                         np.array([0, 1e-12, 0.2, 0.5, 0.9, 0.99, 1 - 1e-9]))
    M = np.concatenate([M, Mg.ravel(), [math.radians(80.0)]])
    e = np.concatenate([e, eg.ravel(), [0.2]])
    E = ref.kepler(M, e)
    t = 50000 + rng.uniform(0, 2000, 5000)
    P = np.exp(rng.uniform(0, np.log(1e5), 5000))
    ee = rng.uniform(0, 1, 5000)
    Tp = 50000 - P * rng.uniform(0, 1, 5000)
    f = ref.true_anomaly(t, P, ee, Tp)
    np.savez_compressed(os.path.join(HERE, "kepler.npz"), M=M, e=e, E=E, t=t, P=P, ecc=ee, Tp=Tp, f=f)
    print("kepler.npz: %d (M,e) pairs, %d true anomalies" % (M.size, t.size))



def main():
    ref = Oracle("reference")

    if not ONLY or "kepler" in ONLY:
        kepler_vectors(ref)
    # ---- the example data sets, through the reference's own loaders ------------------
    def ex(path, *a, multi=False, **k):
        d = ref.ref_load_multi(path, *a) if multi else ref.ref_load(path, *a, **k)
        return rvdata_from_ref(d)

    d51 = ex(EX + "/51Peg/51Peg.rv", "ms", 0)
    save_case(ref, "51Peg", ModelFlags(npmax=1), d51, 32, "mix", 43,
              load_info=dict(file="examples/51Peg/51Peg.rv", units="ms", skip=0))
    dbl = ex(EX + "/BL2009/BL2009_dataset2.kms.rv", "kms", 0)
    save_case(ref, "BL2009", ModelFlags(npmax=5), dbl, 32, "mix", 44,
              load_info=dict(file="examples/BL2009/BL2009_dataset2.kms.rv", units="kms", skip=0))
    dmp = ex(EX + "/many_planets/HD10180.kms.rv", "kms", 1)
    save_case(ref, "many_planets", ModelFlags(npmax=10, hyperpriors=True), dmp, 32, "mix", 45,
              load_info=dict(file="examples/many_planets/HD10180.kms.rv", units="kms", skip=1))
    dst = ex(EX + "/studentT/d1.txt", "ms", 2)
    save_case(ref, "studentT", ModelFlags(npmax=1, studentt=True), dst, 32, "max", 46,
              load_info=dict(file="examples/studentT/d1.txt", units="ms", skip=2))
    files = [EX + "/multi_instrument/HD106252_%s.txt" % s for s in ("ELODIE", "HET", "HJS", "Lick")]
    dmi = rvdata_from_ref(ref.ref_load_multi(files, "ms", 2))
    save_case(ref, "multi_instrument", ModelFlags(npmax=1, multi_instrument=True), dmi, 32, "mix", 47,
              load_info=dict(files=["examples/multi_instrument/" + os.path.basename(f) for f in files],
                             units="ms", skip=2))
    # BASELINE.json's many_planets variant: Student-t + multi_instrument, Np up to 10
    save_case(ref, "many_planets_studentt_multi",
              ModelFlags(npmax=10, hyperpriors=True, studentt=True, multi_instrument=True), dmi, 32,
              "mix", 48)
    dmj = rvdata_from_ref(ref.ref_load_multi(EX + "/multi_instrument/HD106252_joined.txt", "ms", 1))
    save_case(ref, "multi_instrument_joined", ModelFlags(npmax=2, multi_instrument=True), dmj, 16,
              "mix", 49, load_info=dict(file="examples/multi_instrument/HD106252_joined.txt",
                                        units="ms", skip=1))
    dtr = ex(EX + "/trends/d1.txt", "ms", 2)
    save_case(ref, "trends", ModelFlags(npmax=0, trend=True, degree=3), dtr, 16, "max", 50,
              load_info=dict(file="examples/trends/d1.txt", units="ms", skip=2))
    dac = rvdata_from_ref(ref.ref_load(EX + "/activity_correlations/simulated2.txt", "ms", 4,
                                       ["b", "", "indc"]))
    save_case(ref, "activity_correlations", ModelFlags(npmax=2), dac, 16, "mix", 51,
              load_info=dict(file="examples/activity_correlations/simulated2.txt", units="ms",
                             skip=4, indicators=["b", "", "indc"]))
    dco = ex(EX + "/CoRoT7/corot7.txt", "ms", 2)
    save_case(ref, "CoRoT7", ModelFlags(GP=True, npmax=5, hyperpriors=True), dco, 16, "mix", 52,
              load_info=dict(file="examples/CoRoT7/corot7.txt", units="ms", skip=2))

    # ---- synthetic shapes of SURVEY 8(d) ------------------------------------------------
    save_case(ref, "synth_kep_N1000_np10", ModelFlags(npmax=10), make_data(1000), 16, "max", 43)
    save_case(ref, "synth_kep_N1000_np10_studentt", ModelFlags(npmax=10, studentt=True),
              make_data(1000), 8, "mix", 53)
    save_case(ref, "synth_gp_N1000_np10", ModelFlags(GP=True, npmax=10), make_data(1000), 4, "max", 43)
    save_case(ref, "synth_gp_N33_multi", ModelFlags(GP=True, npmax=2, multi_instrument=True),
              make_data(33, 3), 16, "mix", 54, with_cov=True)
    save_case(ref, "synth_gp_N130_trend_ko", ModelFlags(GP=True, npmax=3, trend=True, degree=2,
                                                        known_object=True, n_known_object=1),
              make_data(130), 8, "mix", 55)
    save_case(ref, "synth_all_flags", ModelFlags(npmax=3, studentt=True, multi_instrument=True,
                                                 trend=True, degree=3, known_object=True,
                                                 n_known_object=2, hyperpriors=True),
              make_data(200, 3, 2), 32, "mix", 56)

    # BASELINE.json configs[4], the scale point: N=4096, QP-GP + Np<=10 (12 s per lane on the CPU)
    save_case(ref, "synth_gp_N4096", ModelFlags(GP=True, npmax=10), make_data(4096), 4, "mix", 11)

    # ---- MA(1) term (src/RVmodel.cpp:378-386) ---------------------------------------------
    save_case(ref, "synth_ma_N200", ModelFlags(npmax=3, MA=True), make_data(200), 16, "mix", 58)
    save_case(ref, "synth_ma_studentt_multi", ModelFlags(npmax=2, MA=True, studentt=True, multi_instrument=True,
                                                         trend=True, degree=2),
              make_data(150, 3, 1), 16, "mix", 59)
    save_case(ref, "synth_gp_ma_N100", ModelFlags(GP=True, npmax=2, MA=True), make_data(100), 8, "mix", 60)

    # ---- kernel = celerite (src/RVmodel.cpp:218-266, 875-882) --------------------------------
    save_case(ref, "synth_celerite_N300", ModelFlags(GP=True, kernel=1, npmax=3), make_data(300), 16, "mix", 61)
    save_case(ref, "CoRoT7_celerite", ModelFlags(GP=True, kernel=1, npmax=5, hyperpriors=True), dco, 16, "mix", 62,
              load_info=dict(file="examples/CoRoT7/corot7.txt", units="ms", skip=2))
    save_case(ref, "synth_celerite_ma_trend", ModelFlags(GP=True, kernel=1, npmax=2, MA=True, trend=True, degree=2),
              make_data(120, 1, 1), 8, "mix", 63)

    # ---- edge cases: extreme eccentricities / periods, non-finite outputs ----------------
    def edge(P, nps, L):
        B = P.shape[1]
        ecc = [0.0, 1e-12, 0.5, 0.9, 0.99, 1 - 1e-9]
        for b in range(B):
            P[L["row_ecc"]:L["row_ecc"] + L["npmax"], b] = ecc[b % len(ecc)]
        P[L["row_P"], 0:4] = [0.011, 1.0, 3e5, 1e7]  # P << dt, P >> timespan
        P[L["row_phi"], 4:8] = [0.0, np.pi, 2 * np.pi, 1e-300]
        P[L["row_jit"], 8] = 0.0                      # var = sig^2 exactly
        P[L["row_vsys"], 9] = np.nan                  # NaN -> +inf (src/RVmodel.cpp:938-941)
        P[L["row_K"], 10] = np.inf
        nps[11] = 0
    save_case(ref, "synth_edge", ModelFlags(npmax=2), make_data(64), 16, "max", 57, edge=edge)


if __name__ == "__main__":
    main()
