"""Developer check on a GPU box: the blocked GP path against the CPU oracle + timings."""
import sys, json, os
sys.path.insert(0, ".")
import numpy as np
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params
from oracle.oracle import Oracle

port = Oracle("port")
def relerr(a, b):
    return np.abs(a - b) / np.maximum(np.abs(b), 1e-300)

def parity(N, B, flags, path, ninst=1, seed=7):
    d = make_data(N, ninst, 0, seed=100 + N)
    P, nps = make_params(flags, d, B, "mix", seed=seed)
    port.set_rvdata(d); port.set_flags(flags)
    ref, _ = port.eval_batch(P, nps, os.cpu_count() or 8)
    with LikelihoodEngine(d, flags, B) as eng:
        got = eng.evaluate(P, nps)
    fin = np.isfinite(ref)
    ok_inf = bool(np.array_equal(np.isfinite(got), fin))
    r = relerr(got[fin], ref[fin])
    print(json.dumps(dict(kind="parity", N=N, B=B, path=path, ninst=ninst, max=float(r.max()) if r.size else 0.0,
                          med=float(np.median(r)) if r.size else 0.0, inf_match=ok_inf)), flush=True)

def timing(N, B, path, slots=0, reps=5):
    flags = ModelFlags(GP=True, npmax=10)
    d = make_data(N)
    P, nps = make_params(flags, d, B, "max")
    with LikelihoodEngine(d, flags, B, gp_slots=slots) as eng:
        eng.evaluate(P, nps)
        ms = []
        for _ in range(reps):
            eng.flush_l2(); eng.log_likelihood_device(B); eng.sync(); ms.append(eng.last_kernel_ms())
    F = N ** 3 / 3 + 37.0 * N * N + N * (330.0 * 10 + 45)
    best = min(ms)
    print(json.dumps(dict(kind="timing", N=N, B=B, path=path, ms=best, evals_s=B / best * 1e3,
                          tflops=B / best * 1e3 * F * 1e-12)), flush=True)

if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "parity"):
        for N in (3, 31, 63, 64, 65, 127, 128, 129, 177, 191, 192, 193, 257, 500, 1000):
            parity(N, 24, ModelFlags(GP=True, npmax=2), 2)
        parity(161, 24, ModelFlags(GP=True, npmax=2, multi_instrument=True, trend=True, degree=2), 2, ninst=3)
        parity(1000, 64, ModelFlags(GP=True, npmax=10), 2)
    if which in ("all", "timing"):
        timing(177, 2048, 2)
        timing(1000, 1184, 2)
        timing(384, 2048, 2)
        timing(4096, 148, 2, reps=2)
    if which == "streams":
        for N, B in ((1000, 1184), (4096, 148), (384, 2048)):
            for ns in (1, 2, 3, 4):
                flags = ModelFlags(GP=True, npmax=10)
                d = make_data(N)
                P, nps = make_params(flags, d, B, "max")
                with LikelihoodEngine(d, flags, B, gp_streams=ns) as eng:
                    res = eng.evaluate(P, nps)
                    ms = []
                    for _ in range(4):
                        eng.flush_l2(); eng.log_likelihood_device(B); eng.sync(); ms.append(eng.last_kernel_ms())
                print(json.dumps(dict(N=N, B=B, streams=ns, ms=min(ms), evals_s=B / min(ms) * 1e3, chk=float(np.sum(res)))), flush=True)
