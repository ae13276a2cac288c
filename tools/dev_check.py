"""Developer check on a GPU box: parity against the CPU oracle + kernel timings."""
import sys, time, json
sys.path.insert(0, ".")
import numpy as np
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params
from oracle.oracle import Oracle

port = Oracle("port")
def relerr(a, b):
    return np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300))

def case(name, flags, N, ninst=1, nind=0, B=64, np_mode="mix", mode=0, time_B=0, nthreads=16):
    d = make_data(N, ninst, nind)
    P, nps = make_params(flags, d, B, np_mode)
    port.set_rvdata(d); port.set_flags(flags)
    ref, secs = port.eval_batch(P, nps, nthreads)
    eng = LikelihoodEngine(d, flags, max(B, time_B), kepler_mode=mode)
    got = eng.evaluate(P, nps)
    r = relerr(got, ref)
    out = dict(case=name, mode=mode, B=B, relerr=float(r), cpu_evals_s=B / secs)
    if time_B:
        P2, nps2 = make_params(flags, d, time_B, "max")
        eng.evaluate(P2, nps2)
        for _ in range(3):
            eng.log_likelihood_device(time_B)
        eng.sync()
        ms = []
        for _ in range(5):
            eng.log_likelihood_device(time_B); eng.sync(); ms.append(eng.last_kernel_ms())
        out["kernel_ms"] = min(ms); out["evals_s"] = time_B / (min(ms) * 1e-3)
        t0 = time.perf_counter(); got2 = eng.log_likelihood(time_B); t1 = time.perf_counter()
        out["e2e_evals_s"] = time_B / (t1 - t0)
    eng.close()
    print(json.dumps(out), flush=True)
    return out

if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "kep"):
        for mode in (0, 1):
            case("kep_gauss_N1000_np10", ModelFlags(npmax=10), 1000, B=256, mode=mode, time_B=8192)
        case("kep_small_batch", ModelFlags(npmax=10), 1000, B=64, time_B=512)
        case("kep_student_multi_trend_ko_ind", ModelFlags(npmax=3, studentt=True, multi_instrument=True, trend=True,
             degree=3, known_object=True, n_known_object=2, hyperpriors=True), 200, 3, 1, B=256, time_B=8192)
        case("kep_51peg_like", ModelFlags(npmax=1), 256, B=256, time_B=8192)
    if which in ("all", "gp"):
        case("gp_N177_np5", ModelFlags(GP=True, npmax=5, hyperpriors=True), 177, B=64, time_B=1024)
        case("gp_N1000_np10", ModelFlags(GP=True, npmax=10), 1000, B=32, time_B=592)
        case("gp_N50", ModelFlags(GP=True, npmax=2), 50, B=16)
        case("gp_N33_multi", ModelFlags(GP=True, npmax=2, multi_instrument=True), 33, 3, 0, B=16)
