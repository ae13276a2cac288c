"""Developer check (profile build, KIMA_CUDA_LIB=variants/prof.so): per-CTA entry / exit times of gps_kernel by planet count."""
import ctypes, json, os, sys, io, contextlib
sys.path.insert(0, ".")
import numpy as np
from kima_b200 import LikelihoodEngine, ModelFlags, _lib
from kima_b200.synthetic import make_data, make_params
N, B = 177, 16384
d = make_data(N)
flags = ModelFlags(GP=True, npmax=5, hyperpriors=True)
P, nps_mix = make_params(flags, d, B, "mix")
lib = ctypes.CDLL(os.environ["KIMA_CUDA_LIB"])
for tag, nps in (("mix", nps_mix), ("sorted", np.sort(nps_mix)), ("all5", np.full(B, 5, np.int32))):
    with LikelihoodEngine(d, flags, B) as eng:
        eng.evaluate(P, nps)
        eng.evaluate(P, nps)
        ms = eng.last_kernel_ms()
        buf = np.zeros(3 * B, np.uint64)
        rc = lib.kc_prof_fetch_cta(buf.ctypes.data_as(ctypes.c_void_p), B)
    t0, t1, sm = buf[0::3].astype(np.int64), buf[1::3].astype(np.int64), buf[2::3].astype(np.int64)
    dur = (t1 - t0) / 1e3
    out = dict(tag=tag, ms=ms, rc=rc, mean_us=float(dur.mean()), by_np={int(k): float(dur[nps == k].mean()) for k in range(6)})
    # gaps between consecutive CTAs on the same SM
    gaps = []
    for s in range(int(sm.max()) + 1):
        idx = np.nonzero(sm == s)[0]
        o = np.argsort(t0[idx])
        g = t0[idx][o][1:] - t1[idx][o][:-1]
        gaps.append(g)
    gaps = np.concatenate(gaps) / 1e3
    out.update(gap_mean_us=float(gaps.mean()), gap_p50=float(np.median(gaps)), gap_p95=float(np.percentile(gaps, 95)), ctas_per_sm=float(B / (sm.max() + 1)),
               span_ms=float((t1.max() - t0.min()) / 1e6))
    print(json.dumps(out), flush=True)
