"""Developer check: does the shared-memory-resident GP kernel return the same bits every time?

The same device-resident batch is scored over and over on two contexts whose streams overlap (the pattern of
the host-proposal sampler's two lane groups); every result row is compared on the device with the first one.
    python tools/dev_determinism.py N B seconds [host]
`host`: go through kc_eval (pinned staging block, H2D, D2H) instead of kc_eval_device.
"""
import ctypes, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params

N, B, seconds = int(sys.argv[1]), int(sys.argv[2]), float(sys.argv[3])
host = len(sys.argv) > 4 and sys.argv[4] == "host"
flags = ModelFlags(GP=True, npmax=5, hyperpriors=True)
data = make_data(N)
engs, refs, Ps = [], [], []
for g in range(2):
    P, nps = make_params(flags, data, B, "mix", seed=100 + g)
    e = LikelihoodEngine(data, flags, B)
    refs.append(e.evaluate(P, nps))   # leaves the batch in the device block
    engs.append(e); Ps.append((P, nps))
ROWS = 32
dp = ctypes.POINTER(ctypes.c_double)
t0 = time.time(); reps = 0; bad = 0
if not host:
    bufs = [torch.zeros(ROWS, B, dtype=torch.float64, device="cuda") for _ in range(2)]
    dref = [torch.from_numpy(r).cuda() for r in refs]
    while time.time() - t0 < seconds:
        for r in range(ROWS):
            for g in range(2):
                ptr = ctypes.cast(bufs[g][r].data_ptr(), dp)
                rc = engs[g].lib.kc_eval_device(engs[g]._ctx, B, ptr)
                assert rc == 0
        for g in range(2):
            engs[g].sync()
        reps += ROWS
        for g in range(2):
            # bit compare (NaN-safe): view as int64
            neq = bufs[g].view(torch.int64) != dref[g].view(torch.int64)[None, :]
            if bool(neq.any()):
                idx = neq.nonzero().cpu().numpy()
                for (r, b) in idx[:8]:
                    bad += 1
                    print("MISMATCH ctx %d rep %d lane %d np %d: %.17g vs %.17g" % (g, reps - ROWS + r, b, Ps[g][1][b],
                          float(bufs[g][r, b]), refs[g][b]), flush=True)
else:
    outs = [np.empty(B) for _ in range(2)]
    while time.time() - t0 < seconds:
        for g in range(2):
            engs[g].log_likelihood_async(B, outs[g])
        for g in range(2):
            engs[g].wait()
            neq = outs[g].view(np.int64) != refs[g].view(np.int64)
            if neq.any():
                for b in np.nonzero(neq)[0][:8]:
                    bad += 1
                    print("MISMATCH ctx %d rep %d lane %d np %d: %.17g vs %.17g" % (g, reps, b, Ps[g][1][b], outs[g][b], refs[g][b]), flush=True)
        reps += 1
el = time.time() - t0
print({"N": N, "B": B, "host": host, "reps_per_ctx": reps, "evals": 2 * reps * B, "mismatches": bad, "evals_per_s": 2 * reps * B / el})
