"""One warm evaluation + one measured evaluation of the GP workload (for ncu)."""
import sys
sys.path.insert(0, ".")
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1184
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
flags = ModelFlags(GP=True, npmax=10)
d = make_data(N)
P, nps = make_params(flags, d, B, "max")
with LikelihoodEngine(d, flags, B) as eng:
    eng.evaluate(P, nps)
    for _ in range(reps):
        eng.flush_l2(); eng.log_likelihood_device(B); eng.sync()
    print("ms", eng.last_kernel_ms())
