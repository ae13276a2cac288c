import sys
sys.path.insert(0, ".")
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params
N = int(sys.argv[1]) if len(sys.argv) > 1 else 177
B = int(sys.argv[2]) if len(sys.argv) > 2 else 148
flags = ModelFlags(GP=True, npmax=5)
d = make_data(N)
P, nps = make_params(flags, d, B, "max")
with LikelihoodEngine(d, flags, B) as eng:
    eng.evaluate(P, nps)
