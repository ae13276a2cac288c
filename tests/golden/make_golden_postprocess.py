"""Pin tests/dnest_post.py to pykima's own post-processing (soft spot 1d).

Runs the reference's CPU sampler (host/_build/kima_cpu_BL2009, the unmodified example) for 3000
saves, keeps levels.txt and sample_info.txt (the only inputs of the evidence estimate) under
tests/golden/postprocess_run/, and stores what pykima.classic.postprocess
(/root/reference/pykima/classic.py:20-240; deterministic branch numResampleLogX=1, plot=False,
save=False) returns for them: log Z, the information H and the log X assigned to every sample.
pykima/classic.py is loaded by path inside a stand-in package (its __init__ pulls in matplotlib
and corner, absent here; classic.py itself only needs `plt` when plot=True).
Run in the build container:   python tests/golden/make_golden_postprocess.py
"""
import importlib.util
import json
import os
import shutil
import subprocess
import sys
import tempfile
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
HERE = os.path.dirname(os.path.abspath(__file__))
RUN = os.path.join(HERE, "postprocess_run")
PK = "/root/reference/pykima"
OPTIONS = "2\n500\n50\n100\n0\n10\n100\n3000\n"


def load_classic():
    for name in ("matplotlib", "matplotlib.pyplot"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    pkg = types.ModuleType("pykima_ref")
    pkg.__path__ = [PK]
    sys.modules["pykima_ref"] = pkg
    for sub in ("loading", "classic"):
        spec = importlib.util.spec_from_file_location("pykima_ref." + sub, os.path.join(PK, sub + ".py"))
        m = importlib.util.module_from_spec(spec)
        sys.modules["pykima_ref." + sub] = m
        spec.loader.exec_module(m)
    return sys.modules["pykima_ref.classic"]


def main():
    d = tempfile.mkdtemp(prefix="kima_post_")
    shutil.copy(os.path.join(HERE, "e2e", "BL2009_dataset2.kms.rv"), d)
    with open(os.path.join(d, "OPTIONS"), "w") as fh:
        fh.write(OPTIONS)
    subprocess.run([os.path.join(ROOT, "host", "_build", "kima_cpu_BL2009"), "-t", "4", "-s", "43"], cwd=d, check=True,
                   stdout=subprocess.DEVNULL)
    os.makedirs(RUN, exist_ok=True)
    for f in ("levels.txt", "sample_info.txt"):
        shutil.copy(os.path.join(d, f), RUN)
    classic = load_classic()
    cwd = os.getcwd()
    os.chdir(d)
    try:
        np.random.seed(0)  # only the resampling of posterior rows draws numbers; the estimates do not
        logz, H, logx = classic.postprocess(plot=False, save=False, verbose=False)
    finally:
        os.chdir(cwd)
    np.savez_compressed(os.path.join(RUN, "pykima_postprocess.npz"), logz=logz, H=H, logx=logx[:, 0])
    print("log Z = %.10f  H = %.10f  samples %d" % (logz, H, logx.shape[0]))
    shutil.rmtree(d, ignore_errors=True)


if __name__ == "__main__":
    main()
