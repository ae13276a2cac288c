"""Build libkima_cuda.so (sm_100a) in-tree with nvcc.  No JIT, no torch extension:
the product is a plain C-ABI shared library (include/kima_cuda.h).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libkima_cuda.so")
HOSTCHECK = os.path.join(HERE, "libkc_hostcheck.so")

CU_SOURCES = ["kc_api.cu", "kc_keplerian.cu", "kc_gp.cu", "kc_gp_blocked.cu", "kc_gp_small.cu", "kc_celerite.cu", "kc_predict.cu", "kc_comm.cu", "kc_peak.cu", "kc_sampler.cu"]
NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-Xcompiler", "-fPIC", "-Xptxas", "-v", "-diag-suppress", "550",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libkima_cuda cannot be built")


def _newer(target: str, deps) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(d) <= t for d in deps)


def _deps():
    d = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    for h in ("kima_cuda.h", "kima_sampler.h", "kc_philox.h"):
        d.append(os.path.join(HERE, "..", "include", h))
    d.append(os.path.abspath(__file__))
    return d


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every CUDA translation unit for sm_100a and link the library."""
    deps = _deps()
    if not force and _newer(LIB, deps) and _newer(HOSTCHECK, deps):
        return LIB
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    log_path = os.path.join(OBJ, "ptxas.log")

    def compile_one(src):
        obj = os.path.join(OBJ, src.replace(".cu", ".o"))
        cmd = [nvcc] + NVCC_FLAGS + os.environ.get("KC_NVCC_EXTRA", "").split() + ["-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed on %s:\n%s\n%s" % (src, r.stdout, r.stderr))
        return obj, r.stderr

    with ThreadPoolExecutor(max_workers=len(CU_SOURCES)) as ex:
        results = list(ex.map(compile_one, CU_SOURCES))
    with open(log_path, "w") as fh:
        for obj, log in results:
            fh.write("==== %s\n%s\n" % (os.path.basename(obj), log))
            if verbose:
                sys.stderr.write(log)
    objs = [o for o, _ in results]
    cmd = [nvcc, "-shared", "-o", LIB] + objs + ["-ldl"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    # host instantiation of the device math headers, for CPU unit tests only
    cxx = shutil.which("g++") or "g++"
    r = subprocess.run([cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-o", HOSTCHECK,
                        os.path.join(CSRC, "host_check.cpp"), os.path.join(CSRC, "host_check_sampler.cpp")],
                       capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("host_check build failed:\n%s\n%s" % (r.stdout, r.stderr))
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
