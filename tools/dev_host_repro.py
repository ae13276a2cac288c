"""Developer check: is the host-proposal sampler reproducible run to run when the GPU is fast (51Peg, 4096 lanes)?"""
import filecmp, os, shutil, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from dev_cache_sampler import one, FILES
name, threads, ts, saves, reps = sys.argv[1], int(sys.argv[2]), 100, int(sys.argv[3]), int(sys.argv[4])
env = {"KIMA_CUDA_GP_CACHE": sys.argv[5]} if len(sys.argv) > 5 else {}
env["KIMA_TRACE_HASH"] = "trace.txt"   # host/RVmodelDevice.h: one line per batch, hashes of the inputs and of the results
base = None
for i in range(reps):
    d, res = one(name, threads, ts, saves, env)
    if base is None:
        base = d; print("run 0", res); continue
    diff = [f for f in FILES if not filecmp.cmp(os.path.join(base, f), os.path.join(d, f), shallow=False)]
    print("run", i, "differs in", diff, res.get("steps_per_s_run"))
    if diff:
        ta = open(os.path.join(base, "trace.txt")).read().split("\n"); tb = open(os.path.join(d, "trace.txt")).read().split("\n")
        first = next((j for j in range(min(len(ta), len(tb))) if ta[j] != tb[j]), None)
        print("   trace lines", len(ta), len(tb), "first differing", first)
        if first is not None:
            for j in range(max(0, first - 1), min(len(ta), first + 4)): print("     ", ta[j], "|", tb[j])
            fa, fb = ta[first].split(), tb[first].split()
            print("   -> inputs differ first" if fa[3] != fb[3] else "   -> SAME inputs, results differ (device)")
    for f in diff:
        a = open(os.path.join(base, f)).read().split("\n"); b = open(os.path.join(d, f)).read().split("\n")
        if f == "sampler_state.txt":
            a = a[0].split() if len(a) < 3 else " ".join(a).split(); b = b[0].split() if len(b) < 3 else " ".join(b).split()
            idx = [j for j in range(min(len(a), len(b))) if a[j] != b[j]]
            print("   state tokens", len(a), len(b), "differing", len(idx), "first", [(j, a[j], b[j]) for j in idx[:12]])
        else:
            idx = [j for j in range(min(len(a), len(b))) if a[j] != b[j]]
            print("  ", f, "lines differing", len(idx))
            for j in idx[:6]: print("     ", j, "|", a[j], "|", b[j])
    shutil.rmtree(d, ignore_errors=True)
