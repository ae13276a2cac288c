"""Developer check: repeated CoRoT7 host-proposal runs, cache off and on; which files differ from the first run, and where."""
import filecmp, os, shutil, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from dev_cache_sampler import one, FILES
threads, ts, saves, reps = 4096, 100, int(sys.argv[1]), int(sys.argv[2])
base = None
for tag, cache in [("off0", "0")] + [("dbg%d" % i, "1") for i in range(reps)]:
    d, res = one("CoRoT7", threads, ts, saves, dict(KIMA_CUDA_GP_CACHE=cache, KIMA_CUDA_GP_CACHE_DEBUG=(sys.argv[3] if len(sys.argv) > 3 else "0")))
    if base is None:
        base = d
        print(tag, "reference run", res)
        continue
    diff = [f for f in FILES if not filecmp.cmp(os.path.join(base, f), os.path.join(d, f), shallow=False)]
    print(tag, "differs in", diff)
    for f in diff:
        a = open(os.path.join(base, f)).read().splitlines(); b = open(os.path.join(d, f)).read().splitlines()
        nd = [i for i, (x, y) in enumerate(zip(a, b)) if x != y]
        print("  ", f, "lines", len(a), len(b), "differing", len(nd), "first", nd[:3])
        for i in nd[:2]:
            print("     <", a[i][:200]); print("     >", b[i][:200])
    shutil.rmtree(d, ignore_errors=True)
