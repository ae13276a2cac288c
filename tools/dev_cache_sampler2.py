"""Developer check: is the host-proposal CoRoT7 run reproducible run to run, and where do cache on / off differ?"""
import filecmp, json, os, shutil, subprocess, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sampler_bench as sb
from dev_cache_sampler import one, FILES

threads = int(sys.argv[1]) if len(sys.argv) > 1 else 256
ts, saves = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (20, 20)
runs = {}
for tag, cache in (("off_a", "0"), ("off_b", "0"), ("on_a", "1"), ("on_b", "1"), ("on_c", "1")):
    runs[tag] = one("CoRoT7", threads, ts, saves, dict(KIMA_CUDA_GP_CACHE=cache))
def same(a, b):
    return {f: filecmp.cmp(os.path.join(runs[a][0], f), os.path.join(runs[b][0], f), shallow=False) for f in FILES}
print("off_a vs off_b", same("off_a", "off_b"))
print("on_a vs on_b", same("on_a", "on_b"))
print("off_a vs on_a", same("off_a", "on_a"))
print("off_a vs on_b", same("off_a", "on_b"))
print("off_a vs on_c", same("off_a", "on_c"))
a = open(os.path.join(runs["off_a"][0], "sample_info.txt")).read().splitlines()
b = open(os.path.join(runs["on_a"][0], "sample_info.txt")).read().splitlines()
for i, (x, y) in enumerate(zip(a, b)):
    if x != y:
        print("first difference at line", i, "\n ", x, "\n ", y)
        break
