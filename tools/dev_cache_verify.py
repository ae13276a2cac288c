"""Developer check: the CoRoT7 host-proposal run with KIMA_CUDA_GP_CACHE_VERIFY=1 (every cached batch re-evaluated uncached)."""
import os, shutil, subprocess, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sampler_bench as sb
threads, ts, saves = 4096, 100, int(sys.argv[1]) if len(sys.argv) > 1 else 60
for rep in range(int(sys.argv[2]) if len(sys.argv) > 2 else 3):
    exe = os.path.join(sb.BUILD, "kima_gpu_CoRoT7")
    d = tempfile.mkdtemp(prefix="kima_verify_")
    for f in sb.DATA["CoRoT7"]:
        shutil.copy(os.path.join(sb.E2E, f), d)
    open(os.path.join(d, "OPTIONS"), "w").write(sb.OPTIONS.format(nli=max(2000, threads * ts // 4), ts=ts, saves=saves))
    r = subprocess.run([exe, "-t", str(threads), "-s", "43"], cwd=d, capture_output=True, text=True, timeout=1200,
                       env=dict(os.environ, KIMA_CUDA_GP_CACHE_VERIFY="1"))
    mm = [l for l in r.stderr.splitlines() if "MISMATCH" in l]
    print("rep", rep, "rc", r.returncode, "mismatches", len(mm))
    for l in mm[:10]: print("  ", l)
    shutil.rmtree(d, ignore_errors=True)
