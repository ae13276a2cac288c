#!/usr/bin/env python
"""bench.py — log-likelihood evaluations per second on the BASELINE.json workloads.

    python bench.py --gpus N --steps K --warmup W            # the CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path

Workloads (BASELINE.json `metric`):
  K  Keplerian, N=1000 epochs, Np=10 planets in every lane, Gaussian white noise
  G  quasi-periodic GP (dense Cholesky), N=1000 epochs, Np=10
A "step" scores one batch of lanes (particles) once.  The top-level JSON keys are
workload K; the same keys for workload G sit under "gp".  One rank per GPU; lanes
shard with no data-path collective (weak scaling: lanes per GPU fixed).

Timing: per step, the L2 is flushed (untimed), then the step is timed on the device
with CUDA events on the stream the kernels run on (kc_last_kernel_ms); the K steps
are bracketed by a barrier + device synchronise and the max over ranks is taken.
`e2e` times the public call with host buffers (H2D + kernels + D2H) by wall clock.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from kima_b200.flags import ModelFlags, layout_from_flags  # noqa: E402
from kima_b200.synthetic import make_data, make_params  # noqa: E402

METRIC = "log-likelihood evals/sec: Keplerian (N=1k, Np=10) and QP-GP Cholesky (N=1k)"
N_EPOCHS = 1000
NP = 10
KEP_LANES = 8192          # lanes per GPU, workload K
GP_LANES = 1184           # lanes per GPU, workload G (8 per SM)
REF_KEP_LANES = 4096      # bounded sample per step of the CPU arms
REF_GP_LANES = 48


def flops_keplerian(N, Np):
    """F_K = N (330 Np + 45)."""
    return N * (330.0 * Np + 45.0)


def flops_gp(N, Np):
    """F_G = N^3/3 + 37 N^2 + F_K."""
    return N ** 3 / 3.0 + 37.0 * N * N + flops_keplerian(N, Np)


class ClockSampler:
    """nvidia-smi clocks + throttle reasons while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.rows = []
        self.proc = None
        self.th = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            return

        def pump():
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        self.th = threading.Thread(target=pump, daemon=True)
        self.th.start()

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1]))
                smax = float(r[2])
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if r[5 + k].lower().startswith("active"):
                    reasons.add(nm)
        sm_sorted = sorted(sm)
        # "under load": the upper half of the samples (idle samples bracket the region)
        load = sm_sorted[len(sm_sorted) // 2:] if sm_sorted else []
        return {"sm_mhz": statistics.median(load) if load else None, "sm_max_mhz": smax,
                "reasons": sorted(reasons), "samples": len(sm)}


def init_dist(n_gpus: int):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"   # keep NCCL's version banner off stdout (one JSON line)
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


def cpu_arm(kind_pref: str, flags: ModelFlags, data, lanes: int, threads: int, reps: int = 1):
    """Time the reference's own CPU implementation (oracle/_ref, kind "reference") or,
    where it was not built, the plain-C restatement (kind "port") on `lanes` lanes."""
    from oracle.oracle import Oracle, have_reference
    kind = "reference" if (kind_pref == "reference" and have_reference()) else "port"
    o = Oracle(kind)
    o.set_rvdata(data)
    o.set_flags(flags)
    P, nps = make_params(flags, data, lanes, "max")
    best = None
    for _ in range(reps):
        _, secs = o.eval_batch(P, nps, threads)
        best = secs if best is None else min(best, secs)
    return kind, lanes / best, best


# ncu --set full captures of the dominant kernel of each workload (profiles/, per launch):
# dram__bytes_read.sum + dram__bytes_write.sum, and what the algorithm needs per launch.
NCU_TRAFFIC = {
    "kep": {"kernel": "keplerian_warp_kernel<0,0>", "traffic": 3.50e6,
            "algorithmic_bytes": 8192 * (52 * 8 + 4 + 8) + 4 * 8000,
            "source": "profiles/ncu_kep_r01d.json (8192 lanes: parameters in, logL out, data arrays once)"},
    "gp": {"kernel": "gpb_update_kernel (block column 8 of 16, j0=512, one lane group of 592)",
           "traffic": 1.424e9 + 0.118e9,
           "algorithmic_bytes": 592 * ((425 + 64) * 512 + 425 * 64) * 8,
           "source": "profiles/ncu_gp_r01d.json (592 lanes: 425 panel rows x 512 columns and the 64-row block "
                     "row read once, 425 x 64 tile of L written once)"},
}


def run_workload(eng, lanes: int, steps: int, warmup: int, flops_per_eval: float, peak_tflops: float,
                 exch, sampler_idx: int, key: str = "kep"):
    """W warm-up + K timed steps of one workload on this rank's engine."""
    for _ in range(max(warmup, 3)):
        eng.flush_l2()
        eng.log_likelihood_device(lanes)
    eng.sync()
    sampler = ClockSampler(sampler_idx)
    sampler.start()
    if exch:
        exch.barrier()
    n0 = eng.launch_count()
    dev_ms = []
    t0 = time.perf_counter()
    for _ in range(steps):
        eng.flush_l2()
        eng.log_likelihood_device(lanes)
        eng.sync()
        dev_ms.append(eng.last_kernel_ms())
    eng.sync()
    if exch:
        exch.barrier()
    wall = time.perf_counter() - t0
    launches = eng.launch_count() - n0
    clocks = sampler.stop()
    total_ms = float(sum(dev_ms))
    # end to end through the public call, host buffers: H2D + kernels + D2H inside the clock
    out = np.empty(lanes)
    eng.log_likelihood(lanes, out)
    e2e_t = []
    for _ in range(steps):
        eng.flush_l2()
        eng.sync()
        t1 = time.perf_counter()
        eng.log_likelihood(lanes, out)
        e2e_t.append(time.perf_counter() - t1)
    e2e_s = float(sum(e2e_t))
    if exch:
        total_ms = exch.max_float(total_ms)
        e2e_s = exch.max_float(e2e_s)
        launches = int(exch.sum_float(float(launches)))
    world = exch.world if exch else 1
    ms_per_step = total_ms / steps
    value = world * lanes / (ms_per_step * 1e-3)
    per_gpu = lanes / (ms_per_step * 1e-3)
    achieved = per_gpu * flops_per_eval * 1e-12
    n_rows = eng.layout["n_rows"]
    res = {
        "value": value, "ms_per_step": ms_per_step, "gpu_launches": launches,
        "e2e": {"value": world * lanes / (e2e_s / steps), "unit": "evals/s",
                "h2d_bytes_per_step": int(n_rows * lanes * 8 + lanes * 4),
                "d2h_bytes_per_step": int(lanes * 8)},
        # "tensor" = compute roofline in TFLOP/s (the contract's two classes are hbm | tensor).
        # The denominator is the FP64 peak: DMMA (FP64 tensor pipe) for the GP kernels, which
        # issue mma.sync.m8n8k4.f64; the Keplerian kernel runs on the FP64 vector pipe (DFMA),
        # whose measured peak is 0.91 of the DMMA figure used here.
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak_tflops, "unit": "TFLOP/s",
                     "frac": achieved / peak_tflops, "traffic": NCU_TRAFFIC[key]["traffic"],
                     "traffic_detail": NCU_TRAFFIC[key],
                     "pipe": "FP64 tensor (DMMA.8x8x4)" if key == "gp" else "FP64 vector (DFMA); no tensor-core instructions",
                     "flops_per_eval": flops_per_eval,
                     "per": "one step = one evaluation of all lanes (Keplerian: 1 launch; GP: 68 launches, two "
                            "lane groups x (prep + 16 diag + 16 update + finish); the dominant kernel "
                            "gpb_update_kernel is 82% of the step's kernel time, profiles/launches_r01d.csv)",
                     "peak_source": "kc_measure_fp64_peak in this run (max of DFMA / DMMA m8n8k4); "
                                    "MEASURED_PEAKS.json has no FP64 entry"},
        "clocks": clocks, "wall_s_timed_region": wall, "finite": bool(np.all(np.isfinite(out))),
    }
    return res


def predictive_leg(data, flags, device: int, threads: int, B: int = 1184, nt: int = 1000, cpu_B: int = 4):
    """posterior-predictive curves (deterministic model + GP mean on a time
    grid) per second through kc_predict with host buffers, beside the CPU restatement of
    pykima's per-sample loop (oracle/predictive.py) on a few samples."""
    from kima_b200 import LikelihoodEngine
    from oracle import predictive as opred
    from oracle.oracle import Oracle
    P, nps = make_params(flags, data, B, "max", seed=47)
    tg = np.linspace(data.t.min(), data.t.max(), nt)
    with LikelihoodEngine(data, flags, B, device=device) as eng:
        eng.predict(P, nps, tg)
        secs, kms = [], []
        for _ in range(3):
            t0 = time.perf_counter()
            m, g = eng.predict(P, nps, tg)
            secs.append(time.perf_counter() - t0)
            kms.append(eng.last_kernel_ms())
        launches = eng.launch_count()
    orc = Oracle("port")
    orc.set_rvdata(data)
    orc.set_flags(flags)
    t0 = time.perf_counter()
    err = 0.0
    for b in range(cpu_B):
        want = opred.model(orc, flags, data, P, b, nps[b], tg) + opred.gp_mean(orc, flags, data, P, b, nps[b], tg)
        err = max(err, float(np.max(np.abs(m[b] + g[b] - want)) / (np.abs(want).max() + 1.0)))
    cpu_s = time.perf_counter() - t0
    return {"unit": "predictive curves/s", "workload": "QP-GP N=%d Np=%d, %d posterior samples, %d grid times" % (data.N(), NP, B, nt),
            "value": B / min(secs), "ms_per_call": 1e3 * min(secs), "kernels_ms": min(kms), "gpu_launches": launches,
            "d2h_bytes_per_call": 2 * 8 * B * nt,
            "cpu_baseline": {"value": cpu_B / cpu_s, "unit": "predictive curves/s", "cores": threads, "kind": "port",
                             "sample": "%d samples, numpy/LAPACK restatement of pykima's per-sample loop, %.1f s" % (cpu_B, cpu_s)},
            "max_rel_err_vs_cpu": err}


def sampler_device_leg(local: int, lanes: int = 65536, steps: int = 50, reps: int = 3):
    """SURVEY.md 8 row f1, device-timed: whole MCMC steps (proposal, prior test, likelihood, accept, level
    move, counters) of `lanes` lanes on the GPU through kc_sampler_run, synthetic white-noise data of the
    51Peg shape (N=256, Np in {0,1}); CUDA events on the library's stream around each block of steps."""
    from kima_b200 import LikelihoodEngine
    from kima_b200.sampler import DeviceSampler, SamplerOptions, default_prior_table, draw_from_prior
    flags = ModelFlags(npmax=1)
    data = make_data(256)
    pri = default_prior_table(flags, data)
    with LikelihoodEngine(data, flags, lanes, device=local) as eng:
        ds = DeviceSampler(eng, pri, SamplerOptions(lanes=lanes, num_particles=1, thread_steps=steps, max_levels=64))
        S, nps = draw_from_prior(ds, flags, pri, lanes)
        logl = eng.evaluate(S[:eng.layout["n_rows"]], nps)
        logl = np.where(np.isfinite(logl), logl, -1e300)
        tb = np.random.default_rng(1).uniform(0.01, 0.99, lanes)
        lv_logl = np.concatenate([[-1e300], np.quantile(logl, np.linspace(0.1, 0.8, 7))])
        ds.upload(S, nps, logl, tb, np.zeros(lanes, np.int32))
        tries = np.full(8, 1000, np.uint64)
        ms = []
        launches0 = eng.launch_count()
        for _ in range(reps + 1):  # the first block is the warm-up
            ds.set_levels(lv_logl, np.full(8, 0.5), -np.arange(8.0), tries, True, True)
            ds.run(steps)
            cnt, _ = ds.collect()
            ms.append(ds.last_run_ms())
            tries = tries + cnt[:, 1]
        launches = eng.launch_count() - launches0
        eng.log_likelihood_device(lanes)
        eng.sync()
        lk = eng.last_kernel_ms()
        ds.close()
    best = min(ms[1:])
    return {"unit": "MCMC steps/s", "workload": "white noise N=256, Np<=1 (51Peg shape), %d lanes x 1 particle, "
            "8 levels, blocks of %d lock-step steps" % (lanes, steps), "value": lanes * steps / best * 1e3,
            "ms_per_step": best / steps, "likelihood_kernel_ms_per_step": lk, "gpu_launches": launches,
            "accept_rate": float(cnt[:, 0].sum() / max(1, cnt[:, 1].sum())),
            "kernels": "smp_sorted_propose_kernel<256> (accept of step i + move-sorted proposal of step i+1) + keplerian kernel per step (profiles/launches_sampler_r01e_summary.txt)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="kima_b200", choices=["kima_b200", "reference"])
    ap.add_argument("--kepler-mode", type=int, default=0)
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--workloads", default="kep,gp")
    ap.add_argument("--skip-sampler", action="store_true")
    ap.add_argument("--skip-predictive", action="store_true")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    threads = os.cpu_count() or 1
    data = make_data(N_EPOCHS)
    fk = ModelFlags(npmax=NP)
    fg = ModelFlags(GP=True, npmax=NP)
    base = {"metric": METRIC, "unit": "evals/s", "n_gpus": a.gpus, "steps": a.steps,
            "warmup": max(a.warmup, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic"}

    if a.impl == "reference":
        if rank != 0:
            return 0
        out = dict(base, impl="reference")
        res = {}
        for key, flags, lanes in (("kep", fk, REF_KEP_LANES), ("gp", fg, REF_GP_LANES)):
            for _ in range(1):  # one untimed pass warms the caches and the thread pool
                cpu_arm("reference", flags, data, max(lanes // 8, threads), threads)
            secs = []
            kind = None
            for _ in range(a.steps):
                kind, _, s = cpu_arm("reference", flags, data, lanes, threads)
                secs.append(s)
            v = lanes * len(secs) / sum(secs)
            res[key] = {"value": v, "unit": "evals/s", "ms_per_step": 1e3 * sum(secs) / len(secs),
                        "cpu_baseline": {"value": v, "unit": "evals/s", "cores": threads, "kind": kind,
                                         "sample": "%d lanes per step, %d steps, %d host threads" % (lanes, len(secs), threads)},
                        "e2e": {"value": v, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        out.update(res["kep"])
        out["gp"] = res["gp"]
        out["config"] = {"workload": "K: Keplerian N=1000 Np=10 Gaussian; gp: QP-GP N=1000 Np=10; "
                                     "reference RVmodel::calculate_mu/calculate_C/log_likelihood on host cores",
                         "lanes_per_step": REF_KEP_LANES, "gp_lanes_per_step": REF_GP_LANES}
        print(json.dumps(out))
        return 0

    # stdout carries exactly one JSON line: whatever libraries write there meanwhile (NCCL's version
    # banner at the first collective, ...) goes to stderr
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    rank, world, local = init_dist(a.gpus)
    from kima_b200 import LikelihoodEngine
    from kima_b200.dist import TorchBarrierExchange
    exch = None
    if world > 1:
        import torch
        exch = TorchBarrierExchange(torch.device("cuda", local))
    out = dict(base)
    results = {}
    peak = None
    for key, flags, lanes, fl in (("kep", fk, KEP_LANES, flops_keplerian(N_EPOCHS, NP)),
                                  ("gp", fg, GP_LANES, flops_gp(N_EPOCHS, NP))):
        if key not in a.workloads.split(","):
            continue
        eng = LikelihoodEngine(data, flags, lanes, device=local, kepler_mode=a.kepler_mode)
        if peak is None:
            dfma, dmma = eng.measure_fp64_peak(5)
            peak = max(dfma, dmma)
        # each rank scores its own shard of the global lane set (different seeds per rank)
        P, nps = make_params(flags, data, lanes, "max", seed=43 + rank)
        eng.set_params(P, nps)
        eng.log_likelihood(lanes)  # puts the block on the device
        r = run_workload(eng, lanes, a.steps, a.warmup, fl, peak, exch, local, key)
        if rank == 0 and world == 1 and not a.skip_cpu_baseline:
            ref_lanes = REF_KEP_LANES * 4 if key == "kep" else REF_GP_LANES * 8
            kind, v, secs = cpu_arm("reference", flags, data, ref_lanes, threads)
            r["cpu_baseline"] = {"value": v, "unit": "evals/s", "cores": threads, "kind": kind,
                                 "sample": "%d lanes of the same workload, %d host threads, %.1f s" % (ref_lanes, threads, secs)}
        results[key] = r
        eng.close()
    first = "kep" if "kep" in results else "gp"
    out.update(results[first])
    if first == "kep" and "gp" in results:
        out["gp"] = dict(results["gp"], unit="evals/s")
    out["config"] = {"workload": "K: Keplerian N=1000 Np=10 (Np fixed at npmax), Gaussian white noise, "
                                 "%d lanes/GPU; gp: QP-GP N=1000 Np=10, %d lanes/GPU" % (KEP_LANES, GP_LANES),
                     "lanes_per_gpu": KEP_LANES, "gp_lanes_per_gpu": GP_LANES,
                     "kepler_mode": a.kepler_mode, "parallelism": "lanes sharded over %d GPU(s), no data-path collective" % world,
                     "l2": "flushed between timed iterations (256 MiB memset on the kernel stream)",
                     "fp64_peak_tflops_measured": peak}
    if rank == 0 and world == 1 and not a.skip_sampler:
        # the reference's unmodified examples end to end, MCMC steps/s of
        # the batched GPU sampler (host/BatchedSampler.h) next to the reference's CPU sampler
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import sampler_bench
        os.environ.setdefault("KIMA_CUDA_DEVICE", str(local))
        out["sampler"] = {"unit": "MCMC steps/s", "lanes": 4096, "cpu_threads": threads,
                          "runs": [sampler_bench.compare(n, 4096, threads, gpu_saves=80, cpu_saves=100)
                                   for n in ("CoRoT7", "51Peg")]}
        # the same examples with proposals, prior test, accept logic and level
        # counters on the GPU as well (KIMA_DEVICE_PROPOSALS=1, include/kima_sampler.h): particles stay
        # in HBM, the host only does DNest4's bookkeeping between blocks of thread_steps steps
        dev_env = {"KIMA_DEVICE_PROPOSALS": "1"}
        out["sampler"]["device_proposals"] = [
            sampler_bench.run("gpu", "51Peg", 65536, 100, 20, env=dev_env),
            sampler_bench.run("gpu", "CoRoT7", 16384, 100, 10, env=dev_env)]
        out["sampler"]["device_steps"] = sampler_device_leg(local)
    if rank == 0 and world == 1 and not a.skip_predictive and not a.skip_cpu_baseline:
        out["predictive"] = predictive_leg(data, fg, local, threads)
    if world > 1 and not a.skip_sampler and os.path.exists(os.path.join(ROOT, "host", "_build", "kima_gpu_CoRoT7")):
        # the same end to end at N GPUs: one C++ process per GPU (host/BatchedSampler.h), lanes
        # sharded, the level-creation barrier over NCCL; 4096 lanes per GPU (weak scaling)
        import shutil
        import torch.distributed as dist
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import sampler_bench
        runs = []
        # (example, lanes per GPU, saves, device proposals?): host proposals as in the 1-GPU leg, and the
        # 51Peg example with the whole step on the GPUs (each rank keeps its lanes' particles in HBM)
        for i, (name, lpg, saves, dev) in enumerate((("CoRoT7", 4096, 80, False), ("51Peg", 16384, 20, True))):
            box = [sampler_bench.prepare_dir(name, lpg * world, 100, saves) if rank == 0 else None]
            dist.broadcast_object_list(box, src=0)
            dist.barrier()
            r = sampler_bench.run_rank(name, box[0], lpg * world, rank, world, local,
                                       int(os.environ.get("MASTER_PORT", "29500")) + 11 + i,
                                       max(1, threads // world),
                                       env={"KIMA_DEVICE_PROPOSALS": "1"} if dev else None)
            wall = exch.max_float(float(r.get("wall_s", 0.0)))
            bad = exch.max_float(1.0 if "error" in r else 0.0)
            if rank == 0:
                if bad or not r.get("mcmc_steps"):
                    runs.append({"example": name, "error": r.get("error", "a rank failed")})
                else:
                    runs.append({"example": name, "lanes": lpg * world, "device_proposals": dev, "mcmc_steps": r["mcmc_steps"],
                                 "wall_s": wall, "steps_per_s": r["mcmc_steps"] / wall,
                                 "run_s": r.get("run_s"),
                                 "steps_per_s_run": r["mcmc_steps"] / r["run_s"] if r.get("run_s") else None})
                shutil.rmtree(box[0], ignore_errors=True)
        if rank == 0:
            out["sampler"] = {"unit": "MCMC steps/s", "lanes_per_gpu": 4096, "ranks": world,
                              "host_threads_per_rank": max(1, threads // world), "runs": runs}
    if exch:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    os.close(saved_stdout)
    if rank == 0:
        print(json.dumps(out))
    return 0


if __name__ == "__main__":
    sys.exit(main())
