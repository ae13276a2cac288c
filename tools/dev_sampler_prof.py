"""Developer timing / ncu target of the device-side sampler: synthetic white-noise data of the
51Peg shape (N=256, Np in {0,1}), `lanes` lanes x 1 particle, a fixed 8-level table; prints
MCMC steps/s from the device time of kc_sampler_run and the share of the likelihood kernel."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kima_b200 import LikelihoodEngine, ModelFlags  # noqa: E402
from kima_b200.sampler import DeviceSampler, SamplerOptions, default_prior_table, draw_from_prior  # noqa: E402
from kima_b200.synthetic import make_data  # noqa: E402


def main():
    lanes = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 50
    reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    N = int(sys.argv[4]) if len(sys.argv) > 4 else 256
    npmax = int(sys.argv[5]) if len(sys.argv) > 5 else 1
    gp = len(sys.argv) > 6 and sys.argv[6] == "gp"
    flags = ModelFlags(npmax=npmax, GP=gp)
    data = make_data(N)
    pri = default_prior_table(flags, data)
    with LikelihoodEngine(data, flags, lanes) as eng:
        ds = DeviceSampler(eng, pri, SamplerOptions(lanes=lanes, num_particles=1, thread_steps=steps, max_levels=64))
        S, nps = draw_from_prior(ds, flags, pri, lanes)
        logl = eng.evaluate(S[:eng.layout["n_rows"]], nps)
        logl = np.where(np.isfinite(logl), logl, -1e300)
        tb = np.random.default_rng(1).uniform(0.01, 0.99, lanes)
        q = np.quantile(logl, np.linspace(0.1, 0.8, 7))
        lv_logl = np.concatenate([[-1e300], q])
        level = np.zeros(lanes, np.int32)
        ds.upload(S, nps, logl, tb, level)
        lv_tb = np.full(8, 0.5)
        lv_logX = -np.arange(8.0)
        tries = np.full(8, 1000, np.uint64)
        out = []
        for r in range(reps):
            ds.set_levels(lv_logl, lv_tb, lv_logX, tries, True, True)
            ds.run(steps)
            cnt, _ = ds.collect()
            ms = ds.last_run_ms()
            tries = tries + cnt[:, 1]
            out.append(ms)
        # the likelihood kernel alone over the same proposals
        eng.log_likelihood_device(lanes)
        eng.sync()
        lk = eng.last_kernel_ms()
        best = min(out)
        print(json.dumps({"lanes": lanes, "N": N, "npmax": npmax, "GP": gp, "steps": steps, "run_ms": out,
                          "ms_per_step": best / steps, "steps_per_s": lanes * steps / best * 1e3,
                          "likelihood_ms_per_step": lk, "accept_rate": float(cnt[:, 0].sum() / cnt[:, 1].sum())}))
        ds.close()


if __name__ == "__main__":
    main()
