"""Log-evidence, information and posterior weights of a DNest4 run directory (levels.txt +
sample_info.txt): a restatement of pykima.classic.postprocess
(/root/reference/pykima/classic.py:20-135), deterministic branch (numResampleLogX=1,
temperature=1, cut=0):

  * a sample belongs to the highest level whose (log L, tiebreaker) it exceeds, starting from the
    level it was saved at (:69-72);
  * the N samples of a level, in ascending (log L, tiebreaker), get
    log X = logX_level + log(Umin + (1 - Umin) * linspace(1/(N+1), 1 - 1/(N+1), N)) in descending
    order, Umin = exp(logX_next - logX_level) (:86-104);
  * prior weight of a sample = (X_right - X_left) / 2 between its neighbours in the level (the
    level's edges at the ends) (:106-120); Z = sum of weight * L after normalising the weights
    (:124-129).

Pinned against pykima's own function on a committed run (tests/golden/postprocess_run/, made by
tests/golden/make_golden_postprocess.py; tests/test_host_logic.py).  Test helper: both arms of a
comparison go through it."""
import numpy as np


def _logsumexp(v):
    m = np.max(v)
    return np.log(np.sum(np.exp(v - m))) + m


def _logdiffexp(x1, x2):
    return np.log(1.0 - np.exp(x2 - x1)) + x1


def postprocess(levels, sample_info):
    """(log Z, H, log X of every sample, posterior weights) from the two arrays."""
    levels = np.atleast_2d(np.asarray(levels, np.float64))
    si = np.atleast_2d(np.asarray(sample_info, np.float64))
    nl, ns = levels.shape[0], si.shape[0]
    lvl = [(levels[i, 1], levels[i, 2]) for i in range(nl)]
    smp = [(si[i, 1], si[i, 2], i) for i in range(ns)]
    sandwich = si[:, 0].astype(int).copy()
    for i in range(ns):
        while sandwich[i] < nl - 1 and smp[i] > lvl[sandwich[i] + 1]:
            sandwich[i] += 1
    logx = np.zeros(ns)
    logp = np.zeros(ns)
    for i in range(nl):
        which = np.nonzero(sandwich == i)[0]
        here = sorted(smp[k] for k in which)
        N = len(here)
        logx_max = levels[i, 0]
        logx_min = -1e300 if i == nl - 1 else levels[i + 1, 0]
        Umin = np.exp(logx_min - logx_max)
        U = Umin + (1.0 - Umin) * np.linspace(1.0 / (N + 1), 1.0 - 1.0 / (N + 1), N)
        lx = np.sort(logx_max + np.log(U))[::-1]
        for j in range(N):
            logx[here[j][2]] = lx[j]
            if j != N - 1:
                left = lx[j + 1]
            elif i == nl - 1:
                left = -1e300
            else:
                left = levels[i + 1, 0]
            right = lx[j - 1] if j != 0 else levels[i, 0]
            logp[here[j][2]] = np.log(0.5) + _logdiffexp(right, left)
    logl = si[:, 1]
    logp -= _logsumexp(logp)
    logP = logp + logl
    logz = _logsumexp(logP)
    P = np.exp(logP - logz)
    H = -logz + np.sum(P * logl)
    return float(logz), float(H), logx, P


def log_evidence(run_dir):
    """(log Z, posterior weights, H) of the run in `run_dir`."""
    lv = np.loadtxt(str(run_dir) + "/levels.txt", ndmin=2)
    si = np.loadtxt(str(run_dir) + "/sample_info.txt", ndmin=2)
    logz, H, _, P = postprocess(lv, si)
    return logz, P, H
