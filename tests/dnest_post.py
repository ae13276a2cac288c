"""Log-evidence and posterior weights from a DNest4 run directory (levels.txt + sample_info.txt):
the prior-mass assignment of DNest4's postprocess — samples of a level are spread evenly, in
order of (log-likelihood, tiebreaker), over the level's interval of X; Z is the trapezoid sum of
L dX.  Test helper: both arms of a comparison go through the same estimator."""
import numpy as np


def log_evidence(run_dir):
    lv = np.loadtxt(str(run_dir) + "/levels.txt", ndmin=2)
    si = np.loadtxt(str(run_dir) + "/sample_info.txt", ndmin=2)
    logX = lv[:, 0]
    lx = np.full(len(si), np.nan)
    for j in range(len(lv)):
        idx = np.where(si[:, 0] == j)[0]
        if idx.size == 0:
            continue
        X_hi = np.exp(logX[j])
        X_lo = np.exp(logX[j + 1]) if j + 1 < len(lv) else 0.0
        order = idx[np.lexsort((si[idx, 2], si[idx, 1]))]  # ascending likelihood = descending X
        n = idx.size
        lx[order] = np.log(X_hi - (np.arange(n) + 0.5) / n * (X_hi - X_lo))
    o = np.argsort(-lx)
    X = np.exp(lx[o])
    logL = si[o, 1]
    left = np.concatenate([[1.0], X[:-1]])
    right = np.concatenate([X[1:], [0.0]])
    logw = np.log(0.5 * (left - right)) + logL
    m = logw.max()
    logZ = m + np.log(np.exp(logw - m).sum())
    post = np.exp(logw - logZ)
    w = np.empty_like(post)
    w[o] = post
    H = float(np.sum(post * (logL - logZ)))
    return logZ, w, H
