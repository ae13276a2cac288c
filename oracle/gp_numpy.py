"""TEST INFRASTRUCTURE — numpy/LAPACK restatement of the GP likelihood for sizes where the
plain-C oracle is too slow (N=4096).  Follows RVmodel::calculate_C (src/RVmodel.cpp:186-216)
and the GP branch of log_likelihood (:841-873); the residual comes from the C oracle's mu."""
import numpy as np


def qp_cov(t, sig, jit2_per_epoch, eta1, eta2, eta3, eta4):
    tau = t[:, None] - t[None, :]
    C = eta1 * eta1 * np.exp(-0.5 * (tau / eta2) ** 2 - 2.0 * (np.sin(np.pi * tau / eta3) / eta4) ** 2)
    C[np.diag_indices_from(C)] += sig * sig + jit2_per_epoch
    return C


def gp_loglik(residual, C):
    L = np.linalg.cholesky(C)
    logdet = 2.0 * np.sum(np.log(np.diag(L)))
    z = np.linalg.solve(L, residual) if L.shape[0] < 2000 else __import__("scipy.linalg", fromlist=["x"]).solve_triangular(L, residual, lower=True)
    return -0.5 * residual.size * np.log(2 * np.pi) - 0.5 * logdet - 0.5 * float(z @ z)
