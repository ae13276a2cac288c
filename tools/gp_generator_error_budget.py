"""Error budget of the phase-table covariance generator, against an 80-bit evaluation of the same kernel."""
import sys, json
import numpy as np
LD = np.longdouble

def chol_logl_ld(C, r):
    C = C.astype(LD).copy(); n = len(r)
    L = np.zeros_like(C)
    for j in range(n):
        d = C[j, j] - np.dot(L[j, :j], L[j, :j])
        L[j, j] = np.sqrt(d)
        if j + 1 < n:
            L[j+1:, j] = (C[j+1:, j] - L[j+1:, :j] @ L[j, :j]) / L[j, j]
    z = np.zeros(n, LD); rr = r.astype(LD)
    for i in range(n):
        z[i] = (rr[i] - np.dot(L[i, :i], z[:i])) / L[i, i]
    logdet = 2 * np.sum(np.log(np.diag(L)))
    return -0.5 * n * np.log(2 * LD(np.pi)) - 0.5 * logdet - 0.5 * np.dot(z, z)

def K_ref(t, e1, e2, e3, e4):
    tau = t[:, None] - t[None, :]
    return e1 * e1 * np.exp(-0.5 * (tau / e2) ** 2 - 2.0 * (np.sin(np.pi * tau / e3) / e4) ** 2)

def K_true(t, e1, e2, e3, e4):
    t = t.astype(LD); e1, e2, e3, e4 = map(LD, (e1, e2, e3, e4))
    pi = LD("3.14159265358979323846264338327950288")
    # M_PI is what the reference multiplies by; the truth here is the exact kernel of the fp64 M_PI constant
    pi = LD(np.pi)
    tau = t[:, None] - t[None, :]
    return e1 * e1 * np.exp(-LD(0.5) * (tau / e2) ** 2 - LD(2.0) * (np.sin(pi * tau / e3) / e4) ** 2)

def K_fast(t, e1, e2, e3, e4):
    u = t - t[len(t) // 2]            # exact (Sterbenz) for same-magnitude epochs
    q = u.astype(LD) / LD(e3)          # stands for the double-double quotient
    w = q - np.rint(q)
    sg = np.where(np.rint(q).astype(np.int64) & 1, -1.0, 1.0)
    pi = LD(np.pi)
    s = (np.sin(pi * w)).astype(np.float64) * sg; c = (np.cos(pi * w)).astype(np.float64) * sg
    g1 = -0.5 / (e2 * e2); g2 = -2.0 / (e4 * e4)
    tau = u[:, None] - u[None, :]
    sd = s[:, None] * c[None, :] - c[:, None] * s[None, :]
    return (e1 * e1) * np.exp(tau * tau * g1 + sd * sd * g2)

if __name__ == "__main__":
    name = sys.argv[1]; nl = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    z = np.load("tests/golden/%s.npz" % name, allow_pickle=True)
    L = json.loads(str(z["layout"])); t = z["t"]; y = z["y"]; sig = z["sig"]
    P = z["params"]; re = L["row_eta"]
    rng = np.random.default_rng(1)
    for b in range(min(nl, P.shape[1])):
        e1, e2, e3, e4 = P[re:re + 4, b]; jit = P[L["row_jit"], b]
        r = y - np.mean(y) + 0 * rng.standard_normal(len(y))
        D = np.diag(sig ** 2 + jit ** 2)
        Kt = K_true(t, e1, e2, e3, e4); Kr = K_ref(t, e1, e2, e3, e4); Kf = K_fast(t, e1, e2, e3, e4)
        lt = chol_logl_ld(Kt + D.astype(LD), r); lr = chol_logl_ld(Kr + D, r); lf = chol_logl_ld(Kf + D, r)
        sc = e1 * e1
        print(name, b, "eta", e1, e2, e3, e4, "jit", jit)
        print("  entry err / eta1^2: ref %.2e fast %.2e" % (np.abs(Kr - Kt).max() / sc, np.abs(Kf - Kt).max() / sc))
        print("  logL %.6f  rel err: ref %.2e fast %.2e  fast-vs-ref %.2e" % (lt, abs((lr - lt) / lt), abs((lf - lt) / lt), abs((lf - lr) / lr)))
