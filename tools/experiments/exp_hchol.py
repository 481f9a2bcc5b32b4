"""Probe: is H = dR1 G[c1 rows] + dR2 G[c2 rows] numerically symmetric PSD, and how well does a diagonally pivoted,
truncated Cholesky H ~ Pm L L^T Pm^T reproduce it?  (Basis of the W = L^T Imk form of the frequency GEMM.)"""
import sys, numpy as np, warnings
sys.path.insert(0, '.')
from oracle import meem_oracle as mo
from scipy import linalg
warnings.simplefilter("ignore")


def getH(h, a, d, NMK):
    geo = mo.FlatGeometry(h, a, d, NMK)
    K = geo.NMK[-1]; N = geo.NMK[-2]; NT = geo.size; p = NT - K; M = geo.M
    m0 = mo.wavenumber(1.0, h); mk = mo.m_k_array(K, m0, h); Nk = mo.N_k_array(m0, h, mk)
    A = mo.assemble_A(geo, m0, mk, Nk)
    A11 = A[:p, :p].real; A12 = A[:p, p:].real
    rows = np.where(np.abs(A12).sum(1) > 0)[0]
    E = np.zeros((p, N)); E[rows, np.arange(N)] = 1
    G = linalg.solve(A11, E)
    c1 = p - 2 * N if M > 1 else 0
    aM = geo.a[M - 1]
    dR1 = np.array([mo.diff_R_1n(n, aM, M - 1, geo) for n in range(N)])
    H = dR1[:, None] * G[c1:c1 + N]
    if M > 1:
        dR2 = np.array([mo.diff_R_2n(n, aM, M - 1, geo) for n in range(N)])
        H = H + dR2[:, None] * G[c1 + N:c1 + 2 * N]
    return H


def pchol(H, tolfac=1.0):
    Hs = 0.5 * (H + H.T); n = Hs.shape[0]; perm = np.arange(n); L = np.zeros((n, n)); dg = np.diag(Hs).copy()
    dmax = dg.max(); r = n
    for j in range(n):
        q = j + np.argmax(dg[j:])
        if dg[q] <= tolfac * n * 2.2e-16 * dmax:
            r = j; break
        perm[[j, q]] = perm[[q, j]]; L[[j, q]] = L[[q, j]]; dg[[j, q]] = dg[[q, j]]
        L[j, j] = np.sqrt(dg[j])
        L[j + 1:, j] = (Hs[perm[j + 1:], perm[j]] - L[j + 1:, :j] @ L[j, :j]) / L[j, j]
        dg[j + 1:] -= L[j + 1:, j] ** 2
    return L[:, :r], perm, r, dg[r:]


if __name__ == "__main__":
    cases = [("mini", (1.001, [0.125, 0.25], [0.25, 0.125], [10] * 3)), ("small", (20.0, [5, 10, 15], [15, 10, 5], [50] * 4)),
             ("single", (3.0, [1.0], [0.7], [9, 11])), ("overhang", (1.001, [0.5, 1.0], [0.25, 0.5], [6, 7, 8])),
             ("config3", (1.9, [0.3, 0.5, 1.0, 1.2, 1.6], [0.5, 0.7, 0.8, 0.2, 0.5], [8, 9, 10, 11, 12, 13])),
             ("uneq", (20.0, [5, 10, 15], [15, 10, 5], [30, 12, 25, 40])), ("uneq2", (20.0, [5, 10, 15], [15, 10, 5], [30, 12, 40, 9])),
             ("big", (25.0, [10, 15, 20], [20, 15, 10], [200] * 4))]
    if len(sys.argv) > 1:
        from bench import c4_geometries
        cases = [("c4_%d" % i, (g["h"], g["a"], g["d"], g["NMK"])) for i, g in enumerate(c4_geometries(int(sys.argv[1])))]
    worst = 0
    for name, args in cases:
        H = getH(*args)
        Hs = 0.5 * (H + H.T)
        ev = np.linalg.eigvalsh(Hs)
        L, perm, r, rest = pchol(H)
        err = np.max(np.abs(L @ L.T - Hs[np.ix_(perm, perm)])) / np.max(np.abs(Hs))
        worst = max(worst, err)
        if len(cases) < 20:
            print(name, "N", H.shape[0], "asym %.1e" % (np.max(np.abs(H - H.T)) / np.max(np.abs(H))),
                  "eig min %.3e max %.3e" % (ev[0], ev[-1]), "rank", r, "recon err %.1e" % err, "rest", np.array(rest))
    print("worst recon err", worst)
