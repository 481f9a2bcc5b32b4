"""Numerical probe: K-side Schur (current) vs N-side (Woodbury with H^-1) formulation."""
import sys, numpy as np, warnings
sys.path.insert(0, '.')
from oracle import meem_oracle as mo
from scipy import linalg
warnings.simplefilter("ignore")
def probe(name, h, a, d, n, om):
    geo = mo.FlatGeometry(h, a, d, [n]*(len(a)+1))
    M = geo.M; K = geo.NMK[-1]; N = geo.NMK[-2]; NT = geo.size; p = NT-K
    m0 = mo.wavenumber(om, h); mk = mo.m_k_array(K, m0, h); Nk = mo.N_k_array(m0, h, mk)
    A = mo.assemble_A(geo, m0, mk, Nk); flags=[1]*M
    b = mo.assemble_b(geo, flags, m0, mk, Nk)
    X = linalg.solve(A, b)
    c = mo.c_vector(geo, flags)
    ref = np.dot(c, X[:-K])
    A11 = A[:p,:p].real; A12 = A[:p,p:].real; A21 = A[p:,:p].real; A22 = A[p:,p:]
    rows = np.where(np.abs(A12).sum(1)>0)[0]; cols = np.where(np.abs(A21).sum(0)>0)[0]
    Imk = -A12[rows]            # N x K
    E = np.zeros((p,N)); E[rows, np.arange(N)] = 1
    lu = linalg.lu_factor(A11)
    G = linalg.lu_solve(lu, E); y0 = linalg.lu_solve(lu, b[:p].real)
    # A21 = -Imk^T diag(R') on cols: A21[k, col] = -dR[col] Imk[n(col), k]
    # define Hfull: A21 @ G = -Imk^T Hm   with Hm (N x N)
    # A21 = Bm @ P where Bm (K x p). find Hm via least squares-free: A21 G = -(Imk^T) Hm
    # build dR per col: A21[:,col] = -dR * Imk[n,:]
    Hm = np.zeros((N,N)); u = np.zeros(N)
    for col in cols:
        # find n: which Imk row is proportional
        kk = np.argmax(np.abs(A21[:,col]))
        ratios = -A21[kk,col]/Imk[:,kk]
        # choose n with best match across all k
        errs = [np.max(np.abs(-r*Imk[n,:]-A21[:,col]))/np.max(np.abs(A21[:,col])) for n,r in enumerate(ratios)]
        n_ = int(np.argmin(errs)); dR = ratios[n_]
        Hm[n_] += dR*G[col]; u[n_] += dR*y0[col]
    D = np.diag(A22)
    # K-side: S = D + Imk^T Hm Imk ; rhs = b2 + Imk^T u      (A21 x1 = -Imk^T(u + Hm v))
    S = np.diag(D) + Imk.T @ Hm @ Imk
    # check S vs direct Schur
    S_direct = A22 - A21 @ linalg.lu_solve(lu, A12)
    # x1 = y0 - A11^-1 A12 x2 = y0 + G Imk x2
    b2 = b[p:]
    x2 = linalg.solve(S, b2 + Imk.T @ u)
    v = Imk @ x2
    coefK = c@y0 + (G.T@c)@v
    # N-side: v = Imk D^-1 (b2 + Imk^T (u + Hm v))  =>  (I - Q Hm) v = Imk D^-1 b2 + Q u,  Q = Imk D^-1 Imk^T
    Q = (Imk / D) @ Imk.T
    r = (Imk/D) @ b2 + Q @ u
    v1 = linalg.solve(np.eye(N) - Q@Hm, r)
    coefN1 = c@y0 + (G.T@c)@v1
    # with H^-1:  (Hm^-1 - Q) w = r, w = Hm v
    condH = np.linalg.cond(Hm)
    try:
        Hinv = linalg.inv(Hm)
        w = linalg.solve(Hinv - Q, r); v2 = Hinv @ w
        coefN2 = c@y0 + (G.T@c)@v2
    except Exception as e:
        coefN2 = np.nan
    rel = lambda x: abs(x-ref)/abs(ref)
    print(f"{name:22s} n={n:3d} om={om:4.1f} Sdiff={np.max(np.abs(S-S_direct))/np.max(np.abs(S_direct)):.1e} K-side {rel(coefK):.1e}  N-side(I-QH) {rel(coefN1):.1e}  N-side(Hinv) {rel(coefN2):.1e}  cond(H)={condH:.1e} cond(S)={np.linalg.cond(S):.1e} cond(I-QH)={np.linalg.cond(np.eye(N)-Q@Hm):.1e}")
for om in (0.5, 1.5, 3.0):
    probe("mini_bicylinder", 1.001, [0.125,0.25],[0.25,0.125], 10, om)
    probe("small_tricylinder", 20.0, [5,10,15],[15,10,5], 50, om)
    probe("big_tricylinder", 25.0, [10,15,20],[20,15,10], 200, om)
