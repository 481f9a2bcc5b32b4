import sys, numpy as np, warnings
sys.path.insert(0, '.')
from oracle import meem_oracle as mo
from scipy import linalg
warnings.simplefilter("ignore")
def probe(name, h, a, d, n, om):
    geo = mo.FlatGeometry(h, a, d, [n]*(len(a)+1))
    K = geo.NMK[-1]; N = geo.NMK[-2]; NT = geo.size; p = NT-K
    m0 = mo.wavenumber(om, h); mk = mo.m_k_array(K, m0, h); Nk = mo.N_k_array(m0, h, mk)
    A = mo.assemble_A(geo, m0, mk, Nk)
    A11 = A[:p,:p].real; A12 = A[:p,p:].real; A21 = A[p:,:p].real; A22 = A[p:,p:]
    lu = linalg.lu_factor(A11)
    S = A22 - A21 @ linalg.lu_solve(lu, A12)
    Sr = S.real
    print(name, n, om, "asym(S)=", np.max(np.abs(Sr-Sr.T))/np.max(np.abs(Sr)), "diag dom:", np.min(np.abs(np.diag(Sr)) / (np.abs(Sr).sum(1)-np.abs(np.diag(Sr)))),
          "|D| range", np.abs(np.diag(A22)).min(), np.abs(np.diag(A22)).max(), "offdiag max", np.max(np.abs(Sr-np.diag(np.diag(Sr)))))
probe("mini", 1.001, [0.125,0.25],[0.25,0.125], 10, 1.5)
probe("small", 20.0, [5,10,15],[15,10,5], 50, 1.5)
probe("big", 25.0, [10,15,20],[20,15,10], 200, 1.5)
