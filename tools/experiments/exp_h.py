import sys, numpy as np, warnings
sys.path.insert(0, '.')
from oracle import meem_oracle as mo
from scipy import linalg
from bench import c4_geometries
warnings.simplefilter("ignore")
def getH(h, a, d, NMK):
    geo = mo.FlatGeometry(h, a, d, NMK)
    K = geo.NMK[-1]; N = geo.NMK[-2]; NT = geo.size; p = NT-K; om=1.0
    m0 = mo.wavenumber(om, h); mk = mo.m_k_array(K, m0, h); Nk = mo.N_k_array(m0, h, mk)
    A = mo.assemble_A(geo, m0, mk, Nk)
    A11 = A[:p,:p].real; A12 = A[:p,p:].real; A21 = A[p:,:p].real
    rows = np.where(np.abs(A12).sum(1)>0)[0]
    Imk = -A12[rows]
    lu = linalg.lu_factor(A11)
    E = np.zeros((p,N)); E[rows, np.arange(N)] = 1
    G = linalg.lu_solve(lu, E)
    # A21 G = -Imk^T H  -> H = -(Imk^T)^+ A21 G  (K>=N => full column rank of Imk^T? Imk^T is KxN)
    Hm = -np.linalg.lstsq(Imk.T, A21 @ G, rcond=None)[0]
    return Hm
for name,args in [("mini",(1.001,[0.125,0.25],[0.25,0.125],[10]*3)),("small",(20.0,[5,10,15],[15,10,5],[50]*4)),
                  ("single",(3.0,[1.0],[0.7],[9,11])),("overhang",(1.001,[0.5,1.0],[0.25,0.5],[6,7,8])),
                  ("config3",(1.9,[0.3,0.5,1.0,1.2,1.6],[0.5,0.7,0.8,0.2,0.5],[8,9,10,11,12,13]))]:
    H = getH(*args)
    ev = np.linalg.eigvalsh(0.5*(H+H.T))
    print(name, "asym", np.max(np.abs(H-H.T))/np.max(np.abs(H)), "eig", ev[:3], ev[-3:])
