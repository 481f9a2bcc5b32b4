import sys, numpy as np, warnings
sys.path.insert(0, '.')
from oracle import meem_oracle as mo
from scipy import linalg
from bench import c4_geometries
warnings.simplefilter("ignore")
def probe(name, h, a, d, NMK, om):
    geo = mo.FlatGeometry(h, a, d, NMK)
    K = geo.NMK[-1]; NT = geo.size; p = NT-K
    m0 = mo.wavenumber(om, h); mk = mo.m_k_array(K, m0, h); Nk = mo.N_k_array(m0, h, mk)
    A = mo.assemble_A(geo, m0, mk, Nk)
    A11 = A[:p,:p].real; A12 = A[:p,p:].real; A21 = A[p:,:p].real; A22 = A[p:,p:]
    lu = linalg.lu_factor(A11)
    S = A22 - A21 @ linalg.lu_solve(lu, A12)
    Sr = S.real
    asym = np.max(np.abs(Sr-Sr.T))/np.max(np.abs(Sr))
    ev = np.linalg.eigvalsh(0.5*(Sr+Sr.T)[1:,1:])
    evf = np.linalg.eigvalsh(0.5*(Sr+Sr.T))
    return asym, ev.min(), ev.max(), evf.min()
rows=[]
for om in (0.3, 1.5, 3.0):
    rows.append(("mini",)+probe("mini", 1.001, [0.125,0.25],[0.25,0.125], [10]*3, om))
    rows.append(("small",)+probe("small", 20.0, [5,10,15],[15,10,5], [50]*4, om))
    rows.append(("overhang",)+probe("ov", 1.001, [0.5,1.0],[0.25,0.5], [6,7,8], om))
    rows.append(("single",)+probe("s", 3.0, [1.0],[0.7], [9,11], om))
    rows.append(("config3",)+probe("c3", 1.9, [0.3,0.5,1.0,1.2,1.6],[0.5,0.7,0.8,0.2,0.5],[8,9,10,11,12,13], om))
    rows.append(("uneq",)+probe("u", 20.0, [5,10,15],[15,10,5], [30,12,25,40], om))
    rows.append(("uneq2",)+probe("u", 20.0, [5,10,15],[15,10,5], [30,12,40,9], om))
for r in rows: print("%-10s asym %.1e  eigmin[1:] %.3e eigmax %.3e eigmin(full) %.3e" % r)
worst=0; mn=1e9
for g in c4_geometries(300):
    for om in (0.3,1.2,2.5):
        a_,e0,e1,ef = probe("c4", g["h"], g["a"], g["d"], g["NMK"], om)
        worst=max(worst,a_); mn=min(mn,e0/e1)
print("C4 300 geoms: worst asym", worst, "min eigmin/eigmax", mn)
