"""Accuracy of the angle-addition form of I_mk against the reference-ordered form and against mpmath."""
import sys, numpy as np, warnings
sys.path.insert(0, '.')
from oracle import meem_oracle as mo
import mpmath as mp
warnings.simplefilter("ignore")
mp.mp.dps = 40
def fast(geo, m0, mk, Nk):
    N = geo.NMK[-2]; K = geo.NMK[-1]; i = geo.M-1
    L = geo.h - geo.d[i]
    lam = np.array([mo.lambda_ni(n, i, geo.h, geo.dext) for n in range(N)])
    sk, ck = np.sin(mk*L), np.cos(mk*L); sl, cl = np.sin(lam*L), np.cos(lam*L)
    a = sk[None,:]*cl[:,None]; b = ck[None,:]*sl[:,None]
    sp = mk[None,:]+lam[:,None]; sd = mk[None,:]-lam[:,None]
    small = np.abs(sd*L) < 1.0
    with np.errstate(all='ignore'):
        t = np.where(small, (a+b)/sp + np.sin(sd*L)/sd, ((a+b)*sd+(a-b)*sp)/(sp*sd))
    return (np.sqrt(2/Nk)*0.5)[None,:]*t, lam, L
for name,(h,a,d,n) in dict(big=(25.0,[10,15,20],[20,15,10],200), small=(20.0,[5,10,15],[15,10,5],50)).items():
    geo = mo.FlatGeometry(h,a,d,[n]*4)
    for om in (0.3, 1.5, 3.0):
        m0 = mo.wavenumber(om,h); mk = mo.m_k_array(n,m0,h); Nk = mo.N_k_array(m0,h,mk)
        ref = mo.I_mk_table(geo,m0,mk,Nk)
        f, lam, L = fast(geo,m0,mk,Nk)
        sub = (slice(1,None), slice(1,None))
        dabs = np.abs(f[sub]-ref[sub]); scale = np.abs(ref).max()
        # exact (mpmath, from the same double inputs mk, lam, L, Nk)
        idx = np.unravel_index(np.argsort(dabs.ravel())[-5:], dabs.shape)
        worst=[]
        for r,c in zip(*idx):
            m_,k_ = r+1,c+1
            ex = mp.sqrt(2/mp.mpf(Nk[k_]))*(mp.sin((mp.mpf(mk[k_])+mp.mpf(lam[m_]))*L)/(mp.mpf(mk[k_])+mp.mpf(lam[m_]))+mp.sin((mp.mpf(mk[k_])-mp.mpf(lam[m_]))*L)/(mp.mpf(mk[k_])-mp.mpf(lam[m_])))/2
            worst.append((float(abs(ref[m_,k_]-ex)), float(abs(f[m_,k_]-ex))))
        print(name, om, "max|fast-ref|/max|ref| = %.2e" % (dabs.max()/scale), " (|ref-exact|, |fast-exact|) at worst entries:", ["%.1e/%.1e"%w for w in worst])
