import sys, warnings, numpy as np
from scipy import linalg
sys.path.insert(0,'/root/repo')
from oracle import meem_oracle as mo
warnings.simplefilter("ignore")
sys.path.insert(0,'/root/repo/tools/experiments')
import exp_band as eb

def lu_window_static(W, npiv, nsearch, static_rows, static_cols):
    # static pivots: column static_cols[i] pivots on row static_rows[i] (no search); remaining columns partial pivoting
    done_rows=[]
    order=list(static_cols)+[c for c in range(npiv) if c not in set(static_cols)]
    # work with explicit row set
    nr=W.shape[0]
    avail=np.ones(nr,bool); avail[nsearch:]=False
    prow={}
    for idx,c in enumerate(order):
        if idx < len(static_rows): r=static_rows[idx]
        else:
            cand=np.where(avail)[0]; r=cand[np.argmax(np.abs(W[cand,c]))]
        avail[r]=False; prow[c]=r
        piv=W[r,c]
        others=[i for i in range(nr) if i!=r and (avail[i] or i>=nsearch)]
        l=W[others,c]/piv
        W[others,:]-=np.outer(l,W[r,:])
        W[others,c]=0
    return prow, avail

def test(h,a,d,NMK,om=1.0):
    geo=mo.FlatGeometry(h,a,d,NMK); M=geo.M; K,N,NT=geo.NMK[-1],geo.NMK[-2],geo.size; p=NT-K
    flags=[[1]*M]
    m0=mo.wavenumber(om,h); mk=mo.m_k_array(K,m0,h); Nk=mo.N_k_array(m0,h,mk)
    A=mo.assemble_A(geo,m0,mk,Nk); A11=A[:p,:p].real
    col,pot,vel=mo.block_layout(geo)
    B1=np.stack([mo.assemble_b(geo,fl,m0,mk,Nk)[:p].real for fl in flags],axis=1)
    E=np.zeros((p,N)); E[np.arange(pot[M-1][0],pot[M-1][0]+N),np.arange(N)]=1
    rhs=np.hstack([E,B1])
    ref=linalg.solve(A11,rhs)
    # global static elimination: for each bd, C1^bd columns pivot on the D rows holding them
    W=np.hstack([A11,rhs]).copy()
    dext=geo.dext
    static_rows=[];static_cols=[]
    for bd in range(M):
        Nb=NMK[bd]
        if bd==M-1 or dext[bd]>dext[bd+1]:
            rows=range(pot[bd][0],pot[bd][0]+Nb)   # pot rows have D on region bd
        else:
            rows=range(vel[bd][0],vel[bd][0]+Nb)   # case 2: vel rows have D on region bd (R' -> zero for n=0!)
        static_rows+=list(rows); static_cols+=list(range(col[bd],col[bd]+Nb))
    # check pivots nonzero
    pv=np.array([W[r,c] for r,c in zip(static_rows,static_cols)])
    if np.any(pv==0): return None
    prow,avail=lu_window_static(W,p,p,static_rows,static_cols)
    # solve: each column c has pivot row prow[c]; back substitution in reverse elimination order
    order=list(static_cols)+[c for c in range(p) if c not in set(static_cols)]
    X=np.zeros_like(rhs)
    for c in reversed(order):
        r=prow[c]
        X[c]=(W[r,p:]-W[r,:p]@X+W[r,c]*X[c])/W[r,c]
    n_last = NMK[0] if M==1 else 2*N
    sl=slice(col[M-1],col[M-1]+n_last)
    cvec=mo.c_vector(geo,flags[0]).real
    e1=np.max(np.abs(X[sl]-ref[sl]))/np.max(np.abs(ref[sl]))
    w=cvec@X; wr=cvec@ref
    e2=np.max(np.abs(w-wr))/np.max(np.abs(wr))
    return e1,e2

for name,args in [("mini",(1.001,[0.125,0.25],[0.25,0.125],[10]*3)),("small",(20.0,[5,10,15],[15,10,5],[50]*4)),
                  ("big",(25.0,[10,15,20],[20,15,10],[200]*4)),("bi300",(1.001,[0.5,1.0],[0.5,0.25],[300]*3)),
                  ("tri200",(20.0,[5.,10.,15.],[15.,10.,5.],[200]*4)),("overhang",(1.001,[0.5,1.0],[0.25,0.5],[6,7,8])),
                  ("config3",(1.9,[0.3,0.5,1.0,1.2,1.6],[0.5,0.7,0.8,0.2,0.5],[8,9,10,11,12,13]))]:
    print(name,test(*args))
