"""Probe (round 2): block-bidiagonal ("window") elimination of A11 = A[:p,:p] for the geometry stage.

Unknown blocks u_0 (N_0), u_i = (C1^i, C2^i) (2 N_i); equation blocks B_bd = potential + velocity rows of boundary bd
(reference layout: meem_engine.py:164-276; paper pubs/JFM/FLMguide.tex:1470-1475: "block bi-diagonal ... Thomas algorithm").
Window bd holds the rows left over from window bd-1 plus B_bd, restricted to the columns of u_bd, u_bd+1; partial pivoting
inside the window sees every row that has a non-zero in the pivot column, so the pivots are the ones LAPACK's dense getrf
picks.  The c-vector rows (one per radiating mode) ride through the elimination as extra, never-pivoted rows: when every
unknown is eliminated their right-hand-side part holds -c^T A11^-1 [E | b] = -[w_j | s0_j], so neither G = A11^-1 E nor
the factors are ever stored outside the last window.

Checks H, u, w, s0 of the window algorithm against a dense LAPACK solve of A11 on the golden geometries, staircases and
C4 geometries.   python tools/experiments/exp_band.py [n_c4]
"""
import sys
import warnings

import numpy as np
from scipy import linalg

sys.path.insert(0, '.')
from oracle import meem_oracle as mo  # noqa: E402

warnings.simplefilter("ignore")


def lu_window(W, npiv, nsearch, nstatic=0):
    """In-place partial-pivot elimination of the first npiv columns; pivot rows searched in rows < nsearch only.
    The first nstatic columns take the diagonal entry as pivot without a search (static pivots, k_geom_band / k_geom_coop2)."""
    for k in range(npiv):
        q = k if k < nstatic else k + int(np.argmax(np.abs(W[k:nsearch, k])))
        if q != k:
            W[[k, q]] = W[[q, k]]
        l = W[k + 1:, k] / W[k, k]
        W[k + 1:, k + 1:] -= np.outer(l, W[k, k + 1:])
        W[k + 1:, k] = 0.0


STATIC = False      # windows of case-1 boundaries: potential rows first, C1 block eliminated on its diagonal without a pivot search


def band_reduce(geo, flags_list, A11, B1, cvecs):
    """Window elimination.  Returns (Glast [n_last x (N+R)] = rows of [G | y0] of the last region, w [R x N], s0 [R x R])."""
    M, NMK = geo.M, geo.NMK
    col, pot, vel = mo.block_layout(geo)
    N = NMK[M - 1]
    R = len(flags_list)
    nun = [NMK[0]] + [2 * NMK[i] for i in range(1, M)]
    carried = np.zeros((0, 0))       # rows left over: columns [u_bd | b]
    crow = np.zeros((R, R))          # right-hand-side part of the c rows
    ccur = np.array([c[col[0]:col[0] + nun[0]] for c in cvecs]).reshape(R, nun[0])
    for bd in range(M - 1):
        prow = list(range(pot[bd][0], pot[bd][0] + pot[bd][1]))
        vrow = list(range(vel[bd][0], vel[bd][0] + vel[bd][1]))
        rows = prow + vrow
        n0, n1 = nun[bd], nun[bd + 1]
        nc = carried.shape[0]
        static = STATIC and geo.dext[bd] > geo.dext[bd + 1]
        W = np.zeros((nc + len(rows) + R, n0 + n1 + R))
        r_car, r_new = (len(prow), [0, len(prow) + nc]) if static else (0, [nc, nc + len(prow)])
        if nc:
            W[r_car:r_car + nc, :n0] = carried[:, :n0]
            W[r_car:r_car + nc, n0 + n1:] = carried[:, n0:]
        for rr, r0 in ((prow, r_new[0]), (vrow, r_new[1])):
            W[r0:r0 + len(rr), :n0 + n1] = A11[np.ix_(rr, range(col[bd], col[bd] + n0 + n1))]
            W[r0:r0 + len(rr), n0 + n1:] = B1[rr]
        W[nc + len(rows):, :n0] = ccur
        W[nc + len(rows):, n0:n0 + n1] = np.array([c[col[bd + 1]:col[bd + 1] + n1] for c in cvecs]).reshape(R, n1)
        W[nc + len(rows):, n0 + n1:] = crow
        lu_window(W, n0, nc + len(rows), NMK[bd] if static else 0)
        carried = W[n0:nc + len(rows), n0:].copy()
        ccur = W[nc + len(rows):, n0:n0 + n1].copy()
        crow = W[nc + len(rows):, n0 + n1:].copy()
    bd = M - 1
    rows = list(range(pot[bd][0], pot[bd][0] + pot[bd][1]))
    n0 = nun[bd]
    nc = carried.shape[0]
    assert nc + len(rows) == n0
    W = np.zeros((n0 + R, n0 + N + R))
    r_car, r_pot = (len(rows), 0) if STATIC else (0, nc)
    if nc:
        W[r_car:r_car + nc, :n0] = carried[:, :n0]
        W[r_car:r_car + nc, n0 + N:] = carried[:, n0:]
    W[r_pot:r_pot + len(rows), :n0] = A11[np.ix_(rows, range(col[bd], col[bd] + n0))]
    W[r_pot:r_pot + len(rows), n0:n0 + N] = np.eye(N)
    W[r_pot:r_pot + len(rows), n0 + N:] = B1[rows]
    W[n0:, :n0] = ccur
    W[n0:, n0 + N:] = crow
    lu_window(W, n0, n0, NMK[bd] if STATIC else 0)
    Glast = linalg.solve_triangular(W[:n0, :n0], W[:n0, n0:])
    return Glast, -W[n0:, n0:n0 + N], -W[n0:, n0 + N:]


def check(name, h, a, d, NMK, region_body=None, modes=None):
    geo = mo.FlatGeometry(h, a, d, NMK)
    M = geo.M
    K, N, NT = geo.NMK[-1], geo.NMK[-2], geo.size
    p = NT - K
    region_body = list(range(M)) if region_body is None else region_body
    modes = [0] if modes is None else modes
    flags_list = mo.region_flags_for_modes(region_body, modes)
    m0 = mo.wavenumber(1.0, h)
    mk = mo.m_k_array(K, m0, h)
    Nk = mo.N_k_array(m0, h, mk)
    A = mo.assemble_A(geo, m0, mk, Nk)
    A11 = A[:p, :p].real
    col, pot, vel = mo.block_layout(geo)
    B1 = np.stack([mo.assemble_b(geo, fl, m0, mk, Nk)[:p].real for fl in flags_list], axis=1)
    cvecs = [mo.c_vector(geo, fl).real for fl in flags_list]
    E = np.zeros((p, N))
    E[np.arange(pot[M - 1][0], pot[M - 1][0] + N), np.arange(N)] = 1
    GY = linalg.solve(A11, np.hstack([E, B1]))
    n_last = NMK[0] if M == 1 else 2 * N
    ref_last = GY[col[M - 1]:col[M - 1] + n_last]
    ref_w = np.array([c @ GY[:, :N] for c in cvecs])
    ref_s0 = np.array([c @ GY[:, N:] for c in cvecs])
    Glast, w, s0 = band_reduce(geo, flags_list, A11, B1, cvecs)

    def rel(x, y):
        s = np.max(np.abs(y))
        return float(np.max(np.abs(x - y)) / s) if s > 0 else float(np.max(np.abs(x)))
    errs = (rel(Glast, ref_last), rel(w, ref_w), rel(s0, ref_s0))
    return name, p, errs


if __name__ == "__main__":
    if "--static" in sys.argv:
        sys.argv.remove("--static")
        STATIC = True
        print("static pivots for the C1 block of every case-1 window (potential rows first)")
    cases = [("mini", (1.001, [0.125, 0.25], [0.25, 0.125], [10] * 3)), ("small", (20.0, [5, 10, 15], [15, 10, 5], [50] * 4)),
             ("single", (3.0, [1.0], [0.7], [9, 11])), ("overhang", (1.001, [0.5, 1.0], [0.25, 0.5], [6, 7, 8])),
             ("config3", (1.9, [0.3, 0.5, 1.0, 1.2, 1.6], [0.5, 0.7, 0.8, 0.2, 0.5], [8, 9, 10, 11, 12, 13])),
             ("uneq", (20.0, [5, 10, 15], [15, 10, 5], [30, 12, 25, 40])), ("uneq2", (20.0, [5, 10, 15], [15, 10, 5], [30, 12, 40, 9])),
             ("big", (25.0, [10, 15, 20], [20, 15, 10], [200] * 4))]
    for Ms in (12, 24, 48):
        a = list(np.linspace(1.0, 6.0, Ms))
        d = list(np.linspace(8.0, 1.0, Ms))
        cases.append(("stair%d" % Ms, (10.0, a, d, [8] * (Ms + 1))))
    worst = 0.0
    for name, args in cases:
        nm, p, errs = check(name, *args)
        worst = max(worst, *errs)
        print("%-10s p=%5d  G_last %.2e  w %.2e  s0 %.2e" % (nm, p, *errs))
    # two radiating bodies
    nm, p, errs = check("small_R2", 20.0, [5, 10, 15], [15, 10, 5], [50] * 4, region_body=[0, 1, 1], modes=[0, 1])
    worst = max(worst, *errs)
    print("%-10s p=%5d  G_last %.2e  w %.2e  s0 %.2e" % (nm, p, *errs))
    if len(sys.argv) > 1:
        from bench import c4_geometries
        w4 = 0.0
        for i, g in enumerate(c4_geometries(int(sys.argv[1]))):
            nm, p, errs = check("c4_%d" % i, g["h"], g["a"], g["d"], g["NMK"], region_body=[0] * len(g["a"]))
            w4 = max(w4, *errs)
        print("C4 x%s worst %.2e" % (sys.argv[1], w4))
        worst = max(worst, w4)
    print("worst relative difference vs dense LAPACK solve: %.2e" % worst)
