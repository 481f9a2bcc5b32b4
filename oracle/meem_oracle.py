"""numpy/scipy restatement of OpenFLASH's MEEM hot path.  TEST INFRASTRUCTURE.

This file is the parity oracle: a CPU restatement of the reference algorithm
(dense complex assembly -> LAPACK LU -> reduction to hydrodynamic coefficients)
working on the same flat arrays the reference extracts at its engine boundary
(``h, d[], a[], NMK[], heaving[]``; reference ``meem_engine.py:110-116``).

Parity status: PINNED.  ``tests/test_oracle_golden.py`` checks this module
against (i) fixtures produced by importing the untouched reference in the build
container (``tests/golden/make_golden.py`` -> ``tests/golden/*.npz``), (ii) the
reference's stored known-answer CSVs (``debug_results/DEBUG_*_regenerated_v2.csv``,
copied values in ``tests/golden/debug_nondim_added_mass.json``) and (iii) the
tutorial notebook's printed 2x2 matrices.

Third-party arithmetic: the reference's numerics live in scipy (pinned
``scipy>=1.7,<1.14`` in its pyproject, 1.18.1 installed here): AMOS Bessel
(``scipy.special.ive/kve/hankel1``), ``scipy.optimize.root_scalar`` (newton,
brentq) and LAPACK ``zgetrf/zgetrs``.  The oracle calls the very same scipy
entry points, in the same operation order as the reference formulas, so it
reproduces the reference to rounding noise (<= 1e-14 relative on coefficients).

Only tests/, __graft_entry__.smoke() and bench.py's cpu-baseline legs import this.
Reference line numbers cite /root/reference/package/src/openflash/.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Sequence

import numpy as np
from numpy import pi
from scipy import linalg
from scipy.optimize import root_scalar
from scipy.special import hankel1, ive, kve

G = 9.81            # multi_constants.py:4
RHO = 1023          # multi_constants.py:5
M0_H_THRESH = 14    # multi_equations.py:18


# --------------------------------------------------------------------------- geometry
@dataclass
class FlatGeometry:
    """Arrays the reference flattens out of its domain list (meem_engine.py:110-116).

    ``a`` and ``d`` hold the M interior regions only (the reference appends
    ``a=inf, d=0`` for the exterior); ``NMK`` has M+1 entries.
    """
    h: float
    a: np.ndarray
    d: np.ndarray
    NMK: Sequence[int]

    def __post_init__(self):
        self.a = np.asarray(self.a, dtype=float)
        self.d = np.asarray(self.d, dtype=float)
        self.NMK = [int(x) for x in self.NMK]
        assert len(self.a) == len(self.d) == len(self.NMK) - 1

    @property
    def M(self):
        return len(self.a)

    @property
    def size(self):            # meem_engine.py:116
        N = self.NMK
        return N[0] + N[-1] + 2 * sum(N[1:-1])

    @property
    def dext(self):            # d with the exterior's 0.0 appended
        return np.append(self.d, 0.0)


def lambda_ni(n, i, h, d):
    """multi_equations.py:35-36 — operation order (n*pi)/(h-d_i) matters."""
    return n * pi / (h - d[i])


# --------------------------------------------------------------------------- dispersion
def omega_of_m0(m0, h, g=G):
    """multi_equations.py:20-24."""
    if m0 == np.inf:
        return np.inf
    return np.sqrt(m0 * np.tanh(m0 * h) * g)


def wavenumber(omega, h):
    """multi_equations.py:26-30 (scipy newton, x0=2, finite-difference derivative)."""
    if omega == np.inf:
        return np.inf
    err = lambda m0: m0 * np.tanh(h * m0) - omega ** 2 / G
    return root_scalar(err, x0=2, method="newton").root


def m_k_entry(k, m0, h):
    """multi_equations.py:42-61 (brentq on x tan x + m0h tanh(m0h), scipy defaults)."""
    if k == 0:
        return m0
    if m0 == np.inf:
        return ((k - 1 / 2) * pi) / h
    err = lambda x: x * np.tan(x) + m0 * h * np.tanh(m0 * h)
    lo = np.nextafter(pi * (k - 1 / 2), np.inf)
    hi = np.nextafter(pi * k, np.inf)
    res = root_scalar(err, x0=(hi + lo) / 2, method="brentq", bracket=[lo, hi])
    return res.root / h


def m_k_array(K, m0, h):
    """meem_engine.py:40."""
    return np.array([m_k_entry(k, m0, h) for k in range(K)])


def N_k_array(m0, h, mk):
    """multi_equations.py:379-384 / meem_engine.py:41."""
    K = len(mk)
    out = np.empty(K)
    if m0 == np.inf:
        out[:] = 1 / 2
        return out
    with np.errstate(over="ignore"):
        out[0] = 1 / 2 * (1 + np.sinh(2 * m0 * h) / (2 * m0 * h))
    if K > 1:
        m = mk[1:]
        out[1:] = 1 / 2 * (1 + np.sin(2 * m * h) / (2 * m * h))
    return out


# --------------------------------------------------------------------------- coupling integrals
def I_nm_table(bd, geo: FlatGeometry):
    """multi_equations.py:71-93 evaluated for the whole (N_bd x N_bd+1) table."""
    h, d = geo.h, geo.dext
    N, Mm = geo.NMK[bd], geo.NMK[bd + 1]
    dj = max(d[bd], d[bd + 1])
    L = h - dj
    n = np.arange(N, dtype=float)[:, None]
    m = np.arange(Mm, dtype=float)[None, :]
    lam1 = lambda_ni(n, bd, h, d)
    lam2 = lambda_ni(m, bd + 1, h, d)
    out = np.zeros((N, Mm))
    with np.errstate(divide="ignore", invalid="ignore"):
        s = lam1 + lam2
        frac1 = np.sin(s * L) / s
        df = lam1 - lam2
        frac2 = np.where(lam1 == lam2, L, np.sin(df * L) / df)
        out[:, :] = frac1 + frac2
        # n == 0, m >= 1
        row0 = np.sqrt(2) * np.sin(lam2 * L) / lam2
        out[0:1, :] = 0.0 if dj == d[bd + 1] else row0
        # n >= 1, m == 0
        col0 = np.sqrt(2) * np.sin(lam1 * L) / lam1
        out[:, 0:1] = 0.0 if dj == d[bd] else col0
    out[0, 0] = L
    return out


def I_mk_table(geo: FlatGeometry, m0, mk, Nk):
    """multi_equations.py:96-124 for the (N_{M-1} x K) table of the last boundary."""
    h, d = geo.h, geo.dext
    i = geo.M - 1
    N, K = geo.NMK[i], geo.NMK[-1]
    dj = d[i]
    L = h - dj
    out = np.zeros((N, K))
    lam = lambda_ni(np.arange(N, dtype=float), i, h, d)          # (N,)
    # m >= 1, k >= 1
    if N > 1 and K > 1:
        l1 = lam[1:, None]
        mkk = mk[None, 1:]
        with np.errstate(divide="ignore", invalid="ignore"):
            frac1 = np.sin((mkk + l1) * L) / (mkk + l1)
            frac2 = np.sin((mkk - l1) * L) / (mkk - l1)
            gen = np.sqrt(2 / Nk[None, 1:]) * (frac1 + frac2) / 2
            eq = np.sqrt(2 / Nk[None, 1:]) * L / 2
        out[1:, 1:] = np.where(np.abs(mkk) == l1, eq, gen)
    # m == 0, k >= 1
    if K > 1:
        out[0, 1:] = (1 / np.sqrt(Nk[1:])) * np.sin(mk[1:] * L) / mk[1:]
    # k == 0 column
    if m0 == np.inf:
        out[:, 0] = 0.0
    elif m0 * h < M0_H_THRESH:
        out[0, 0] = (1 / np.sqrt(Nk[0])) * np.sinh(m0 * L) / m0
        if N > 1:
            sgn = (-1.0) ** np.arange(1, N)
            num = sgn * np.sqrt(2) * (1 / np.sqrt(Nk[0])) * m0 * np.sinh(m0 * L)
            out[1:, 0] = num / (m0 ** 2 + lam[1:] ** 2)
    else:
        tail = np.exp(-m0 * dj) - np.exp(m0 * dj - 2 * m0 * h)
        out[0, 0] = np.sqrt(2 * h / m0) * tail
        if N > 1:
            sgn = (-1.0) ** np.arange(1, N)
            num = sgn * 2 * np.sqrt(h * m0 ** 3) * tail
            out[1:, 0] = num / (m0 ** 2 + lam[1:] ** 2)
    return out


# --------------------------------------------------------------------------- radial functions
def R_1n(n, r, i, geo: FlatGeometry):
    """multi_equations.py:192-218."""
    n = np.asarray(n, dtype=float)
    lam = np.where(n == 0, 1.0, lambda_ni(n, i, geo.h, geo.dext))
    ai = geo.a[i]
    with np.errstate(all="ignore"):
        val = (ive(0, lam * r) / ive(0, lam * ai)) * np.exp(lam * (r - ai))
    val = np.where(r == ai, 1.0, val)
    return np.where(n == 0, 0.5, val)


def diff_R_1n(n, r, i, geo: FlatGeometry):
    """multi_equations.py:220-241."""
    n = np.asarray(n, dtype=float)
    lam = np.where(n == 0, 1.0, lambda_ni(n, i, geo.h, geo.dext))
    ai = geo.a[i]
    with np.errstate(all="ignore"):
        val = ((lam * ive(1, lam * r)) / ive(0, lam * ai)) * np.exp(lam * (r - ai))
    return np.where(n == 0, 0.0, val)


def R_2n(n, r, i, geo: FlatGeometry):
    """multi_equations.py:246-279."""
    if i == 0:
        raise ValueError("R_2n function is not defined for the innermost region (i=0).")
    n = np.asarray(n, dtype=float)
    lam = np.where(n == 0, 1.0, lambda_ni(n, i, geo.h, geo.dext))
    ao = geo.a[i]
    with np.errstate(all="ignore"):
        zero = 0.5 * np.log(np.divide(r, ao))
        val = (kve(0, lam * r) / kve(0, lam * ao)) * np.exp(lam * (ao - r))
    val = np.where(r == ao, 1.0, val)
    return np.where(n == 0, zero, val)


def diff_R_2n(n, r, i, geo: FlatGeometry):
    """multi_equations.py:281-297."""
    n = np.asarray(n, dtype=float)
    lam = np.where(n == 0, 1.0, lambda_ni(n, i, geo.h, geo.dext))
    ao = geo.a[i]
    with np.errstate(all="ignore"):
        val = ((-lam * kve(1, lam * r)) / kve(0, lam * ao)) * np.exp(lam * (ao - r))
    return np.where(n == 0, 1.0 / (2 * r), val)


def diff_Lambda_k_at_aM(geo: FlatGeometry, m0, mk):
    """multi_equations.py:352-374 at r = a_{M-1} (the only radius the engine uses)."""
    K = geo.NMK[-1]
    aM = geo.a[-1]
    out = np.zeros(K, dtype=complex)
    if m0 == np.inf:
        out[0] = 1.0
    else:
        out[0] = -(m0 * hankel1(1, m0 * aM)) / hankel1(0, m0 * aM)
    if K > 1:
        m = mk[1:]
        with np.errstate(all="ignore"):
            out[1:] = (-(m * kve(1, m * aM)) / kve(0, m * aM)) * np.exp(m * (aM - aM))
    return out


# --------------------------------------------------------------------------- layout
def block_layout(geo: FlatGeometry):
    """Row/column bookkeeping of meem_engine.py:164-276.

    Returns (col_start[region], pot_rows[bd]=(row0,height), vel_rows[bd]=(row0,height)).
    Column order: [C1^0 | C1^1 C2^1 | ... | C1^{M-1} C2^{M-1} | C^e].
    """
    NMK, M, d = geo.NMK, geo.M, geo.dext
    col = [0]
    for i in range(M):
        col.append(col[-1] + (NMK[0] if i == 0 else 2 * NMK[i]))
    pot, vel = [], []
    row = 0
    for bd in range(M):
        if bd == M - 1:
            hgt = NMK[bd]
        else:
            hgt = NMK[bd] if d[bd] > d[bd + 1] else NMK[bd + 1]
        pot.append((row, hgt)); row += hgt
    for bd in range(M):
        if bd == M - 1:
            hgt = NMK[bd + 1]
        else:
            hgt = NMK[bd] if d[bd] <= d[bd + 1] else NMK[bd + 1]
        vel.append((row, hgt)); row += hgt
    assert row == geo.size
    return col, pot, vel


def assemble_A(geo: FlatGeometry, m0, mk=None, Nk=None):
    """A matrix of meem_engine.py:100-276 + assemble_A_multi (:44-55), complex (N_T,N_T)."""
    h, d, a, NMK, M = geo.h, geo.dext, geo.a, geo.NMK, geo.M
    if mk is None:
        mk = m_k_array(NMK[-1], m0, h)
    if Nk is None:
        Nk = N_k_array(m0, h, mk)
    A = np.zeros((geo.size, geo.size), dtype=complex)
    col, pot, vel = block_layout(geo)
    Inm = [I_nm_table(bd, geo) for bd in range(M - 1)]
    Imk = I_mk_table(geo, m0, mk, Nk)

    def rad(fn, region, bd):          # radial vector of `region` at r = a[bd]
        return fn(np.arange(NMK[region]), a[bd], region, geo)

    def diag_block(fn, region, bd, sign):        # me.py:503-509 / :538-545
        return sign * (h - d[region]) * np.diag(rad(fn, region, bd))

    def dense_block(fn, bd, left, sign):         # me.py:511-527 / :547-561
        I = Inm[bd]
        if left:
            region, I = bd, I.T
        else:
            region = bd + 1
        return sign * np.outer(np.ones(I.shape[0]), rad(fn, region, bd)) * I

    for bd in range(M):
        N, Mm = NMK[bd], NMK[bd + 1]
        # ---------------- potential matching rows (meem_engine.py:164-219)
        r0, hgt = pot[bd]
        c1L = col[bd]                       # C1 of region bd
        c2L = c1L + N                       # C2 of region bd (bd >= 1)
        cR = col[bd + 1]                    # first column of region bd+1 (or exterior)
        if bd == M - 1:
            A[r0:r0 + N, c1L:c1L + N] = diag_block(R_1n, bd, bd, +1)
            if bd > 0:
                A[r0:r0 + N, c2L:c2L + N] = diag_block(R_2n, bd, bd, +1)
            A[r0:r0 + N, cR:cR + Mm] = (-1) * np.outer(np.ones(N), np.ones(Mm)) * Imk   # me.py:529-536, Lambda_k(a_M)=1
        else:
            left_diag = d[bd] > d[bd + 1]
            if left_diag:
                A[r0:r0 + N, c1L:c1L + N] = diag_block(R_1n, bd, bd, +1)
                if bd > 0:
                    A[r0:r0 + N, c2L:c2L + N] = diag_block(R_2n, bd, bd, +1)
                A[r0:r0 + N, cR:cR + Mm] = dense_block(R_1n, bd, False, -1)
                A[r0:r0 + N, cR + Mm:cR + 2 * Mm] = dense_block(R_2n, bd, False, -1)
            else:
                A[r0:r0 + Mm, c1L:c1L + N] = dense_block(R_1n, bd, True, +1)
                if bd > 0:
                    A[r0:r0 + Mm, c2L:c2L + N] = dense_block(R_2n, bd, True, +1)
                A[r0:r0 + Mm, cR:cR + Mm] = diag_block(R_1n, bd + 1, bd, -1)
                A[r0:r0 + Mm, cR + Mm:cR + 2 * Mm] = diag_block(R_2n, bd + 1, bd, -1)
        # ---------------- velocity matching rows (meem_engine.py:221-276)
        r0, hgt = vel[bd]
        if bd == M - 1:
            dR1 = rad(diff_R_1n, bd, bd)
            A[r0:r0 + Mm, c1L:c1L + N] = (-1) * np.outer(np.ones(Mm), dR1) * Imk.T      # me.py:570-576
            if bd > 0:
                dR2 = rad(diff_R_2n, bd, bd)
                A[r0:r0 + Mm, c2L:c2L + N] = (-1) * np.outer(np.ones(Mm), dR2) * Imk.T
            A[r0:r0 + Mm, cR:cR + Mm] = h * np.diag(diff_Lambda_k_at_aM(geo, m0, mk))  # me.py:563-568
        else:
            left_diag = d[bd] <= d[bd + 1]
            if left_diag:
                A[r0:r0 + N, c1L:c1L + N] = diag_block(diff_R_1n, bd, bd, -1)
                if bd > 0:
                    A[r0:r0 + N, c2L:c2L + N] = diag_block(diff_R_2n, bd, bd, -1)
                A[r0:r0 + N, cR:cR + Mm] = dense_block(diff_R_1n, bd, False, +1)
                A[r0:r0 + N, cR + Mm:cR + 2 * Mm] = dense_block(diff_R_2n, bd, False, +1)
            else:
                A[r0:r0 + Mm, c1L:c1L + N] = dense_block(diff_R_1n, bd, True, -1)
                if bd > 0:
                    A[r0:r0 + Mm, c2L:c2L + N] = dense_block(diff_R_2n, bd, True, -1)
                A[r0:r0 + Mm, cR:cR + Mm] = diag_block(diff_R_1n, bd + 1, bd, +1)
                A[r0:r0 + Mm, cR + Mm:cR + 2 * Mm] = diag_block(diff_R_2n, bd + 1, bd, +1)
    return A


def assemble_b(geo: FlatGeometry, heaving, m0, mk=None, Nk=None):
    """b vector: multi_equations.py:129-174, meem_engine.py:278-301, problem_cache.py:93-153."""
    h, d, a, NMK, M = geo.h, geo.dext, geo.a, geo.NMK, geo.M
    hv = [float(x) for x in heaving] + [0.0]
    if mk is None:
        mk = m_k_array(NMK[-1], m0, h)
    if Nk is None:
        Nk = N_k_array(m0, h, mk)
    b = np.zeros(geo.size, dtype=complex)
    idx = 0
    sq2 = np.sqrt(2)
    for bd in range(M):                                   # potential rows
        if bd == M - 1:
            i = bd
            const = -hv[i] / (h - d[i])
            for n in range(NMK[bd]):
                if n == 0:
                    b[idx] = const * 1 / 2 * ((h - d[i]) ** 3 / 3 - (h - d[i]) * a[i] ** 2 / 2)
                else:
                    b[idx] = sq2 * (h - d[i]) * const * ((-1) ** n) / (lambda_ni(n, i, h, d) ** 2)
                idx += 1
        else:
            i = bd
            j = i + (d[i] <= d[i + 1])
            const = hv[i + 1] / (h - d[i + 1]) - hv[i] / (h - d[i])
            for n in range(NMK[j]):
                if n == 0:
                    b[idx] = const * 1 / 2 * ((h - d[j]) ** 3 / 3 - (h - d[j]) * a[i] ** 2 / 2)
                else:
                    b[idx] = sq2 * (h - d[j]) * const * ((-1) ** n) / (lambda_ni(n, j, h, d) ** 2)
                idx += 1
    for bd in range(M):                                   # velocity rows
        i = bd
        if bd == M - 1:
            const = -hv[i] * a[i] / (2 * (h - d[i]))
            for k in range(NMK[-1]):
                if k == 0:
                    if m0 == np.inf:
                        val = 0.0
                    elif m0 * h < M0_H_THRESH:
                        val = const * (1 / np.sqrt(Nk[0])) * np.sinh(m0 * (h - d[i])) / m0
                    else:
                        val = const * np.sqrt(2 * h / m0) * (np.exp(-m0 * d[i]) - np.exp(m0 * d[i] - 2 * m0 * h))
                else:
                    val = const * (1 / np.sqrt(Nk[k])) * np.sin(mk[k] * (h - d[i])) / mk[k]
                b[idx] = val
                idx += 1
        else:
            nrows = NMK[bd + (d[bd] > d[bd + 1])]
            for n in range(nrows):
                if n == 0:
                    val = (hv[i + 1] - hv[i]) * (a[i] / 2)
                elif d[i] > d[i + 1]:
                    if hv[i]:
                        lam = lambda_ni(n, i + 1, h, d)
                        val = (-sq2 * a[i] * np.sin(lam * (h - d[i]))) / (2 * (h - d[i]) * lam)
                    else:
                        val = 0.0
                else:
                    if hv[i + 1]:
                        lam = lambda_ni(n, i, h, d)
                        val = (sq2 * a[i] * np.sin(lam * (h - d[i + 1]))) / (2 * (h - d[i + 1]) * lam)
                    else:
                        val = 0.0
                b[idx] = val
                idx += 1
    return b


# --------------------------------------------------------------------------- reduction
def int_R_1n(i, n, geo: FlatGeometry):
    """multi_equations.py:423-433."""
    a, h, d = geo.a, geo.h, geo.dext
    if n == 0:
        inner = 0 if i == 0 else a[i - 1]
        return a[i] ** 2 / 4 - inner ** 2 / 4
    lam = lambda_ni(n, i, h, d)
    bottom = lam * ive(0, lam * a[i])
    inner_term = 0 if i == 0 else (a[i - 1] * ive(1, lam * a[i - 1]) / bottom) * np.exp(lam * (a[i - 1] - a[i]))
    outer_term = (a[i] * ive(1, lam * a[i]) / bottom) * np.exp(lam * (a[i] - a[i]))
    return outer_term - inner_term


def int_R_2n(i, n, geo: FlatGeometry):
    """multi_equations.py:436-448."""
    a, h, d = geo.a, geo.h, geo.dext
    if i == 0:
        raise ValueError("i cannot be 0")
    if n == 0:
        return (a[i - 1] ** 2 * (2 * np.log(a[i] / a[i - 1]) + 1) - a[i] ** 2) / 8
    lam = lambda_ni(n, i, h, d)
    outer_term = a[i] * kve(1, lam * a[i])
    inner_term = a[i - 1] * kve(1, lam * a[i - 1])
    bottom = -lam * kve(0, lam * a[i])
    with np.errstate(all="ignore"):
        return (outer_term / bottom) * np.exp(lam * (a[i] - a[i])) - (inner_term / bottom) * np.exp(lam * (a[i] - a[i - 1]))


def int_phi_p_i(i, geo: FlatGeometry):
    """multi_equations.py:451-457."""
    a, h, d = geo.a, geo.h, geo.dext
    denom = 16 * (h - d[i])
    if i == 0:
        num = a[i] ** 2 * (4 * (h - d[i]) ** 2 - a[i] ** 2)
    else:
        num = (a[i] ** 2 * (4 * (h - d[i]) ** 2 - a[i] ** 2) - a[i - 1] ** 2 * (4 * (h - d[i]) ** 2 - a[i - 1] ** 2))
    return num / denom


def z_n_d(n):
    """multi_equations.py:460-464."""
    return 1 if n == 0 else np.sqrt(2) * (-1) ** n


def c_vector(geo: FlatGeometry, heaving):
    """meem_engine.py:65-98."""
    NMK, M = geo.NMK, geo.M
    c = np.zeros(geo.size - NMK[-1], dtype=complex)
    for n in range(NMK[0]):
        c[n] = heaving[0] * int_R_1n(0, n, geo) * z_n_d(n)
    colp = NMK[0]
    for i in range(1, M):
        Ni = NMK[i]
        for m in range(Ni):
            c[colp + m] = heaving[i] * int_R_1n(i, m, geo) * z_n_d(m)
            c[colp + Ni + m] = heaving[i] * int_R_2n(i, m, geo) * z_n_d(m)
        colp += 2 * Ni
    return c


def excitation_phase(X, geo: FlatGeometry, m0):
    """multi_equations.py:467-471."""
    if m0 == np.inf:
        return -(pi / 2)
    coeff = X[-geo.NMK[-1]]
    return -(pi / 2) + np.angle(coeff) - np.angle(hankel1(0, m0 * geo.a[-1]))


def excitation_force(damping, m0, h):
    """multi_equations.py:473-475 (uses the module constants rho, g)."""
    const = np.tanh(m0 * h) + m0 * h * (1 - (np.tanh(m0 * h)) ** 2)
    with np.errstate(all="ignore"):
        return ((2 * const * RHO * (G ** 2) * damping) / (omega_of_m0(m0, h, G) * m0)) ** (1 / 2)


def hydro_coefficients(geo: FlatGeometry, X, m0, mode_flags, rho=RHO):
    """meem_engine.py:326-393: one dict per mode j for a solution X of mode i."""
    out = []
    for flags in mode_flags:
        c = c_vector(geo, flags)
        phi = sum(flags[i] * int_phi_p_i(i, geo) for i in range(geo.M))
        coef = 2 * pi * (np.dot(c, X[:-geo.NMK[-1]]) + phi)
        re = coef.real * rho
        im = 0 if m0 == np.inf else coef.imag * omega_of_m0(m0, geo.h, G) * rho
        out.append(dict(real=re, imag=im,
                        excitation_phase=excitation_phase(X, geo, m0),
                        excitation_force=excitation_force(im, m0, geo.h),
                        c_vector=c))
    return out


def region_flags_for_modes(region_body: Sequence[int], modes: Sequence[int]):
    """Per-mode heaving flags per region: body b moves => all its regions move
    (meem_engine.py:350-367, :618-627)."""
    rb = np.asarray(region_body)
    return [[1 if rb[r] == b else 0 for r in range(len(rb))] for b in modes]


# --------------------------------------------------------------------------- the sweep
def solve_single(geo: FlatGeometry, heaving, m0):
    """meem_engine.py:307-313 (scipy.linalg.solve)."""
    mk = m_k_array(geo.NMK[-1], m0, geo.h)
    Nk = N_k_array(m0, geo.h, mk)
    A = assemble_A(geo, m0, mk, Nk)
    b = assemble_b(geo, heaving, m0, mk, Nk)
    return linalg.solve(A, b)


def run_sweep(geo: FlatGeometry, omegas, region_body, modes, rho=RHO, want_X=False, m0_override=None):
    """Numerics of MEEMEngine.run_and_store_results (meem_engine.py:555-690).

    Returns dict(added_mass[F,R,R], damping[F,R,R], excitation_force[F,R],
    excitation_phase[F,R], m0[F], X[F,R,N_T] if want_X) with the reference's
    index convention [freq, j (force on body j), i (body i moving)].
    """
    omegas = np.asarray(omegas, dtype=float)
    F, R = len(omegas), len(modes)
    flags = region_flags_for_modes(region_body, modes)
    K = geo.NMK[-1]
    AM = np.full((F, R, R), np.nan)
    DP = np.full((F, R, R), np.nan)
    EF = np.full((F, R), np.nan)
    EP = np.full((F, R), np.nan)
    M0 = np.full(F, np.nan)
    X_all = np.full((F, R, geo.size), np.nan + 0j, dtype=complex) if want_X else None
    cvecs = [c_vector(geo, f) for f in flags]
    phis = [sum(f[i] * int_phi_p_i(i, geo) for i in range(geo.M)) for f in flags]
    for fi, om in enumerate(omegas):
        m0 = wavenumber(om, geo.h) if m0_override is None else m0_override[fi]
        M0[fi] = m0
        mk = m_k_array(K, m0, geo.h)
        Nk = N_k_array(m0, geo.h, mk)
        A = assemble_A(geo, m0, mk, Nk)
        with np.errstate(all="ignore"):
            try:
                import warnings
                with warnings.catch_warnings():
                    warnings.simplefilter("ignore")
                    lu = linalg.lu_factor(A, check_finite=False)
            except Exception:
                continue
            for i in range(R):
                b = assemble_b(geo, flags[i], m0, mk, Nk)
                X = linalg.lu_solve(lu, b, check_finite=False)
                if want_X:
                    X_all[fi, i] = X
                for j in range(R):
                    coef = 2 * pi * (np.dot(cvecs[j], X[:-K]) + phis[j])
                    AM[fi, j, i] = coef.real * rho
                    im = 0 if m0 == np.inf else coef.imag * omega_of_m0(m0, geo.h, G) * rho
                    DP[fi, j, i] = im
                    if i == j:
                        EF[fi, j] = excitation_force(im, m0, geo.h)
                        EP[fi, j] = excitation_phase(X, geo, m0)
    out = dict(added_mass=AM, damping=DP, excitation_force=EF, excitation_phase=EP, m0=M0)
    if want_X:
        out["X"] = X_all
    return out


# --------------------------------------------------------------------------- fields
def make_R_Z(a, h, d, sharp, spatial_res, R_range=None, Z_range=None):
    """multi_equations.py:478-501."""
    a = np.asarray(a, dtype=float)
    d = np.asarray(d, dtype=float)
    if R_range is not None:
        r_vec = np.asarray(R_range, dtype=float)
    else:
        rmin = (2 * a[-1] / spatial_res) if sharp else 0.0
        r_vec = np.linspace(rmin, 2 * a[-1], spatial_res)
    z_vec = np.asarray(Z_range, dtype=float) if Z_range is not None else np.linspace(0, -h, spatial_res)
    if sharp:
        a_eps = 1.0e-4
        for i in range(len(a)):
            r_vec = np.append(r_vec, a[i] * (1 - a_eps))
            r_vec = np.append(r_vec, a[i] * (1 + a_eps))
        r_vec = np.unique(r_vec)
        for i in range(len(d)):
            z_vec = np.append(z_vec, -d[i])
        z_vec = np.unique(z_vec)
    return np.meshgrid(r_vec, z_vec, indexing="ij")


def _Z_n_i(n, z, i, geo):
    """multi_equations.py:301-308."""
    n = np.asarray(n, dtype=float)
    lam = np.where(n == 0, 0.0, lambda_ni(n, i, geo.h, geo.dext))
    return np.where(n == 0, 1.0, np.sqrt(2) * np.cos(lam * (z + geo.h)))


def _diff_Z_n_i(n, z, i, geo):
    """multi_equations.py:310-318."""
    n = np.asarray(n, dtype=float)
    lam = np.where(n == 0, 0.0, lambda_ni(n, i, geo.h, geo.dext))
    return np.where(n == 0, 0.0, -lam * np.sqrt(2) * np.sin(lam * (z + geo.h)))


def _Lambda_k(k, r, m0, geo, mk, diff=False):
    """multi_equations.py:322-374 (finite m0)."""
    k = np.asarray(k)
    aM = geo.a[-1]
    safe = np.where(k == 0, 1.0, mk[k])
    with np.errstate(all="ignore"):
        if diff:
            zero = -(m0 * hankel1(1, m0 * r)) / hankel1(0, m0 * aM)
            non = ((-(safe * kve(1, safe * r))) / kve(0, safe * aM)) * np.exp(safe * (aM - r))
            return np.where(k == 0, zero, non)
        zero = hankel1(0, m0 * r) / hankel1(0, m0 * aM)
        non = (kve(0, safe * r) / kve(0, safe * aM)) * np.exp(safe * (aM - r))
    return np.where(r == aM, 1.0, np.where(k == 0, zero, non))


def _Z_k_e(k, z, m0, h, mk, Nk, diff=False):
    """multi_equations.py:388-416."""
    k = np.asarray(k)
    if not diff:
        non = (1 / np.sqrt(Nk[k])) * np.cos(mk[k] * (z + h))
        if m0 * h < M0_H_THRESH:
            zero = (1 / np.sqrt(Nk[0])) * np.cosh(m0 * (z + h))
        else:
            zero = np.sqrt(2 * m0 * h) * (np.exp(m0 * z) + np.exp(-m0 * (z + 2 * h)))
    else:
        non = -(1 / np.sqrt(Nk[k])) * mk[k] * np.sin(mk[k] * (z + h))
        if m0 * h < M0_H_THRESH:
            zero = (1 / np.sqrt(Nk[0])) * m0 * np.sinh(m0 * (z + h))
        else:
            zero = m0 * np.sqrt(2 * h * m0) * (np.exp(m0 * z) - np.exp(-m0 * (z + 2 * h)))
    return np.where(k == 0, zero, non)


def fields(geo: FlatGeometry, X, m0, heaving, R, Z):
    """Potential and velocity fields at the points (R, Z) (any shape): MEEMEngine.calculate_potentials
    (meem_engine.py:395-467) and calculate_velocities (:478-553) with the reference's masks.
    `heaving` = per-region flags.  Returns dict(phiH, phiP, phi, vrH, vzH, vrP, vzP, vr, vz)."""
    h, a, d, NMK, M = geo.h, geo.a, geo.d, geo.NMK, geo.M
    mk = m_k_array(NMK[-1], m0, h)
    Nk = N_k_array(m0, h, mk)
    R = np.asarray(R, dtype=float)
    Z = np.asarray(Z, dtype=float)
    Cs, col = [], 0
    for i in range(M + 1):
        w = NMK[i] if i in (0, M) else 2 * NMK[i]
        Cs.append(np.asarray(X)[col:col + w])
        col += w
    rad = [R <= a[0]] + [(R > a[i - 1]) & (R <= a[i]) for i in range(1, M)] + [R > a[-1]]
    nanc = np.nan + 1j * np.nan
    phiH = np.full(R.shape, nanc, dtype=complex)
    phiP = np.full(R.shape, nanc, dtype=complex)
    vrH = np.full(R.shape, np.nan, dtype=complex)
    vzH = np.full(R.shape, np.nan, dtype=complex)
    vrP = np.zeros(R.shape, dtype=complex)
    vzP = np.zeros(R.shape, dtype=complex)
    for i in range(M + 1):
        reg = rad[i]
        if not np.any(reg):
            continue
        r, z = R[reg][None, :], Z[reg][None, :]
        n = np.arange(NMK[i])[:, None]
        if i == M:
            c = Cs[i][:, None]
            pH = np.sum(c * _Lambda_k(n, r, m0, geo, mk) * _Z_k_e(n, z, m0, h, mk, Nk), axis=0)
            vr = np.sum(c * _Lambda_k(n, r, m0, geo, mk, diff=True) * _Z_k_e(n, z, m0, h, mk, Nk), axis=0)
            vz = np.sum(c * _Lambda_k(n, r, m0, geo, mk) * _Z_k_e(n, z, m0, h, mk, Nk, diff=True), axis=0)
            pP = np.zeros_like(pH)
            pmask = np.ones(r.shape[1], dtype=bool)
        else:
            c1 = Cs[i][:NMK[i], None]
            t, dt = c1 * R_1n(n, r, i, geo), c1 * diff_R_1n(n, r, i, geo)
            if i > 0:
                c2 = Cs[i][NMK[i]:, None]
                t, dt = t + c2 * R_2n(n, r, i, geo), dt + c2 * diff_R_2n(n, r, i, geo)
            pH = np.sum(t * _Z_n_i(n, z, i, geo), axis=0)
            vr = np.sum(dt * _Z_n_i(n, z, i, geo), axis=0)
            vz = np.sum(t * _diff_Z_n_i(n, z, i, geo), axis=0)
            pP = heaving[i] * ((1 / (2 * (h - d[i]))) * ((z[0] + h) ** 2 - (r[0] ** 2) / 2))      # me.py:179-180
            if heaving[i]:
                vrP[reg] = heaving[i] * (-r[0] / (2 * (h - d[i])))                                # me.py:182-183
                vzP[reg] = heaving[i] * ((z[0] + h) / (h - d[i]))                                 # me.py:185-186
            pmask = z[0] < -d[i]
        idx = np.where(reg)
        sel = tuple(ix[pmask] for ix in idx)
        phiH[sel] = pH[pmask]
        phiP[sel] = pP[pmask]
        vrH[reg], vzH[reg] = vr, vz
    vr_t, vz_t = vrH + vrP, vzH + vzP
    for i in range(M):
        body = rad[i] & (Z > -d[i])
        vr_t[body] = np.nan
        vz_t[body] = np.nan
    return dict(phiH=phiH, phiP=phiP, phi=phiH + phiP, vrH=vrH, vzH=vzH, vrP=vrP, vzP=vzP, vr=vr_t, vz=vz_t)
