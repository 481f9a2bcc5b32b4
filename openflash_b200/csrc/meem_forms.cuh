// Closed forms of the MEEM system on flat geometry arrays: coupling integrals, radial tables,
// block layout, A / b fill, c-vector ingredients.  Follows the reference's operation order where
// exact-equality branches depend on it.  Line numbers cite
// /root/reference/package/src/openflash/{multi_equations.py = me.py, meem_engine.py = eng.py}.
//
// All fill_* routines are written as thread-strided loops over (tid, nthr) with no internal
// synchronisation, so the same code runs as a CTA-cooperative device routine and, with
// tid = 0 / nthr = 1, as plain C++ in the CPU unit tests.
#pragma once
#include "meem_math.cuh"

namespace meem {

constexpr int kMaxRegions = 64;   // interior regions per geometry (M)
constexpr int kTabVecs = 10;      // radial-table vectors per region

struct cplx {
    double re, im;
};
MEEM_HD cplx cmake(double re, double im) { cplx z; z.re = re; z.im = im; return z; }
MEEM_HD cplx operator+(cplx a, cplx b) { return cmake(a.re + b.re, a.im + b.im); }
MEEM_HD cplx operator-(cplx a, cplx b) { return cmake(a.re - b.re, a.im - b.im); }
MEEM_HD cplx operator*(cplx a, cplx b) { return cmake(a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re); }
MEEM_HD cplx operator/(cplx a, cplx b) {   // Smith's algorithm (what numpy / LAPACK zladiv-style division does)
    if (fabs(b.re) >= fabs(b.im)) {
        double r = b.im / b.re, den = b.re + b.im * r;
        return cmake((a.re + a.im * r) / den, (a.im - a.re * r) / den);
    }
    double r = b.re / b.im, den = b.re * r + b.im;
    return cmake((a.re * r + a.im) / den, (a.im * r - a.re) / den);
}
MEEM_HD double cabs1(cplx a) { return fabs(a.re) + fabs(a.im); }   // LAPACK izamax pivot measure
MEEM_HD double cabs1(double a) { return fabs(a); }
MEEM_HD void set_real(double& dst, double v) { dst = v; }
MEEM_HD void set_real(cplx& dst, double v) { dst.re = v; dst.im = 0.0; }

// One geometry, flattened exactly as the reference does at its engine boundary (eng.py:110-116).
struct Geom {
    int M;              // interior regions = boundaries
    double h;
    const double* a;    // [M+1] radii, a[M] = inf (unused)
    const double* d;    // [M+1] drafts, d[M] = 0
    const int* nmk;     // [M+1] harmonics per region, nmk[M] = K (exterior)
};

// Row/column bookkeeping of eng.py:164-276 and the per-region table offsets.
struct Layout {
    int M, NT, p, K;
    int col[kMaxRegions + 1];      // first column of region i (col[M] = first exterior column = p)
    int pot_row[kMaxRegions], pot_h[kMaxRegions];
    int vel_row[kMaxRegions], vel_h[kMaxRegions];
    int tab[kMaxRegions + 1];      // offset (doubles) of region i's radial table; tab[M] = int_phi[]
};

MEEM_HDX void make_layout(const Geom& g, Layout& L) {
    const int M = g.M;
    L.M = M;
    int c = 0, t = 0;
    for (int i = 0; i < M; ++i) {
        L.col[i] = c;
        L.tab[i] = t;
        c += (i == 0 ? 1 : 2) * g.nmk[i];
        t += kTabVecs * g.nmk[i];
    }
    L.col[M] = c;
    L.tab[M] = t;
    L.K = g.nmk[M];
    L.p = c;
    L.NT = c + L.K;
    int r = 0;
    for (int bd = 0; bd < M; ++bd) {
        int hgt = (bd == M - 1) ? g.nmk[bd] : (g.d[bd] > g.d[bd + 1] ? g.nmk[bd] : g.nmk[bd + 1]);
        L.pot_row[bd] = r; L.pot_h[bd] = hgt; r += hgt;
    }
    for (int bd = 0; bd < M; ++bd) {
        int hgt = (bd == M - 1) ? g.nmk[bd + 1] : (g.d[bd] <= g.d[bd + 1] ? g.nmk[bd] : g.nmk[bd + 1]);
        L.vel_row[bd] = r; L.vel_h[bd] = hgt; r += hgt;
    }
}
MEEM_HDX int table_doubles(const Layout& L) { return L.tab[L.M] + L.M; }

// me.py:35-36 — (n*pi)/(h-d_i), in that order.
MEEM_HD double lambda_ni(int n, int i, const Geom& g) { return (double)n * kPi / (g.h - g.d[i]); }

// me.py:71-93
MEEM_HD double inm(int n, int m, int i, const Geom& g) {
    const double di = g.d[i], di1 = g.d[i + 1];
    const double dj = fmax(di, di1);
    const double L = g.h - dj;
    if (n == 0 && m == 0) return L;
    const double l1 = lambda_ni(n, i, g), l2 = lambda_ni(m, i + 1, g);
    if (n == 0) return (dj == di1) ? 0.0 : kSqrt2 * sin(l2 * L) / l2;
    if (m == 0) return (dj == di) ? 0.0 : kSqrt2 * sin(l1 * L) / l1;
    const double s = l1 + l2;
    double frac1 = sin(s * L) / s;
    double frac2;
    if (l1 == l2) frac2 = L;
    else { double df = l1 - l2; frac2 = sin(df * L) / df; }
    return frac1 + frac2;
}

// Per-frequency scalars of the exterior region.
struct Exterior {
    double m0, h, m0h;
    bool inf_m0, lowbranch;    // m0*h < 14 (me.py:103,111,169)
    double tail;               // exp(-m0 d) - exp(m0 d - 2 m0 h), d = d[M-1]
    double sinh_m0L;           // sinh(m0 (h - d[M-1]))
};

MEEM_HD void make_exterior(const Geom& g, double m0, Exterior& e) {
    const double dj = g.d[g.M - 1];
    e.m0 = m0; e.h = g.h; e.m0h = m0 * g.h;
    e.inf_m0 = isinf(m0);
    e.lowbranch = e.m0h < kM0HThresh;
    e.tail = exp(-m0 * dj) - exp(m0 * dj - 2.0 * m0 * g.h);
    e.sinh_m0L = sinh(m0 * (g.h - dj));
}

// me.py:96-124.  mk/Nk are this frequency's arrays (K entries).
MEEM_HD double imk(int m, int k, const Geom& g, const Exterior& e, const double* mk, const double* Nk) {
    const int i = g.M - 1;
    const double dj = g.d[i];
    const double L = g.h - dj;
    const double m0 = e.m0;
    if (k == 0) {
        if (e.inf_m0) return 0.0;
        if (m == 0) {
            if (e.lowbranch) return (1.0 / sqrt(Nk[0])) * e.sinh_m0L / m0;
            return sqrt(2.0 * g.h / m0) * e.tail;
        }
        const double sgn = (m & 1) ? -1.0 : 1.0;
        double num;
        if (e.lowbranch) num = sgn * kSqrt2 * (1.0 / sqrt(Nk[0])) * m0 * e.sinh_m0L;
        else num = sgn * 2.0 * sqrt(g.h * (m0 * m0 * m0)) * e.tail;
        const double lam = lambda_ni(m, i, g);
        return num / (m0 * m0 + lam * lam);
    }
    const double mkk = mk[k];
    if (m == 0) return (1.0 / sqrt(Nk[k])) * sin(mkk * L) / mkk;
    const double l1 = lambda_ni(m, i, g);
    if (fabs(mkk) == l1) return sqrt(2.0 / Nk[k]) * L / 2.0;
    const double sp = mkk + l1, sm = mkk - l1;
    double frac1 = sin(sp * L) / sp;
    double frac2 = sin(sm * L) / sm;
    return sqrt(2.0 / Nk[k]) * (frac1 + frac2) / 2.0;
}

// h * Lambda'_k(a_M) (me.py:352-374, :563-568): complex for k = 0, real otherwise.
MEEM_HD cplx ext_diag(int k, const Geom& g, const Exterior& e, const double* mk) {
    const double aM = g.a[g.M - 1];
    if (k == 0) {
        if (e.inf_m0) return cmake(g.h, 0.0);
        double j0, y0, j1, y1;
        bessel_jy01(e.m0 * aM, &j0, &y0, &j1, &y1);
        cplx q = cmake(-(e.m0 * j1), -(e.m0 * y1)) / cmake(j0, y0);
        return cmake(g.h * q.re, g.h * q.im);
    }
    const double m = mk[k];
    const double x = m * aM;
    double v = (-(m * k1e(x)) / k0e(x)) * exp(m * (aM - aM));
    return cmake(g.h * v, 0.0);
}

// arg H0^(1)(m0 a_M) for the excitation phase (me.py:467-471)
MEEM_HD double hankel0_angle(double x) {
    double j0, y0, j1, y1;
    bessel_jy01(x, &j0, &y0, &j1, &y1);
    return atan2(y0, j0);
}

// Radial table of region i, harmonic n (me.py:192-297, :423-448), written at tab[v*N + n]:
//  v=0 R1(a_i) 1 R2(a_i) 2 R1'(a_i) 3 R2'(a_i) 4 R1(a_{i-1}) 5 R2(a_{i-1}) 6 R1'(a_{i-1}) 7 R2'(a_{i-1})
//  v=8 int_R1 * z_n   9 int_R2 * z_n
MEEM_HD void radial_entry(const Geom& g, int i, int n, double* tab) {
    const int N = g.nmk[i];
    const double ai = g.a[i];
    const double ri = (i > 0) ? g.a[i - 1] : 0.0;
    double v[kTabVecs];
    if (n == 0) {
        v[0] = 0.5; v[1] = (i > 0) ? 0.5 * log(ai / ai) : 0.0;
        v[2] = 0.0; v[3] = 1.0 / (2.0 * ai);
        v[4] = 0.5; v[5] = (i > 0) ? 0.5 * log(ri / ai) : 0.0;
        v[6] = 0.0; v[7] = (i > 0) ? 1.0 / (2.0 * ri) : 0.0;
        v[8] = ai * ai / 4.0 - ri * ri / 4.0;
        v[9] = (i > 0) ? (ri * ri * (2.0 * log(ai / ri) + 1.0) - ai * ai) / 8.0 : 0.0;
    } else {
        const double lam = lambda_ni(n, i, g);
        const double zn = (n & 1) ? -kSqrt2 : kSqrt2;
        const double xo = lam * ai;
        const double i0o = i0e(xo), i1o = i1e(xo);
        v[0] = 1.0;
        v[2] = (lam * i1o) / i0o;                       // * exp(lam*(a_i-a_i)) = 1
        const double bottom1 = lam * i0o;
        double int1 = (ai * i1o / bottom1);
        if (i > 0) {
            const double k0o = k0e(xo), k1o = k1e(xo);
            const double xi = lam * ri;
            v[1] = 1.0;
            v[3] = (-lam * k1o) / k0o;
            v[4] = (i0e(xi) / i0o) * exp(lam * (ri - ai));
            v[5] = (k0e(xi) / k0o) * exp(lam * (ai - ri));
            v[6] = ((lam * i1e(xi)) / i0o) * exp(lam * (ri - ai));
            v[7] = ((-lam * k1e(xi)) / k0o) * exp(lam * (ai - ri));
            int1 = int1 - (ri * i1e(xi) / bottom1) * exp(lam * (ri - ai));
            const double bottom2 = -lam * k0o;
            v[9] = ((ai * k1o / bottom2) - (ri * k1e(xi) / bottom2) * exp(lam * (ai - ri))) * zn;
        } else {
            v[1] = v[3] = v[4] = v[5] = v[6] = v[7] = v[9] = 0.0;
        }
        v[8] = int1 * zn;
    }
    for (int q = 0; q < kTabVecs; ++q) tab[q * N + n] = v[q];
}

// me.py:451-457
MEEM_HD double int_phi_p(const Geom& g, int i) {
    const double L = g.h - g.d[i];
    const double ai = g.a[i];
    double num = ai * ai * (4.0 * (L * L) - ai * ai);
    if (i > 0) {
        const double ri = g.a[i - 1];
        num = num - ri * ri * (4.0 * (L * L) - ri * ri);
    }
    return num / (16.0 * L);
}

MEEM_HD void fill_tables(const Geom& g, const Layout& L, double* tab, int tid, int nthr) {
    for (int i = 0; i < g.M; ++i)
        for (int n = tid; n < g.nmk[i]; n += nthr) radial_entry(g, i, n, tab + L.tab[i]);
    for (int i = tid; i < g.M; i += nthr) tab[L.tab[g.M] + i] = int_phi_p(g, i);
}

// ---------------------------------------------------------------------------------- A fill
// omega-independent blocks (eng.py:164-276 minus the closures registered at :177-188, :229-246).
// Only entries with row < rlim and col < rlim are written (rlim = p for the Schur path's A11,
// N_T for the dense matrix).  The target must be zeroed beforehand.
// Rows of ONE boundary: potential rows at row rp, velocity rows at row rv, columns shifted by -cshift (the window
// elimination of k_geom_band places boundary bd's rows into a window whose column 0 is region bd's first column;
// the full matrix uses rp = L.pot_row[bd], rv = L.vel_row[bd], cshift = 0).  The last boundary writes only its
// omega-independent potential-row diagonal blocks.
template <class T>
MEEM_HD void fill_A_boundary_at(const Geom& g, const Layout& L, const double* tab, int bd, T* A, long ld, int rp, int rv,
                                int c1L, int c2L, int cR, int tid, int nthr) {
    // c1L / c2L: first column of C1^bd / C2^bd, cR: first column of region bd+1 (C1^{bd+1} followed by C2^{bd+1})
    const int M = g.M;
    const int N = g.nmk[bd];
    const double* TL = tab + L.tab[bd];
    const double hdL = g.h - g.d[bd];
    if (bd == M - 1) {
        for (int n = tid; n < N; n += nthr) {
            set_real(A[(long)(rp + n) * ld + c1L + n], hdL * TL[0 * N + n]);
            if (bd > 0) set_real(A[(long)(rp + n) * ld + c2L + n], hdL * TL[1 * N + n]);
        }
        return;
    }
    const int Mm = g.nmk[bd + 1];
    const double* TR = tab + L.tab[bd + 1];
    const double hdR = g.h - g.d[bd + 1];
    const bool pot_left_diag = g.d[bd] > g.d[bd + 1];
    // diagonal blocks
    if (pot_left_diag) {
        for (int n = tid; n < N; n += nthr) {
            set_real(A[(long)(rp + n) * ld + c1L + n], hdL * TL[0 * N + n]);
            if (bd > 0) set_real(A[(long)(rp + n) * ld + c2L + n], hdL * TL[1 * N + n]);
        }
        for (int m = tid; m < Mm; m += nthr) {       // velocity rows use region bd+1's modes
            set_real(A[(long)(rv + m) * ld + cR + m], hdR * TR[6 * Mm + m]);
            set_real(A[(long)(rv + m) * ld + cR + Mm + m], hdR * TR[7 * Mm + m]);
        }
    } else {
        for (int m = tid; m < Mm; m += nthr) {
            set_real(A[(long)(rp + m) * ld + cR + m], -(hdR * TR[4 * Mm + m]));
            set_real(A[(long)(rp + m) * ld + cR + Mm + m], -(hdR * TR[5 * Mm + m]));
        }
        for (int n = tid; n < N; n += nthr) {
            set_real(A[(long)(rv + n) * ld + c1L + n], -(hdL * TL[2 * N + n]));
            if (bd > 0) set_real(A[(long)(rv + n) * ld + c2L + n], -(hdL * TL[3 * N + n]));
        }
    }
    // dense blocks
    for (int e = tid; e < N * Mm; e += nthr) {
        const int n = e / Mm, m = e - n * Mm;
        const double I = inm(n, m, bd, g);
        if (pot_left_diag) {
            set_real(A[(long)(rp + n) * ld + cR + m], -(TR[4 * Mm + m] * I));
            set_real(A[(long)(rp + n) * ld + cR + Mm + m], -(TR[5 * Mm + m] * I));
            set_real(A[(long)(rv + m) * ld + c1L + n], -(TL[2 * N + n] * I));
            if (bd > 0) set_real(A[(long)(rv + m) * ld + c2L + n], -(TL[3 * N + n] * I));
        } else {
            set_real(A[(long)(rp + m) * ld + c1L + n], TL[0 * N + n] * I);
            if (bd > 0) set_real(A[(long)(rp + m) * ld + c2L + n], TL[1 * N + n] * I);
            set_real(A[(long)(rv + n) * ld + cR + m], TR[6 * Mm + m] * I);
            set_real(A[(long)(rv + n) * ld + cR + Mm + m], TR[7 * Mm + m] * I);
        }
    }
}

template <class T>
MEEM_HD void fill_A_boundary(const Geom& g, const Layout& L, const double* tab, int bd, T* A, long ld, int rp, int rv,
                             int cshift, int tid, int nthr) {
    const int c1 = L.col[bd] - cshift;
    fill_A_boundary_at<T>(g, L, tab, bd, A, ld, rp, rv, c1, c1 + g.nmk[bd], L.col[bd + 1] - cshift, tid, nthr);
}

template <class T>
MEEM_HD void fill_A_static(const Geom& g, const Layout& L, const double* tab, T* A, long ld, int tid, int nthr) {
    for (int bd = 0; bd < g.M; ++bd) fill_A_boundary<T>(g, L, tab, bd, A, ld, L.pot_row[bd], L.vel_row[bd], 0, tid, nthr);
}

// omega-dependent blocks of the last boundary (eng.py:177-188, :229-246; me.py:529-536, :563-576).
MEEM_HD void fill_A_omega(const Geom& g, const Layout& L, const double* tab, const Exterior& ex,
                          const double* mk, const double* Nk, cplx* A, long ld, int tid, int nthr) {
    const int bd = g.M - 1;
    const int N = g.nmk[bd], K = L.K;
    const double* TL = tab + L.tab[bd];
    const int c1L = L.col[bd], c2L = c1L + N, cE = L.col[g.M];
    const int rp = L.pot_row[bd], rv = L.vel_row[bd];
    for (int e = tid; e < N * K; e += nthr) {
        const int n = e / K, k = e - n * K;
        const double I = imk(n, k, g, ex, mk, Nk);
        set_real(A[(long)(rp + n) * ld + cE + k], -I);
        set_real(A[(long)(rv + k) * ld + c1L + n], -(TL[2 * N + n] * I));
        if (bd > 0) set_real(A[(long)(rv + k) * ld + c2L + n], -(TL[3 * N + n] * I));
    }
    for (int k = tid; k < K; k += nthr) A[(long)(rv + k) * ld + cE + k] = ext_diag(k, g, ex, mk);
}

// ---------------------------------------------------------------------------------- b fill
// omega-independent rows (me.py:129-159; eng.py:278-301; problem_cache.py:93-153).  hv[M+1] are the
// heaving flags of the regions for this radiating mode (hv[M] = 0 for the exterior).
MEEM_HD double b_static_entry(const Geom& g, const Layout& L, const unsigned char* hv, int bd, bool vel, int n) {
    const int M = g.M;
    const double h = g.h;
    const double hi = hv[bd] ? 1.0 : 0.0;
    if (!vel) {
        int j;
        double constant;
        if (bd == M - 1) {
            j = bd;
            constant = -hi / (h - g.d[bd]);
        } else {
            j = bd + (g.d[bd] <= g.d[bd + 1] ? 1 : 0);
            const double hi1 = hv[bd + 1] ? 1.0 : 0.0;
            constant = hi1 / (h - g.d[bd + 1]) - hi / (h - g.d[bd]);
        }
        const double Lj = h - g.d[j];
        if (n == 0) return constant * 1.0 / 2.0 * ((Lj * Lj * Lj) / 3.0 - Lj * (g.a[bd] * g.a[bd]) / 2.0);
        const double lam = lambda_ni(n, j, g);
        const double sgn = (n & 1) ? -1.0 : 1.0;
        return kSqrt2 * Lj * constant * sgn / (lam * lam);
    }
    // velocity rows, bd < M-1
    const double hi1 = hv[bd + 1] ? 1.0 : 0.0;
    const double ai = g.a[bd];
    if (n == 0) return (hi1 - hi) * (ai / 2.0);
    if (g.d[bd] > g.d[bd + 1]) {
        if (!hv[bd]) return 0.0;
        const double lam = lambda_ni(n, bd + 1, g);
        const double Li = h - g.d[bd];
        return (-kSqrt2 * ai * sin(lam * Li)) / (2.0 * Li * lam);
    }
    if (!hv[bd + 1]) return 0.0;
    const double lam = lambda_ni(n, bd, g);
    const double Li1 = h - g.d[bd + 1];
    return (kSqrt2 * ai * sin(lam * Li1)) / (2.0 * Li1 * lam);
}

// rows [0, p): every potential row group and the velocity row groups of bd < M-1
template <class T>
MEEM_HD void fill_b_static(const Geom& g, const Layout& L, const unsigned char* hv, T* b, long stride, int tid, int nthr) {
    for (int bd = 0; bd < g.M; ++bd) {
        for (int n = tid; n < L.pot_h[bd]; n += nthr)
            set_real(b[(long)(L.pot_row[bd] + n) * stride], b_static_entry(g, L, hv, bd, false, n));
        if (bd < g.M - 1)
            for (int n = tid; n < L.vel_h[bd]; n += nthr)
                set_real(b[(long)(L.vel_row[bd] + n) * stride], b_static_entry(g, L, hv, bd, true, n));
    }
}

// me.py:163-174: b_velocity_end_entry(k) = constant * I_mk(0, k)
MEEM_HD double b_end_constant(const Geom& g, const unsigned char* hv) {
    const int i = g.M - 1;
    return -(hv[i] ? 1.0 : 0.0) * g.a[i] / (2.0 * (g.h - g.d[i]));
}

// c vector (eng.py:65-98): entry for column `col` < p under flags hv
MEEM_HD double c_entry(const Geom& g, const Layout& L, const double* tab, const unsigned char* hv, int i, int which, int n) {
    if (!hv[i]) return 0.0;
    const int N = g.nmk[i];
    return tab[L.tab[i] + (8 + which) * N + n];
}

template <class T>
MEEM_HD void fill_c(const Geom& g, const Layout& L, const double* tab, const unsigned char* hv, T* c, int tid, int nthr) {
    for (int i = 0; i < g.M; ++i) {
        const int N = g.nmk[i];
        for (int n = tid; n < N; n += nthr) {
            set_real(c[L.col[i] + n], c_entry(g, L, tab, hv, i, 0, n));
            if (i > 0) set_real(c[L.col[i] + N + n], c_entry(g, L, tab, hv, i, 1, n));
        }
    }
}

MEEM_HD double phi_sum(const Geom& g, const Layout& L, const double* tab, const unsigned char* hv) {
    double s = 0.0;
    for (int i = 0; i < g.M; ++i)
        if (hv[i]) s += tab[L.tab[g.M] + i];
    return s;
}

// me.py:473-475
MEEM_HD double excitation_force(double damping, double m0, double h, double rho_ref) {
    const double T = tanh(m0 * h);
    const double c = T + m0 * h * (1.0 - T * T);
    return sqrt((2.0 * c * rho_ref * (kG * kG) * damping) / (omega_of_m0(m0, h) * m0));
}

}  // namespace meem
