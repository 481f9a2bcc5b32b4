// FP64 special functions and dispersion roots for the MEEM hot path.
//
// Replaces the third-party numerics the reference calls:
//   scipy.special.ive / kve      -> meem::i0e, i1e, k0e, k1e      (multi_equations.py:213,237,273,293)
//   scipy.special.hankel1        -> meem::bessel_jy01             (multi_equations.py:361-362, :471)
//   scipy.optimize.root_scalar   -> meem::wavenumber, mk_root     (multi_equations.py:26-30, :42-61)
//
// The same source compiles as device code (nvcc) and as plain C++ (g++, for the CPU
// unit tests in tests/host_math/); MEEM_HD selects the qualifier.
#pragma once
#include <math.h>

#ifdef __CUDACC__
#define MEEM_HD __device__ __forceinline__
#define MEEM_HDX __host__ __device__ inline      // table-free helpers the host-side planner shares
#define MEEM_CONST_TABLE static __constant__ const
#else
#define MEEM_HD static inline
#define MEEM_HDX static inline
#define MEEM_CONST_TABLE static const
#endif

#include "bessel_coeffs.h"

namespace meem {

constexpr double kPi = 3.141592653589793238462643383279502884;
constexpr double kG = 9.81;            // multi_constants.py:4
constexpr double kRhoDefault = 1023.0; // multi_constants.py:5
constexpr double kM0HThresh = 14.0;    // multi_equations.py:18
constexpr double kSqrt2 = 1.41421356237309504880168872420969808;
constexpr double kTwoOverPi = 0.636619772367581343075535053490057448;

// f(t) = c0 + sum_{k>=1} c_k T_k(t), Clenshaw recurrence.
MEEM_HD double cheb_eval(const double* c, int n, double t) {
    double b1 = 0.0, b2 = 0.0;
    const double t2 = 2.0 * t;
    for (int k = n - 1; k >= 1; --k) {
        double b0 = fma(t2, b1, c[k]) - b2;
        b2 = b1;
        b1 = b0;
    }
    return fma(t, b1, c[0]) - b2;
}

// exp(-x) I0(x), x >= 0
MEEM_HD double i0e(double x) {
    x = fabs(x);
    if (x <= 8.0) return cheb_eval(MEEM_I0E_A, MEEM_N_I0E_A, 0.25 * x - 1.0);
    return cheb_eval(MEEM_I0E_B, MEEM_N_I0E_B, 16.0 / x - 1.0) / sqrt(x);
}

// exp(-x) I1(x), x >= 0
MEEM_HD double i1e(double x) {
    double ax = fabs(x), r;
    if (ax <= 8.0) r = ax * cheb_eval(MEEM_I1E_A, MEEM_N_I1E_A, 0.25 * ax - 1.0);
    else r = cheb_eval(MEEM_I1E_B, MEEM_N_I1E_B, 16.0 / ax - 1.0) / sqrt(ax);
    return x < 0 ? -r : r;
}

// exp(x) K0(x), x > 0  (x = 0 -> +inf)
MEEM_HD double k0e(double x) {
    if (x <= 2.0) {
        double t = 0.5 * x * x - 1.0;
        double k0 = cheb_eval(MEEM_K0_S, MEEM_N_K0_S, t) - log(0.5 * x) * cheb_eval(MEEM_I0_S, MEEM_N_I0_S, t);
        return k0 * exp(x);
    }
    return cheb_eval(MEEM_K0E_B, MEEM_N_K0E_B, 4.0 / x - 1.0) / sqrt(x);
}

// exp(x) K1(x), x > 0
MEEM_HD double k1e(double x) {
    if (x <= 2.0) {
        double t = 0.5 * x * x - 1.0;
        double k1 = cheb_eval(MEEM_K1_S, MEEM_N_K1_S, t) / x + log(0.5 * x) * x * cheb_eval(MEEM_I1_S, MEEM_N_I1_S, t);
        return k1 * exp(x);
    }
    return cheb_eval(MEEM_K1E_B, MEEM_N_K1E_B, 4.0 / x - 1.0) / sqrt(x);
}

// J0, Y0, J1, Y1 at x > 0 (absolute accuracy ~1e-16 relative to |H_nu(x)|).
MEEM_HD void bessel_jy01(double x, double* j0, double* y0, double* j1, double* y1) {
    if (x <= 8.0) {
        double t = x * x * (1.0 / 32.0) - 1.0;
        double lx = log(x);
        double J0 = cheb_eval(MEEM_J0_S, MEEM_N_J0_S, t);
        double J1 = x * cheb_eval(MEEM_J1_S, MEEM_N_J1_S, t);
        *j0 = J0;
        *j1 = J1;
        *y0 = kTwoOverPi * lx * J0 + cheb_eval(MEEM_Y0_S, MEEM_N_Y0_S, t);
        *y1 = kTwoOverPi * (lx * J1 - 1.0 / x) + x * cheb_eval(MEEM_Y1_S, MEEM_N_Y1_S, t);
        return;
    }
    double z = 64.0 / (x * x);
    double t = 2.0 * z - 1.0;
    double P0 = cheb_eval(MEEM_P0_L, MEEM_N_P0_L, t);
    double Q0 = cheb_eval(MEEM_Q0_L, MEEM_N_Q0_L, t) / x;
    double P1 = cheb_eval(MEEM_P1_L, MEEM_N_P1_L, t);
    double Q1 = cheb_eval(MEEM_Q1_L, MEEM_N_Q1_L, t) / x;
    double s = sin(x), c = cos(x);
    const double r2 = 0.70710678118654752440084436210484904;
    double c0 = (c + s) * r2, s0 = (s - c) * r2;     // cos/sin(x - pi/4)
    double c1 = s0, s1 = -c0;                        // cos/sin(x - 3pi/4)
    double amp = sqrt(kTwoOverPi / x);
    *j0 = amp * (P0 * c0 - Q0 * s0);
    *y0 = amp * (P0 * s0 + Q0 * c0);
    *j1 = amp * (P1 * c1 - Q1 * s1);
    *y1 = amp * (P1 * s1 + Q1 * c1);
}

// ------------------------------------------------------------------ dispersion relation
// multi_equations.py:20-24
MEEM_HD double omega_of_m0(double m0, double h) {
    if (isinf(m0)) return m0;
    return sqrt(m0 * tanh(m0 * h) * kG);
}

// Root of m tanh(h m) = omega^2/g (multi_equations.py:26-30).  The reference runs scipy's
// Newton with a finite-difference slope and stops at |dx| <= 1.48e-8, leaving up to ~2e-12
// relative error; this is a bracket-safeguarded Newton run to convergence.
MEEM_HD double wavenumber(double omega, double h) {
    if (isinf(omega)) return omega;
    const double c0 = omega * omega / kG;
    if (!(c0 > 0.0)) return c0 == 0.0 ? 0.0 : NAN;
    double lo = c0;                                                  // tanh <= 1
    double ch = c0 * h;
    double hi = (ch + sqrt(ch * ch + 4.0 * ch)) / (2.0 * h);         // tanh(x) >= x/(1+x)
    double m = (ch > 1.0) ? c0 : sqrt(c0 / h);
    if (m < lo) m = lo;
    if (m > hi) m = hi;
    for (int it = 0; it < 60; ++it) {
        double T = tanh(h * m);
        double f = m * T - c0;
        if (f > 0.0) hi = m; else lo = m;
        double fp = T + h * m * (1.0 - T * T);
        double mn = m - f / fp;
        if (!(mn > lo && mn < hi)) mn = 0.5 * (lo + hi);
        double dm = fabs(mn - m);
        m = mn;
        if (dm <= 4.0e-16 * m) break;
    }
    return m;
}

// k-th root (k >= 1) of x tan x + c = 0 in ((k-1/2)pi, k pi), c = m0 h tanh(m0 h) >= 0
// (multi_equations.py:42-61; the reference uses brentq with xtol 2e-12).  Solved in the pole-free
// form G(phi) = x cos(phi) - c sin(phi), x = (k-1/2)pi + phi, phi in (0, pi/2].
MEEM_HD double mk_root_x(int k, double c) {
    const double x0 = (k - 0.5) * kPi;
    if (!(c > 0.0)) return (c == 0.0) ? k * kPi : NAN;
    if (isinf(c)) return x0;
    double lo = 0.0, hi = 0.5 * kPi;
    double phi = atan2(x0 + 0.5, c);            // contraction phi <- atan((x0+phi)/c)
    phi = atan2(x0 + phi, c);
    phi = atan2(x0 + phi, c);
    for (int it = 0; it < 60; ++it) {
        double s = sin(phi), co = cos(phi);
        double x = x0 + phi;
        double G = x * co - c * s;
        if (G > 0.0) lo = phi; else hi = phi;
        double Gp = (1.0 - c) * co - x * s;
        double pn = phi - G / Gp;
        if (!(pn > lo && pn < hi)) pn = 0.5 * (lo + hi);
        double dp = fabs(pn - phi);
        phi = pn;
        if (dp <= 2.0e-16 * (x0 + phi)) break;
    }
    return x0 + phi;
}


// ------------------------------------------------------------------ reference-emulating roots
// The reference's roots are not converged to the last bit, so the closest parity
// comes from replaying its iterations: scipy's Newton with a 2-point finite-difference slope
// (root_scalar(method="newton", x0=2): step sqrt(eps)*max(1,|x|), stop when |dx| <= 1.48e-8,
// 50 iterations) and scipy's brentq (xtol 2e-12, rtol 4 eps, 100 iterations) on the bracket
// [nextafter((k-1/2)pi), nextafter(k pi)].  Device libm differs from glibc in the last ulp, so the
// replay agrees to ~1e-15 relative rather than bit-for-bit.
MEEM_HD double wavenumber_ref(double omega, double h) {
    if (isinf(omega)) return omega;
    const double c0 = omega * omega / kG;
    const double rstep = 1.4901161193847656e-08;   // sqrt(DBL_EPSILON)
    double p0 = 2.0;
    for (int it = 0; it < 50; ++it) {
        double fval = p0 * tanh(h * p0) - c0;
        if (fval == 0.0) return p0;
        double hs = rstep * (p0 >= 0.0 ? 1.0 : -1.0) * fmax(1.0, fabs(p0));
        double x1 = p0 + hs;
        double dx = x1 - p0;
        double fder = ((x1 * tanh(h * x1) - c0) - fval) / dx;
        if (fder == 0.0) return p0;
        double p = p0 - fval / fder;
        if (fabs(p - p0) <= 1.48e-8) return p;
        p0 = p;
    }
    return NAN;   // scipy raises "failed to converge"; the engine turns that into NaN rows
}

MEEM_HD double mk_ftan(double x, double c) { return x * tan(x) + c; }

MEEM_HD double mk_root_x_ref(int k, double c) {
    const double xtol = 2e-12, rtol = 8.881784197001252e-16;
    double xpre = nextafter(kPi * (k - 0.5), HUGE_VAL), xcur = nextafter(kPi * (double)k, HUGE_VAL);
    double xblk = 0.0, fblk = 0.0, spre = 0.0, scur = 0.0;
    double fpre = mk_ftan(xpre, c), fcur = mk_ftan(xcur, c);
    if (fpre == 0.0) return xpre;
    if (fcur == 0.0) return xcur;
    if (signbit(fpre) == signbit(fcur)) return NAN;
    for (int i = 0; i < 100; ++i) {
        if (fpre != 0.0 && fcur != 0.0 && (signbit(fpre) != signbit(fcur))) {
            xblk = xpre; fblk = fpre;
            spre = scur = xcur - xpre;
        }
        if (fabs(fblk) < fabs(fcur)) {
            xpre = xcur; xcur = xblk; xblk = xpre;
            fpre = fcur; fcur = fblk; fblk = fpre;
        }
        double delta = (xtol + rtol * fabs(xcur)) / 2.0;
        double sbis = (xblk - xcur) / 2.0;
        if (fcur == 0.0 || fabs(sbis) < delta) return xcur;
        if (fabs(spre) > delta && fabs(fcur) < fabs(fpre)) {
            double stry;
            if (xpre == xblk) {
                stry = -fcur * (xcur - xpre) / (fcur - fpre);
            } else {
                double dpre = (fpre - fcur) / (xpre - xcur);
                double dblk = (fblk - fcur) / (xblk - xcur);
                stry = -fcur * (fblk * dblk - fpre * dpre) / (dblk * dpre * (fblk - fpre));
            }
            if (2.0 * fabs(stry) < fmin(fabs(spre), 3.0 * fabs(sbis) - delta)) {
                spre = scur; scur = stry;
            } else {
                spre = sbis; scur = sbis;
            }
        } else {
            spre = sbis; scur = sbis;
        }
        xpre = xcur; fpre = fcur;
        if (fabs(scur) > delta) xcur += scur;
        else xcur += (sbis > 0.0 ? delta : -delta);
        fcur = mk_ftan(xcur, c);
    }
    return NAN;
}

enum RootMode { kRootsReference = 0, kRootsConverged = 1 };

MEEM_HD double wavenumber_mode(double omega, double h, int mode) {
    return mode == kRootsConverged ? wavenumber(omega, h) : wavenumber_ref(omega, h);
}

// m_k for k >= 1 (k = 0 is m0 itself), multi_equations.py:42-61
MEEM_HD double mk_entry(int k, double m0, double h, int mode) {
    if (k == 0) return m0;
    if (isinf(m0)) return ((k - 0.5) * kPi) / h;
    double c = m0 * h * tanh(m0 * h);
    return (mode == kRootsConverged ? mk_root_x(k, c) : mk_root_x_ref(k, c)) / h;
}

// multi_equations.py:379-384
MEEM_HD double nk_entry(int k, double m0, double h, double mk) {
    if (isinf(m0)) return 0.5;
    if (k == 0) return 0.5 * (1.0 + sinh(2.0 * m0 * h) / (2.0 * m0 * h));
    return 0.5 * (1.0 + sin(2.0 * mk * h) / (2.0 * mk * h));
}

}  // namespace meem
