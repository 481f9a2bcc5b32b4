"""Generate openflash_b200/csrc/bessel_coeffs.h: Chebyshev coefficients for the FP64 Bessel kernels.

    python tools/gen_bessel_coeffs.py

Each table is a Chebyshev interpolant (in a mapped variable t in [-1,1]) of a smooth auxiliary
function, computed with 50-digit mpmath and truncated where the tail drops below 2e-18 of the
leading coefficient.  The device code (meem_math.cuh) evaluates them with Clenshaw's recurrence.

  exponentially scaled modified Bessel (scipy.special.ive / kve equivalents)
    I0E_A : exp(-x) I0(x)            x in [0,8],   t = x/4 - 1
    I0E_B : sqrt(x) exp(-x) I0(x)    x in [8,inf), t = 16/x - 1
    I1E_A : exp(-x) I1(x) / x        x in [0,8],   t = x/4 - 1
    I1E_B : sqrt(x) exp(-x) I1(x)    x in [8,inf), t = 16/x - 1
    I0_S  : I0(x)    as f(y=x^2)     x in [0,2],   t = y/2 - 1
    I1_S  : I1(x)/x  as f(y)         x in [0,2]
    K0_S  : K0(x) + ln(x/2) I0(x)    as f(y), x in (0,2]
    K1_S  : x (K1(x) - ln(x/2) I1(x)) as f(y), x in (0,2]
    K0E_B : sqrt(x) exp(x) K0(x)     x in [2,inf), t = 4/x - 1
    K1E_B : sqrt(x) exp(x) K1(x)     x in [2,inf), t = 4/x - 1
  ordinary Bessel for the Hankel ratio H1/H0 and arg H0 (scipy.special.hankel1 equivalents)
    J0_S  : J0(x)    as f(y=x^2)     x in [0,8],   t = y/32 - 1
    J1_S  : J1(x)/x  as f(y)
    Y0_S  : Y0(x) - (2/pi) ln(x) J0(x)             as f(y)
    Y1_S  : (Y1(x) - (2/pi)(ln(x) J1(x) - 1/x))/x  as f(y)
    P0_L, Q0_L, P1_L, Q1_L : modulus/phase functions for x >= 8,
          P + iQ = sqrt(pi x/2) exp(-i(x - pi/4 - nu pi/2)) H_nu(x), stored as P(z), x*Q(z),
          z = 64/x^2, t = 2z - 1
"""
import os
import sys

import mpmath as mp

mp.mp.dps = 50
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "openflash_b200", "csrc", "bessel_coeffs.h")


def cheb_fit(f, n=96):
    """Chebyshev coefficients c_0..c_{n-1} of f on [-1,1] (first-kind nodes); f(t)=sum' c_k T_k(t)."""
    nodes = [mp.cos(mp.pi * (j + mp.mpf(1) / 2) / n) for j in range(n)]
    vals = [f(t) for t in nodes]
    coeffs = []
    for k in range(n):
        s = mp.mpf(0)
        for j in range(n):
            s += vals[j] * mp.cos(mp.pi * k * (j + mp.mpf(1) / 2) / n)
        coeffs.append(2 * s / n)
    coeffs[0] /= 2
    return coeffs


def truncate(c, tol=mp.mpf("2e-18")):
    scale = max(abs(x) for x in c[:4])
    last = 0
    for k, x in enumerate(c):
        if abs(x) > tol * scale:
            last = k
    return c[: last + 1]


def lim(f, x, eps=mp.mpf("1e-30")):
    return f(eps) if x == 0 else f(x)


TABLES = {}


def add(name, f, n=96):
    c = truncate(cheb_fit(f, n))
    if len(c) > n - 8:
        print(f"WARNING {name}: not converged with n={n} (kept {len(c)})", file=sys.stderr)
    TABLES[name] = c
    print(f"{name:8s} {len(c):3d} terms, tail {float(abs(c[-1])):.2e}")


def safe(f):
    """guard t -> x maps that hit x=0 or x=inf exactly."""
    return f


two_over_pi = 2 / mp.pi

# ---- I0e / I1e
add("I0E_A", lambda t: (lambda x: mp.exp(-x) * mp.besseli(0, x))(4 * (t + 1)))
add("I1E_A", lambda t: (lambda x: mp.exp(-x) * mp.besseli(1, x) / x if x != 0 else mp.mpf(1) / 2)(4 * (t + 1)))


def big_x(scale):
    def wrap(g):
        def f(t):
            u = t + 1
            if u < mp.mpf("1e-40"):
                u = mp.mpf("1e-40")
            return g(scale / u)
        return f
    return wrap


def i_large(nu):
    def g(x):
        if x > 5000:   # asymptotic expansion, avoids overflow and is exact to 50 digits here
            mu = 4 * nu * nu
            s, term = mp.mpf(1), mp.mpf(1)
            for k in range(1, 60):
                term *= -(mu - (2 * k - 1) ** 2) / (k * 8 * x)
                s += term
                if abs(term) < mp.mpf("1e-48"):
                    break
            return s / mp.sqrt(2 * mp.pi)
        return mp.sqrt(x) * mp.exp(-x) * mp.besseli(nu, x)
    return g


def k_large(nu):
    def g(x):
        if x > 5000:
            mu = 4 * nu * nu
            s, term = mp.mpf(1), mp.mpf(1)
            for k in range(1, 60):
                term *= (mu - (2 * k - 1) ** 2) / (k * 8 * x)
                s += term
                if abs(term) < mp.mpf("1e-48"):
                    break
            return s * mp.sqrt(mp.pi / 2)
        return mp.sqrt(x) * mp.exp(x) * mp.besselk(nu, x)
    return g


add("I0E_B", big_x(16)(i_large(0)))
add("I1E_B", big_x(16)(i_large(1)))

# ---- small-x K0/K1 (x <= 2), as functions of y = x^2 in [0,4], t = y/2 - 1
def of_y(g):
    def f(t):
        y = 2 * (t + 1)
        if y < mp.mpf("1e-60"):
            y = mp.mpf("1e-60")
        return g(mp.sqrt(y))
    return f


add("I0_S", of_y(lambda x: mp.besseli(0, x)))
add("I1_S", of_y(lambda x: mp.besseli(1, x) / x))
add("K0_S", of_y(lambda x: mp.besselk(0, x) + mp.log(x / 2) * mp.besseli(0, x)))
add("K1_S", of_y(lambda x: x * (mp.besselk(1, x) - mp.log(x / 2) * mp.besseli(1, x))))
add("K0E_B", big_x(4)(k_large(0)))
add("K1E_B", big_x(4)(k_large(1)))

# ---- J/Y small x (x <= 8), functions of y = x^2 in [0,64], t = y/32 - 1
def of_y64(g):
    def f(t):
        y = 32 * (t + 1)
        if y < mp.mpf("1e-60"):
            y = mp.mpf("1e-60")
        return g(mp.sqrt(y))
    return f


add("J0_S", of_y64(lambda x: mp.besselj(0, x)))
add("J1_S", of_y64(lambda x: mp.besselj(1, x) / x))
add("Y0_S", of_y64(lambda x: mp.bessely(0, x) - two_over_pi * mp.log(x) * mp.besselj(0, x)))
add("Y1_S", of_y64(lambda x: (mp.bessely(1, x) - two_over_pi * (mp.log(x) * mp.besselj(1, x) - 1 / x)) / x))


# ---- modulus/phase for x >= 8: z = 64/x^2 in (0,1], t = 2z - 1
def pq(nu, part):
    def f(t):
        z = (t + 1) / 2
        if z < mp.mpf("1e-12"):
            # asymptotic series limit
            z = mp.mpf("1e-12")
        x = 8 / mp.sqrt(z)
        if x > 2000:
            mu = 4 * nu * nu
            P, Q = mp.mpf(1), mp.mpf(0)
            term = mp.mpf(1)
            for k in range(1, 80):
                term *= (mu - (2 * k - 1) ** 2) / (k * 8 * x)
                if k % 2 == 1:
                    Q += term * (1 if (k // 2) % 2 == 0 else -1)
                else:
                    P += term * (-1 if (k // 2) % 2 == 1 else 1)
                if abs(term) < mp.mpf("1e-48"):
                    break
            val = P if part == "P" else x * Q
            return val
        H = mp.hankel1(nu, x)
        w = mp.sqrt(mp.pi * x / 2) * mp.exp(-1j * (x - mp.pi / 4 - nu * mp.pi / 2)) * H
        return mp.re(w) if part == "P" else x * mp.im(w)
    return f


add("P0_L", pq(0, "P"), 128)
add("Q0_L", pq(0, "Q"), 128)
add("P1_L", pq(1, "P"), 128)
add("Q1_L", pq(1, "Q"), 128)

with open(OUT, "w") as fh:
    fh.write("// GENERATED by tools/gen_bessel_coeffs.py (50-digit mpmath Chebyshev fits) - do not edit.\n")
    fh.write("#pragma once\n\n")
    for name, c in TABLES.items():
        fh.write(f"#define MEEM_N_{name} {len(c)}\n")
        fh.write(f"MEEM_CONST_TABLE double MEEM_{name}[{len(c)}] = {{\n")
        for x in c:
            fh.write("    " + mp.nstr(x, 20, min_fixed=0, max_fixed=0) + ",\n")
        fh.write("};\n\n")
print("wrote", os.path.normpath(OUT))
