"""CPU unit tests of the device math / closed-form headers (openflash_b200/csrc/*.cuh) compiled
as plain C++ with g++ (tests/host_math/host_math.cpp) and compared with the oracle and mpmath.
The same headers run as CUDA device code in the product; this pins their formulas without a GPU."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from oracle import meem_oracle as mo
from helpers import load_npz, relerr

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "host_math", "host_math.cpp")
LIB = os.path.join(HERE, "host_math", "libhost_math.so")
dp = ctypes.POINTER(ctypes.c_double)
ip = ctypes.POINTER(ctypes.c_int)
up = ctypes.POINTER(ctypes.c_ubyte)


@pytest.fixture(scope="module")
def lib():
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-o", LIB, SRC])
    L = ctypes.CDLL(LIB)
    L.hm_mk.argtypes = [ctypes.c_int, ctypes.c_double, ctypes.c_double, dp, dp, ctypes.c_int]
    L.hm_assemble.argtypes = [ctypes.c_int, ctypes.c_double, dp, dp, ip, up, ctypes.c_double, dp, dp, dp, dp, dp, dp]
    return L


def _bes(lib, kind, x):
    x = np.ascontiguousarray(x, float)
    out = np.empty_like(x)
    lib.hm_bessel(kind, len(x), x.ctypes.data_as(dp), out.ctypes.data_as(dp))
    return out


def test_scaled_bessel_vs_scipy(lib):
    from scipy.special import ive, kve, hankel1
    rng = np.random.default_rng(0)
    x = np.exp(rng.uniform(np.log(1e-3), np.log(3000), 4000))
    x = np.concatenate([x, [2.0, 8.0, np.nextafter(2.0, 0), np.nextafter(2.0, 3), np.nextafter(8.0, 0), np.nextafter(8.0, 9)]])
    assert relerr(_bes(lib, 0, x), ive(0, x)) < 3e-15
    assert relerr(_bes(lib, 1, x), ive(1, x)) < 3e-15
    assert relerr(_bes(lib, 2, x), kve(0, x)) < 6e-15      # AMOS kve itself is off by up to 21 ulp
    assert relerr(_bes(lib, 3, x), kve(1, x)) < 3e-15
    xs = np.exp(rng.uniform(np.log(1e-4), np.log(500), 4000))
    H0 = _bes(lib, 4, xs) + 1j * _bes(lib, 5, xs)
    H1 = _bes(lib, 6, xs) + 1j * _bes(lib, 7, xs)
    assert np.max(np.abs(H0 - hankel1(0, xs)) / np.abs(H0)) < 3e-15
    assert np.max(np.abs(H1 - hankel1(1, xs)) / np.abs(H1)) < 3e-15
    assert np.max(np.abs(H1 / H0 - hankel1(1, xs) / hankel1(0, xs)) / np.abs(H1 / H0)) < 4e-15


def test_scaled_bessel_vs_mpmath(lib):
    mp = pytest.importorskip("mpmath")
    mp.mp.dps = 40
    rng = np.random.default_rng(1)
    x = np.exp(rng.uniform(np.log(1e-3), np.log(3000), 300))
    fns = [lambda t: mp.exp(-t) * mp.besseli(0, t), lambda t: mp.exp(-t) * mp.besseli(1, t),
           lambda t: mp.exp(t) * mp.besselk(0, t), lambda t: mp.exp(t) * mp.besselk(1, t)]
    for kind, fn in enumerate(fns):
        got = _bes(lib, kind, x)
        worst = max(float(abs((mp.mpf(float(g)) - fn(mp.mpf(float(xi)))) / fn(mp.mpf(float(xi))))) for xi, g in zip(x, got))
        assert worst < 8 * 2.0 ** -52, (kind, worst)


def test_roots_replay_reference(lib):
    """mode 0 replays scipy's newton/brentq iterations (meem_math.cuh); mode 1 converges fully."""
    for h in [0.5, 1.001, 20, 25, 50]:
        om = np.concatenate([np.linspace(0.05, 3, 60), [5, 10]])
        hh = np.full_like(om, h)
        ref = np.array([mo.wavenumber(o, h) for o in om])
        for mode, tol in ((0, 1e-15), (1, 5e-12)):
            out = np.empty_like(om)
            lib.hm_wavenumber(len(om), om.ctypes.data_as(dp), hh.ctypes.data_as(dp), out.ctypes.data_as(dp), mode)
            assert relerr(out, ref) <= tol, (h, mode)
    for h in [1.001, 20, 50]:
        for m0 in [1e-6, 0.05, 1.0, 5.0, 1e3]:
            K = 120
            ref = mo.m_k_array(K, m0, h)
            nref = mo.N_k_array(m0, h, ref)
            for mode, tol in ((0, 1e-15), (1, 5e-12)):
                mk, nk = np.empty(K), np.empty(K)
                lib.hm_mk(K, m0, h, mk.ctypes.data_as(dp), nk.ctypes.data_as(dp), mode)
                assert relerr(mk, ref) <= tol
                # mode indexing: root k lies in ((k-1/2)pi, k pi]/h
                x = mk[1:] * h
                k = np.arange(1, K)
                assert np.all(x > (k - 0.5) * np.pi) and np.all(x <= k * np.pi * (1 + 5e-16))
                if m0 * h < 300:
                    assert relerr(nk, nref) <= 1e-9 if mode else relerr(nk, nref) <= 1e-13


Z1, META1 = load_npz("ref_single.npz")


@pytest.mark.parametrize("name", sorted(META1))
def test_assembly_matches_oracle_and_reference(lib, name):
    c = META1[name]
    M = len(c["a"])
    a = np.array(list(c["a"]) + [np.inf], float)
    d = np.array(list(c["d"]) + [0.0], float)
    nmk = np.array(c["NMK"], np.int32)
    geo = mo.FlatGeometry(c["h"], c["a"], c["d"], c["NMK"])
    NT = geo.size
    nb = max(c["body_map"]) + 1
    flags_all = mo.region_flags_for_modes(c["body_map"], list(range(nb)))
    for body in range(nb):
        hv = np.array(list(flags_all[body]) + [0], np.uint8)
        for mi, m0 in enumerate(list(c["m0"]) + [np.inf]):
            mk = mo.m_k_array(c["NMK"][-1], m0, c["h"])
            Nk = mo.N_k_array(m0, c["h"], mk)
            A = np.zeros((NT, NT), complex)
            b = np.zeros(NT, complex)
            cv = np.zeros(NT - c["NMK"][-1], complex)
            phi = ctypes.c_double()
            nt = lib.hm_assemble(M, c["h"], a.ctypes.data_as(dp), d.ctypes.data_as(dp), nmk.ctypes.data_as(ip),
                                 hv.ctypes.data_as(up), m0, mk.ctypes.data_as(dp), Nk.ctypes.data_as(dp),
                                 A.ctypes.data_as(dp), b.ctypes.data_as(dp), cv.ctypes.data_as(dp), ctypes.byref(phi))
            assert nt == NT
            with np.errstate(all="ignore"):
                Ao = mo.assemble_A(geo, m0, mk, Nk)
                bo = mo.assemble_b(geo, flags_all[body], m0, mk, Nk)
            assert np.array_equal(A != 0, Ao != 0)
            assert relerr(A, Ao) <= 2e-13
            assert relerr(b, bo) <= 2e-13
            assert relerr(cv, mo.c_vector(geo, flags_all[body])) <= 1e-13
            assert abs(phi.value - sum(f * mo.int_phi_p_i(i, geo) for i, f in enumerate(flags_all[body]))) <= 1e-15 * max(1, abs(phi.value))
            if np.isfinite(m0):
                key = f"{name}/b{body}/m{mi}"
                assert relerr(b, Z1[key + "/b"]) <= 2e-13           # the untouched reference's b
                if body == 0 and key + "/A" in Z1:
                    assert relerr(A, Z1[key + "/A"]) <= 2e-13       # ... and A
