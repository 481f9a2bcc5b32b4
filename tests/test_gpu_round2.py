"""Round-2 GPU parity tests: the window (block-bidiagonal) elimination of the geometry stage in all its variants, the
mode-count sweep up to 400 modes/region, the host-buffer C-ABI call meemb200_solve_host exactly as INTEGRATION.md's stub
makes it, the c vector as an output of the batch call, and tighter root / full-size checks."""
import ctypes as C

import numpy as np
import pytest

from oracle import meem_oracle as mo
from helpers import load_npz, relerr, phase_err

pytestmark = pytest.mark.gpu

Z1, META1 = load_npz("ref_single.npz")
Z2, META2 = load_npz("ref_sweeps.npz")


@pytest.fixture(scope="module")
def ofb():
    import openflash_b200 as ofb
    return ofb


def _check(out, ref, tol=1e-10):
    for var in ("added_mass", "damping", "excitation_force"):
        r = ref[var]
        ok = np.isfinite(r)
        assert np.array_equal(np.isfinite(out[var]), ok), var
        assert relerr(out[var][ok], r[ok]) <= tol, (var, relerr(out[var][ok], r[ok]))
    okp = np.isfinite(ref["excitation_phase"])
    assert np.array_equal(np.isfinite(out["excitation_phase"]), okp)
    assert phase_err(out["excitation_phase"][okp], ref["excitation_phase"][okp]) <= tol


BAND_ENVS = [{"MEEMB200_BAND_VARIANT": "0"}, {"MEEMB200_BAND_VARIANT": "1"}, {"MEEMB200_BAND_VARIANT": "2"}, {"MEEMB200_NO_BAND": "1"},
             {"MEEMB200_NO_COOP": "1"}]
_ids = lambda e: "-".join(f"{k[9:]}={v}" for k, v in e.items())  # noqa: E731


@pytest.mark.parametrize("env", BAND_ENVS, ids=_ids)
@pytest.mark.parametrize("name", sorted(META2))
def test_window_elimination_variants_vs_reference(ofb, monkeypatch, name, env):
    """k_geom_band (window in shared memory at 128 / 256 threads, window in the global slot) and the dense-LU kernel it
    replaces, on the golden sweeps of the untouched reference (coefficient-only path: X keeps the dense elimination)."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    s = META2[name]
    geo = ofb.FlatGeometry.from_body_map(s["h"], s["a"], s["d"], s["NMK"], s["body_map"], s["modes"])
    res = ofb.solve_batch([geo, geo], s["omegas"])
    ref = {v: Z2[f"{name}/{v}"] for v in ("added_mass", "damping", "excitation_force", "excitation_phase")}
    for r in res:
        assert np.all(r["status"] == 0)
        _check(r, ref)


@pytest.mark.parametrize("env", BAND_ENVS[:4], ids=_ids)
def test_window_elimination_heterogeneous_batch(ofb, monkeypatch, env):
    """C4 generator (mixed M = 2..8, 10 % overhanging depth orders) plus single cylinders, unequal mode counts per region and
    two radiating bodies in ONE batch: every window shape (first / interior / last, case 1 and case 2 row counts)."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    import bench
    cases = [(g["h"], g["a"], g["d"], g["NMK"], g["body_map"], g["modes"]) for g in bench.c4_geometries(40)]
    cases += [(3.0, [1.0], [0.7], [9, 11], [0], [0]),
              (1.001, [0.5, 1.0], [0.25, 0.5], [6, 7, 8], [0, 1], [0, 1]),
              (1.9, [0.3, 0.5, 1.0, 1.2, 1.6], [0.5, 0.7, 0.8, 0.2, 0.5], [8, 9, 10, 11, 12, 13], [0, 1, 2, 3, 4], [0, 2, 4]),
              (20.0, [5, 10, 15], [15, 10, 5], [30, 12, 25, 40], [0, 1, 1], [0, 1]),
              (20.0, [5, 10, 15], [15, 10, 5], [30, 12, 40, 9], [0, 0, 1], [1])]
    om = np.array([0.3, 1.1, 2.5])
    geos = [ofb.FlatGeometry.from_body_map(*c) for c in cases]
    res = ofb.solve_batch(geos, om)
    for c, r in zip(cases, res):
        ref = mo.run_sweep(mo.FlatGeometry(*c[:4]), om, c[4], c[5])
        assert np.all(r["status"] == 0)
        _check(r, ref)


def test_window_elimination_c4_bulk_vs_dense_kernel(ofb):
    """2 000 C4 geometries x 8 omega through the product path (k_geom_band + k_freq_small) against the independent
    reference-shaped complex LU on the device (k_dense): 16 000 systems, 1e-9 (two device paths, each 1e-10 of the oracle)."""
    import bench
    gs = bench.c4_geometries(2000)
    om = np.linspace(0.3, 2.5, 8)
    geos = [ofb.FlatGeometry.from_body_map(g["h"], g["a"], g["d"], g["NMK"], g["body_map"], g["modes"]) for g in gs]
    rs = ofb.solve_batch(geos, om)
    rd = ofb.solve_batch(geos, om, algo=ofb.ALGO_DENSE)
    worst = 0.0
    for a, b in zip(rs, rd):
        assert np.all(a["status"] == 0) and np.all(b["status"] == 0)
        for v in ("added_mass", "damping", "excitation_force"):
            worst = max(worst, relerr(a[v], b[v]))
    assert worst <= 1e-9, worst


@pytest.mark.parametrize("M", [12, 24, 48])
def test_many_region_staircase_window_elimination(ofb, M):
    """staircases of 12 / 24 / 48 regions (dev/python/slants/helpers.py:80-125), cone and overhanging
    variant, through the window elimination (O(M N^3): the working set is one 25 x 33 window whatever M)."""
    h, a_out, n = 10.0, 6.0, 8
    a = [a_out * (i + 1) / M for i in range(M)]
    d_cone = [6.0 - 5.0 * (i + 0.5) / M for i in range(M)]
    d_over = [x if i < 2 * M // 3 else d_cone[2 * M // 3 - 1] + 0.05 * (i - 2 * M // 3 + 1) for i, x in enumerate(d_cone)]
    om = np.array([0.6, 1.4, 2.2])
    for d in (d_cone, d_over):
        ref = mo.run_sweep(mo.FlatGeometry(h, a, d, [n] * (M + 1)), om, [0] * M, [0])
        geo = ofb.FlatGeometry.from_body_map(h, a, d, [n] * (M + 1), [0] * M, [0])
        res = ofb.solve_batch([geo], om)[0]
        assert np.all(res["status"] == 0)
        _check(res, ref)


@pytest.mark.parametrize("geom", ["small_bicylinder", "small_tricylinder"])
def test_mode_count_convergence_sweep_c5_full(ofb, geom):
    """BASELINE.json configs[4] with its full list of mode counts 10 .. 400 on both geometries.  Where the reference's
    exp() overflows (NaN coefficients) the device must answer NaN + status, never refuse the plan."""
    h, a, d, bm = {"small_bicylinder": (1.001, [0.5, 1.0], [0.5, 0.25], [0, 0]),
                   "small_tricylinder": (20.0, [5.0, 10.0, 15.0], [15.0, 10.0, 5.0], [0, 0, 0])}[geom]
    ns = [10, 15, 20, 30, 40, 50, 75, 100, 150, 200, 300, 400]
    om = np.array([0.5, 3.0])
    geos = [ofb.FlatGeometry.from_body_map(h, a, d, [n] * (len(a) + 1), bm, [0]) for n in ns]
    res = ofb.solve_batch(geos, om)
    for n, r in zip(ns, res):
        ref = mo.run_sweep(mo.FlatGeometry(h, a, d, [n] * (len(a) + 1)), om, bm, [0])
        _check(r, ref)
        ok = np.isfinite(ref["added_mass"]).reshape(len(om), -1).all(axis=1)
        assert np.array_equal(r["status"] == 0, ok), n


def test_big_tricylinder_full_size_16_omega(ofb):
    """C3 at full size against the ORACLE (not another device path) at 16 frequencies across the sweep, incl. both sides of
    the m0 h = 14 branch."""
    h, a, d, NMK = 25.0, [10, 15, 20], [20, 15, 10], [200] * 4
    om = np.linspace(0.1, 3.0, 16)
    ref = mo.run_sweep(mo.FlatGeometry(h, a, d, NMK), om, [0, 0, 0], [0])
    geo = ofb.FlatGeometry.from_body_map(h, a, d, NMK, [0, 0, 0], [0])
    for env_nocoop in (False, True):
        if env_nocoop:
            import os
            os.environ["MEEMB200_NO_COOP"] = "1"
        try:
            res = ofb.solve_batch([geo], om)[0]
        finally:
            if env_nocoop:
                del os.environ["MEEMB200_NO_COOP"]
        assert np.all(res["status"] == 0)
        _check(res, ref)


def test_roots_tight(ofb):
    """N_k to 1e-12 and the mode index <-> bracket check with the reference's own nextafter bracket
    (multi_equations.py:42-61: root k of x tan x = -m0h tanh m0h in [nextafter((k-1/2)pi), nextafter(k pi)])."""
    for h in (1.001, 20.0, 50.0):
        om = np.linspace(0.1, 3.0, 64)
        m0, mk, nk = ofb.device_roots(om, h, 50)
        k = np.arange(1, 50)
        lo = np.nextafter((k - 0.5) * np.pi, np.inf)
        hi = np.nextafter(k * np.pi, np.inf)
        for i in range(len(om)):
            ref_m0 = mo.wavenumber(om[i], h)
            ref_mk = mo.m_k_array(50, ref_m0, h)
            assert relerr(nk[i], mo.N_k_array(ref_m0, h, ref_mk)) <= 1e-12
            x = mk[i, 1:] * h
            # m_k = x / h is rounded once more, so allow one ulp of x around the closed bracket
            assert np.all(x >= lo * (1 - 2.3e-16)) and np.all(x <= hi * (1 + 2.3e-16))
            assert np.array_equal(np.floor(x / np.pi + 0.5).astype(int), k)


# ------------------------------------------------------------------------------------------------ solve_host
def _raw_solve_host(ofb, h, a, d, NMK, flags, freqs, want_x, null_optional=False):
    """INTEGRATION.md's reference-side stub, verbatim in spirit: raw ctypes, host numpy buffers, no torch."""
    from openflash_b200 import _lib
    lib = _lib.load()
    ctx = C.c_void_p()
    assert lib.meemb200_create(0, C.byref(ctx)) == 0
    M = len(a)
    flags = np.asarray(flags, dtype=np.uint8).reshape(-1, M)
    R = flags.shape[0]
    reg_off = np.array([0, M + 1], dtype=np.int32)
    hh = np.array([h], dtype=np.float64)
    aa = np.array(list(a) + [np.inf], dtype=np.float64)
    dd = np.array(list(d) + [0.0], dtype=np.float64)
    nmk = np.array(NMK, dtype=np.int32)
    n_modes = np.array([R], dtype=np.int32)
    flag_off = np.array([0, R * (M + 1)], dtype=np.int64)
    fl = np.zeros((R, M + 1), dtype=np.uint8)
    fl[:, :M] = flags
    fr = np.ascontiguousarray(np.asarray(freqs, dtype=np.float64))
    b = _lib.MeemBatch()
    b.n_geom, b.n_freq = 1, len(fr)
    b.reg_off, b.h, b.a, b.d, b.nmk = (x.ctypes.data for x in (reg_off, hh, aa, dd, nmk))
    b.n_modes, b.flag_off, b.heaving, b.freq = n_modes.ctypes.data, flag_off.ctypes.data, fl.ctypes.data, fr.ctypes.data
    b.freq_kind, b.algo, b.root_mode, b.want_x, b.rho = 0, 0, 0, int(want_x), 1023.0
    plan = C.c_void_p()
    rc = lib.meemb200_plan_create(ctx, C.byref(b), C.byref(plan))
    if rc != 0:
        return rc, None
    F, NT = len(fr), int(nmk[0] + nmk[-1] + 2 * nmk[1:-1].sum())
    out = dict(added_mass=np.full((F, R, R), -7.0), damping=np.full((F, R, R), -7.0), excitation_force=np.full((F, R), -7.0),
               excitation_phase=np.full((F, R), -7.0), m0=np.full(F, -7.0), status=np.full(F, -7, dtype=np.int32))
    if want_x:
        out["X"] = np.zeros((F, R, NT), dtype=np.complex128)
    o = _lib.MeemOutputs()
    o.added_mass, o.damping = out["added_mass"].ctypes.data, out["damping"].ctypes.data
    o.exc_force = None if null_optional else out["excitation_force"].ctypes.data
    o.exc_phase = None if null_optional else out["excitation_phase"].ctypes.data
    o.m0 = None if null_optional else out["m0"].ctypes.data
    o.status = out["status"].ctypes.data
    o.x = out["X"].ctypes.data if want_x else None
    rc = lib.meemb200_solve_host(plan, C.byref(o), None)
    lib.meemb200_plan_destroy(plan)
    return rc, out


@pytest.mark.parametrize("want_x", [False, True])
@pytest.mark.parametrize("name", ["tutorial_2body", "C2_small_tricylinder_R2"])
def test_solve_host_stub_vs_golden(ofb, name, want_x):
    """meemb200_solve_host driven exactly as the reference-side ctypes stub of INTEGRATION.md drives it (R = 2 sweeps
    of the untouched reference), with and without X."""
    if name not in META2:
        name = [k for k in sorted(META2) if len(META2[k]["modes"]) == 2][0]
    s = META2[name]
    bm = np.asarray(s["body_map"])
    flags = [[1 if bm[i] == m else 0 for i in range(len(bm))] for m in s["modes"]]
    rc, out = _raw_solve_host(ofb, s["h"], s["a"], s["d"], s["NMK"], flags, s["omegas"], want_x)
    assert rc == 0
    assert np.all(out["status"] == 0)
    ref = {v: Z2[f"{name}/{v}"] for v in ("added_mass", "damping", "excitation_force", "excitation_phase")}
    _check(out, ref)
    assert relerr(out["m0"], Z2[f"{name}/m0"]) <= 2e-14
    if want_x:
        pot = Z2[f"{name}/potentials"]
        NMK, col = s["NMK"], 0
        for dom in range(len(NMK)):
            width = NMK[dom] if dom in (0, len(NMK) - 1) else 2 * NMK[dom]
            refp = pot[:, :, dom, :width]
            assert np.max(np.abs(out["X"][:, :, col:col + width] - refp)) <= 1e-9 * np.max(np.abs(refp))
            col += width


def test_solve_host_single_mode_null_outputs_and_einval(ofb):
    """R = 1, optional host outputs NULL (only added mass / damping / status wanted), and the EINVAL paths of the stub:
    non-increasing radii (basic_region_geometry.py:29-30), non-positive frequency (meem_problem.py:31), NULL outputs."""
    from openflash_b200 import _lib
    lib = _lib.load()
    h, a, d, NMK = 1.001, [0.5, 1.0], [0.5, 0.25], [15, 15, 15]
    om = [0.7, 2.0]
    ref = mo.run_sweep(mo.FlatGeometry(h, a, d, NMK), om, [0, 0], [0])
    rc, out = _raw_solve_host(ofb, h, a, d, NMK, [[1, 1]], om, False, null_optional=True)
    assert rc == 0
    assert relerr(out["added_mass"], ref["added_mass"]) <= 1e-10 and relerr(out["damping"], ref["damping"]) <= 1e-10
    assert np.all(out["excitation_force"] == -7.0) and np.all(out["m0"] == -7.0)      # untouched
    rc, _ = _raw_solve_host(ofb, h, [1.0, 0.5], d, NMK, [[1, 1]], om, False)
    assert rc == -1 and b"radii" in lib.meemb200_last_error()
    rc, _ = _raw_solve_host(ofb, h, a, d, NMK, [[1, 1]], [0.0, 1.0], False)
    assert rc == -1 and b"positive" in lib.meemb200_last_error()
    assert lib.meemb200_solve_host(None, None, None) == -1
