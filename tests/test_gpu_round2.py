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
             {"MEEMB200_NO_COOP": "1"}, {"MEEMB200_NO_BAND_STATIC": "1"}, {"MEEMB200_NO_BAND_STATIC": "1", "MEEMB200_BAND_VARIANT": "2"}]
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


@pytest.mark.parametrize("env", BAND_ENVS[:4] + BAND_ENVS[5:], ids=_ids)
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


# ------------------------------------------------------------------------------------------------ c vector, reduction
@pytest.mark.parametrize("name", sorted(META2))
def test_c_vector_output_values_vs_reference(ofb, name):
    """MeemOutputs.c_vector (one [R][p] block per geometry out of the batch call) against the c_vector variable of the
    untouched reference's run_and_store_results (frequency, mode, p), <= 1e-12."""
    key = f"{name}/c_vector"
    if key not in Z2:
        pytest.skip("fixture without c_vector")
    s = META2[name]
    geo = ofb.FlatGeometry.from_body_map(s["h"], s["a"], s["d"], s["NMK"], s["body_map"], s["modes"])
    res = ofb.BatchPlan([geo, geo], s["omegas"], want_c=True).run()
    cref = Z2[key]
    assert np.all(cref.imag == 0)
    for r in res:
        assert r["c_vector"].shape == cref.shape[1:]
        for f in range(cref.shape[0]):
            assert relerr(r["c_vector"], cref[f].real) <= 1e-12


def test_engine_dataset_c_vector_and_assemble_c_values(ofb):
    """MEEMEngine.run_and_store_results -> dataset['c_vector'] values (not only shapes), assemble_c_multi values, and
    compute_hydrodynamic_coefficients for a given X through the device reduction (k_reduce)."""
    name = "tutorial_2body"
    s = META2[name]
    geo = ofb.BasicRegionGeometry.from_vectors(a=np.asarray(s["a"]), d=np.asarray(s["d"]), h=s["h"], NMK=list(s["NMK"]),
                                               body_map=list(s["body_map"]), heaving_map=[False] * (max(s["body_map"]) + 1))
    prob = ofb.MEEMProblem(geo)
    for bi, body in enumerate(prob.geometry.body_arrangement.bodies):
        body.heaving = bi in s["modes"]
    prob.set_frequencies(np.asarray(s["omegas"], float))
    eng = ofb.MEEMEngine([prob])
    ds = eng.run_and_store_results(0).dataset
    cref = Z2[f"{name}/c_vector"]
    got = np.asarray(ds["c_vector"].values)
    assert got.shape == cref.shape and relerr(got.real, cref.real) <= 1e-12 and np.all(got.imag == 0)
    # assemble_c_multi for region flags [1, 0]: the c vector of mode 0
    c0 = eng.assemble_c_multi(prob, [1, 0])
    assert relerr(c0.real, cref[0, 0].real) <= 1e-12
    # reduction of the reference's own X (golden single-solve fixtures)
    c = META1["small_bicylinder_2body"]
    geo1 = ofb.BasicRegionGeometry.from_vectors(a=np.asarray(c["a"]), d=np.asarray(c["d"]), h=c["h"], NMK=list(c["NMK"]),
                                                body_map=list(c["body_map"]), heaving_map=[True, False])
    p1 = ofb.MEEMProblem(geo1)
    e1 = ofb.MEEMEngine([p1])
    for mi, m0 in enumerate(c["m0"]):
        key = f"small_bicylinder_2body/b0/m{mi}"
        out = e1.compute_hydrodynamic_coefficients(p1, Z1[key + "/X"], m0)
        assert relerr([o["real"] for o in out], Z1[key + "/real"]) <= 1e-11
        assert relerr([o["imag"] for o in out], Z1[key + "/imag"]) <= 1e-11
        assert phase_err([o["excitation_phase"] for o in out], Z1[key + "/phase"]) <= 1e-11
        fr = Z1[key + "/force"]
        okf = np.isfinite(fr)
        assert relerr(np.array([o["excitation_force"] for o in out])[okf], fr[okf]) <= 1e-11
        if key + "/c" in Z1:
            assert relerr(np.array([o["c_vector"] for o in out]).real, Z1[key + "/c"].real) <= 1e-12


def test_run_batch_many_geometries_is_one_plan(ofb):
    """run_batch over 300 problems: c_vector comes out of the batch call (no per-geometry plan), values equal to the
    single-problem run, and the whole call stays within a small multiple of the bare BatchPlan.run()."""
    import time
    import bench
    gs = bench.c4_geometries(300)
    om = np.linspace(0.3, 2.5, 8)
    probs = []
    for g in gs:
        geo = ofb.BasicRegionGeometry.from_vectors(a=np.asarray(g["a"]), d=np.asarray(g["d"]), h=g["h"], NMK=list(g["NMK"]),
                                                   body_map=list(g["body_map"]), heaving_map=[True])
        p = ofb.MEEMProblem(geo)
        p.set_frequencies(om)
        probs.append(p)
    eng = ofb.MEEMEngine(probs)
    eng.run_batch(probs[:4], want_potentials=False)            # warm-up (context, kernels)
    t0 = time.perf_counter()
    out = eng.run_batch(probs, want_potentials=False)
    t_engine = time.perf_counter() - t0
    flats = [ofb.flatten_problem(p, modes=list(p.modes)) for p in probs]
    t0 = time.perf_counter()
    raw = ofb.BatchPlan(flats, om, want_c=True).run()
    t_plan = time.perf_counter() - t0
    for r, q in zip(out, raw):
        assert np.array_equal(np.asarray(r.dataset["added_mass"].values), q["added_mass"])
        assert np.array_equal(np.asarray(r.dataset["c_vector"].values)[0].real, q["c_vector"])
    assert t_engine < 4.0 * t_plan + 0.5, (t_engine, t_plan)


# ------------------------------------------------------------------------------------------------ waves, depths
def test_multi_wave_frequency_stage_is_bit_identical(ofb, monkeypatch):
    """The frequency stage in 3+ waves (per-wave counters, item offsets, reuse of the system buffers) reproduces the
    single-wave run bit for bit - CTA-per-system kernels (small_tricylinder_n20, 3 bodies) and the one-warp kernel."""
    cases = [(20.0, [5.0, 10.0, 15.0], [15.0, 10.0, 5.0], [20, 20, 20, 20], [0, 1, 2], [0, 1, 2], 40),
             (1.001, [0.125, 0.25], [0.25, 0.125], [10, 10, 10], [0, 0], [0], 300)]
    for h, a, d, NMK, bm, modes, F in cases:
        geo = ofb.FlatGeometry.from_body_map(h, a, d, NMK, bm, modes)
        om = np.linspace(0.2, 3.0, F)
        one = ofb.solve_batch([geo, geo, geo], om)
        monkeypatch.setenv("MEEMB200_WAVE_ITEMS", str(F))          # 3 geometries x F systems -> 3 waves
        many = ofb.solve_batch([geo, geo, geo], om)
        monkeypatch.delenv("MEEMB200_WAVE_ITEMS")
        for r1, r2 in zip(one, many):
            assert np.all(r1["status"] == 0) and np.all(r2["status"] == 0)
            for v in ("added_mass", "damping", "excitation_force", "excitation_phase"):
                assert np.array_equal(r1[v], r2[v]), v


def test_heterogeneous_depths_batch(ofb):
    """Every geometry its own water depth: one root set per geometry (hashed by (h, K) in the planner, k_mk folds the
    set index), results equal to solving each geometry alone and to the oracle for a few."""
    rng = np.random.default_rng(11)
    hs = rng.uniform(5.0, 80.0, 600)
    om = np.array([0.4, 1.2, 2.4])
    geos = [ofb.FlatGeometry(float(h), [1.0, 2.5], [0.3 * h, 0.1 * h], [8, 8, 8]) for h in hs]
    res = ofb.solve_batch(geos, om)
    for i in (0, 299, 599):
        alone = ofb.solve_batch([geos[i]], om)[0]
        for v in ("added_mass", "damping", "excitation_force", "excitation_phase", "m0"):
            assert np.array_equal(res[i][v], alone[v]), v
        ref = mo.run_sweep(mo.FlatGeometry(float(hs[i]), [1.0, 2.5], [0.3 * hs[i], 0.1 * hs[i]], [8, 8, 8]), om, [0, 0], [0])
        _check(res[i], ref)


def test_plan_argument_errors(ofb):
    """n_geom * n_freq beyond int32 and omega = inf in set_frequencies are refused with EINVAL (ValueError)."""
    geo = ofb.FlatGeometry(1.001, [0.5, 1.0], [0.5, 0.25], [4, 4, 4])
    plan = ofb.BatchPlan([geo], [1.0, 2.0])
    with pytest.raises(ValueError, match="finite"):
        plan.set_frequencies([1.0, np.inf])
    with pytest.raises(ValueError, match="positive"):
        plan.set_frequencies([1.0, -2.0])


def test_product_imk_table_entries(ofb):
    """Entry-level check of the coupling table the frequency stage really uses (angle-addition form + sinc series +
    exact-equality branch) against the oracle's reference-ordered I_mk_table (multi_equations.py:96-124): bicylinder with
    near-resonant (m_k - lambda) L < 1 columns, the asymptotic m0 h >= 14 branch, and a geometry built so that
    |m_k| == lambda exactly for one (m, k) pair."""
    cases = [(1.001, [0.5, 1.0], [0.5, 0.25], [30, 30, 30], [0.5, 2.0, 12.0]),
             (20.0, [5.0, 10.0, 15.0], [15.0, 10.0, 5.0], [50, 50, 50, 50], [0.3, 2.7, 3.0]),
             (50.0, [2.0, 5.0], [20.0, 3.0], [16, 16, 16], [0.3, 1.4, 2.5])]
    for h, a, d, NMK, oms in cases:
        geo = ofb.FlatGeometry(h, a, d, NMK)
        plan = ofb.BatchPlan([geo], oms)
        og = mo.FlatGeometry(h, a, d, NMK)
        near = 0
        for f, w in enumerate(oms):
            m0 = mo.wavenumber(w, h)
            mk = mo.m_k_array(NMK[-1], m0, h)
            Nk = mo.N_k_array(m0, h, mk)
            ref = mo.I_mk_table(og, m0, mk, Nk)
            got = plan.imk_table(0, f)
            assert got.shape == ref.shape
            assert np.max(np.abs(got - ref)) <= 1e-13 * np.max(np.abs(ref))
            lam = np.arange(NMK[-2]) * np.pi / (h - d[-1])
            near += int(np.sum(np.abs(mk[None, 1:] - lam[1:, None]) * (h - d[-1]) < 1.0))
        assert near > 0          # the sinc-series branch was exercised
    # exact equality |m_k| == lambda: with m0 = inf, m_k = (k - 1/2) pi / h; choose d so that lambda_n = n pi / (h - d) hits it:
    # h = 3, d = 1 -> lambda_1 = pi / 2 = m_2 ... (k - 1/2) / 3 = n / 2 -> k = 2, n = 1
    geo = ofb.FlatGeometry(3.0, [1.0], [1.0], [6, 6])
    plan = ofb.BatchPlan([geo], [np.inf], freq_kind=ofb.FREQ_M0)
    mk = mo.m_k_array(6, np.inf, 3.0)
    Nk = mo.N_k_array(np.inf, 3.0, mk)
    ref = mo.I_mk_table(mo.FlatGeometry(3.0, [1.0], [1.0], [6, 6]), np.inf, mk, Nk)
    got = plan.imk_table(0, 0)
    assert abs(mk[2]) == 1 * np.pi / 2.0                       # the branch is really hit
    assert np.max(np.abs(got - ref)) <= 1e-13 * np.max(np.abs(ref))


@pytest.mark.parametrize("env", [{"MEEMB200_FORCE_COOP": "1"}, {"MEEMB200_FORCE_COOP": "1", "MEEMB200_NO_COOP2": "1"}], ids=_ids)
@pytest.mark.parametrize("name", sorted(META2))
def test_cooperative_static_window_elimination(ofb, monkeypatch, name, env):
    """k_geom_coop2 (grid-cooperative window elimination, C1 blocks on their diagonal pivot blocks without a pivot search)
    forced on every golden sweep of the untouched reference - geometries with an overhang fall back to k_geom_coop - and
    the pivoted cooperative kernel on the same inputs."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    s = META2[name]
    geo = ofb.FlatGeometry.from_body_map(s["h"], s["a"], s["d"], s["NMK"], s["body_map"], s["modes"])
    plan = ofb.BatchPlan([geo, geo], s["omegas"])
    case1 = all(x > y for x, y in zip(s["d"][:-1], s["d"][1:]))
    assert plan.geom_kernel == ("k_geom_coop2" if case1 and "MEEMB200_NO_COOP2" not in env else "k_geom_coop")
    ref = {v: Z2[f"{name}/{v}"] for v in ("added_mass", "damping", "excitation_force", "excitation_phase")}
    for r in plan.run():
        assert np.all(r["status"] == 0)
        _check(r, ref)


def test_cooperative_static_unequal_modes_and_many_regions(ofb, monkeypatch):
    """k_geom_coop2 with unequal mode counts per region (odd block sizes, panels that are not multiples of 16), R = 2, a
    single cylinder (static step only) and an 8-region staircase, against the oracle."""
    monkeypatch.setenv("MEEMB200_FORCE_COOP", "1")
    cases = [(20.0, [5, 10, 15], [15, 10, 5], [30, 12, 25, 40], [0, 1, 1], [0, 1]),
             (20.0, [5, 10, 15], [15, 10, 5], [33, 47, 21, 9], [0, 0, 1], [0, 1]),
             (3.0, [1.0], [0.7], [9, 11], [0], [0]),
             (10.0, [0.75 * (i + 1) for i in range(8)], [6.0 - 0.6 * i for i in range(8)], [14] * 9, [0] * 8, [0])]
    om = np.array([0.4, 1.3, 2.6])
    for c in cases:
        geo = ofb.FlatGeometry.from_body_map(*c)
        plan = ofb.BatchPlan([geo], om)
        assert plan.geom_kernel == "k_geom_coop2"
        res = plan.run()[0]
        assert np.all(res["status"] == 0)
        _check(res, mo.run_sweep(mo.FlatGeometry(*c[:4]), om, c[4], c[5]))


def test_geometry_batch_and_stacked_download(ofb):
    """Packed GeometryBatch (design-sweep input) through BatchPlan / ShardedPlan: same bits as the list-of-FlatGeometry path,
    stacked [G, F, R, R] outputs without a per-geometry loop."""
    import bench
    from openflash_b200.distributed import ShardedPlan
    gs = bench.c4_geometries(300)
    om = np.linspace(0.3, 2.5, 6)
    geos = [ofb.FlatGeometry.from_body_map(g["h"], g["a"], g["d"], g["NMK"], g["body_map"], g["modes"]) for g in gs]
    ref = ofb.solve_batch(geos, om)
    gb = ofb.GeometryBatch.from_dicts(gs)
    plan = ofb.BatchPlan(gb, om)
    plan.upload()
    plan.solve()
    st = plan.download_stacked(copy=True)
    sp = ShardedPlan(gb, om)                       # world size 1: the product's multi-GPU entry degenerates to one BatchPlan
    sp.upload()
    sp.solve()
    st2 = sp.download_stacked()
    assert st["added_mass"].shape == (300, 6, 1, 1)
    for gi, r in enumerate(ref):
        for v in ("added_mass", "damping", "excitation_force", "excitation_phase", "m0", "status"):
            assert np.array_equal(st[v][gi], r[v]), v
            assert np.array_equal(st2[v][gi], r[v]), v
    sub = ofb.solve_batch(gb.take([5, 200, 77]), om)
    for k, gi in enumerate([5, 200, 77]):
        assert np.array_equal(sub[k]["added_mass"], ref[gi]["added_mass"])


STREAM_OFF = {"MEEMB200_NO_STREAM": "1", "MEEMB200_NO_STREAM_SYRK": "1", "MEEMB200_NO_BATCH_INV": "1"}


@pytest.mark.parametrize("modes,heaving", [(64, [0]), (72, [0]), (75, [0, 1]), (90, [2]), (99, [1]), (100, [0]), (104, [0, 2]), (128, [1]), (150, [0, 1]), (200, [0]),
                                           (200, [0, 1, 2]), (224, [0]), (248, [0, 1]), (256, [2])])
def test_chunk_stream_frequency_gemm_vs_classic_path_and_oracle(ofb, monkeypatch, modes, heaving):
    """k_freq_gemm's chunk-stream form (v19: cta_trmm_stream / cta_syrk_stream, batched inversion in the coupling table)
    against the tile-by-tile form it replaces (switched back on by the environment) and against the oracle: column-block
    counts with 0..7 left-over blocks (extra blocks dealt over the warps), 8- and 16-row chunks, mode counts that are not a
    multiple of 8 (zero-padded last chunk), odd mode counts (the pad column next to K - 1), 1-3 radiating modes."""
    h, a, d = 25.0, [10.0, 15.0, 20.0], [20.0, 15.0, 10.0]
    NMK = [modes] * 4
    oms = np.linspace(0.2, 2.9, 24)
    geo = ofb.FlatGeometry.from_body_map(h, a, d, NMK, [0, 1, 2], heaving)
    new = ofb.BatchPlan([geo], oms)
    assert "k_freq_gemm" in new.freq_kernel
    out_new = new.run()[0]
    for k, v in STREAM_OFF.items():
        monkeypatch.setenv(k, v)
    out_old = ofb.BatchPlan([geo], oms).run()[0]
    for k in STREAM_OFF:
        monkeypatch.delenv(k)
    assert np.all(out_new["status"] == 0) and np.all(out_old["status"] == 0)
    _check(out_new, out_old, tol=1e-12)
    og = mo.FlatGeometry(h, a, d, NMK)
    sel = [0, 11, 23]
    ref = mo.run_sweep(og, [float(oms[i]) for i in sel], [0, 1, 2], heaving)
    _check({k: out_new[k][sel] for k in ("added_mass", "damping", "excitation_force", "excitation_phase")}, ref)


def test_chunk_stream_heterogeneous_batch_ring_state(ofb):
    """One plan whose systems alternate between mode counts with different ring geometries (16- and 8-row chunks, 3 to 6 stages,
    with and without extra blocks) and the classic path (60 modes: fewer than 8 column blocks): the persistent CTAs carry the
    ring's per-stage parity bits from one system to the next, so every geometry must come out exactly as when solved alone."""
    h, a, d = 25.0, [10.0, 15.0, 20.0], [20.0, 15.0, 10.0]
    oms = np.linspace(0.25, 2.75, 40)
    modes = [200, 72, 248, 60, 104, 150, 232, 90]
    geos = [ofb.FlatGeometry.from_body_map(h, a, d, [m] * 4, [0, 1, 2], [0]) for m in modes]
    together = ofb.BatchPlan(geos, oms).run()
    for m, geo, got in zip(modes, geos, together):
        alone = ofb.BatchPlan([geo], oms).run()[0]
        assert np.all(got["status"] == 0), m
        for k in ("added_mass", "damping", "excitation_force", "excitation_phase"):
            assert np.array_equal(got[k], alone[k]), (m, k, float(np.max(np.abs(got[k] - alone[k]))))


@pytest.mark.parametrize("NMK", [[64, 48, 40, 96], [30, 72, 100, 64], [90, 90, 200, 104]])
def test_chunk_stream_unequal_mode_counts(ofb, monkeypatch, NMK):
    """N (modes of the last interior region) != K (exterior modes): the stream's row groups follow N and rank(H), its column
    blocks and ring follow K - against the classic path (1e-12) and the oracle (1e-10)."""
    h, a, d = 25.0, [10.0, 15.0, 20.0], [20.0, 15.0, 10.0]
    oms = np.linspace(0.3, 2.6, 12)
    geo = ofb.FlatGeometry.from_body_map(h, a, d, NMK, [0, 1, 2], [0, 2])
    out_new = ofb.BatchPlan([geo], oms).run()[0]
    for k, v in STREAM_OFF.items():
        monkeypatch.setenv(k, v)
    out_old = ofb.BatchPlan([geo], oms).run()[0]
    for k in STREAM_OFF:
        monkeypatch.delenv(k)
    assert np.all(out_new["status"] == 0) and np.all(out_old["status"] == 0)
    _check(out_new, out_old, tol=1e-12)
    sel = [0, 6, 11]
    ref = mo.run_sweep(mo.FlatGeometry(h, a, d, NMK), [float(oms[i]) for i in sel], [0, 1, 2], [0, 2])
    _check({k: out_new[k][sel] for k in ("added_mass", "damping", "excitation_force", "excitation_phase")}, ref)
