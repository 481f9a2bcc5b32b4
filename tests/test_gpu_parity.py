"""GPU parity tests: the CUDA path (through the C ABI) against the oracle and the committed golden
fixtures of the untouched reference.  Tolerances follow SURVEY §8c / BASELINE.json: coefficients
1e-10 relative, phase 1e-10 absolute (mod 2 pi), wavenumbers 5e-12 (roots are replayed, typically 1e-15)."""
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


def _algo(ofb, name):
    return {"schur": ofb.ALGO_SCHUR, "schur_lu": ofb.ALGO_SCHUR_LU, "dense": ofb.ALGO_DENSE}[name]


def _check(out, ref, tol=1e-10):
    for var in ("added_mass", "damping", "excitation_force"):
        r = ref[var]
        ok = np.isfinite(r)
        assert np.array_equal(np.isfinite(out[var]), ok), var
        assert relerr(out[var][ok], r[ok]) <= tol, (var, relerr(out[var][ok], r[ok]))
    assert phase_err(out["excitation_phase"], ref["excitation_phase"]) <= tol


def test_device_bessel_and_roots(ofb):
    from scipy.special import ive, kve, hankel1
    rng = np.random.default_rng(3)
    x = np.exp(rng.uniform(np.log(1e-3), np.log(3000), 20000))
    assert relerr(ofb.device_bessel(0, x), ive(0, x)) < 3e-15
    assert relerr(ofb.device_bessel(1, x), ive(1, x)) < 3e-15
    assert relerr(ofb.device_bessel(2, x), kve(0, x)) < 6e-15
    assert relerr(ofb.device_bessel(3, x), kve(1, x)) < 3e-15
    xs = np.exp(rng.uniform(np.log(1e-4), np.log(500), 20000))
    H0 = ofb.device_bessel(4, xs) + 1j * ofb.device_bessel(5, xs)
    H1 = ofb.device_bessel(6, xs) + 1j * ofb.device_bessel(7, xs)
    assert np.max(np.abs(H0 - hankel1(0, xs)) / np.abs(H0)) < 3e-15
    assert np.max(np.abs(H1 / H0 - hankel1(1, xs) / hankel1(0, xs)) / np.abs(H1 / H0)) < 4e-15
    for h in (1.001, 20.0, 50.0):
        om = np.linspace(0.1, 3.0, 64)
        for mode, tol in ((ofb.ROOTS_REFERENCE, 2e-14), (ofb.ROOTS_CONVERGED, 5e-12)):
            m0, mk, nk = ofb.device_roots(om, h, 50, root_mode=mode)
            ref_m0 = np.array([mo.wavenumber(o, h) for o in om])
            assert relerr(m0, ref_m0) <= tol
            for i in (0, 17, 63):
                ref_mk = mo.m_k_array(50, ref_m0[i], h)
                assert relerr(mk[i], ref_mk) <= tol
                k = np.arange(1, 50)   # mode indexing: root k in ((k-1/2)pi, k pi]/h
                assert np.all(mk[i, 1:] * h > (k - 0.5) * np.pi) and np.all(mk[i, 1:] * h <= k * np.pi * (1 + 5e-16))
                assert relerr(nk[i], mo.N_k_array(ref_m0[i], h, ref_mk)) <= 1e-9


@pytest.mark.parametrize("algo", ["schur", "schur_lu", "dense"])
@pytest.mark.parametrize("name", sorted(META1))
def test_single_solves_vs_reference(ofb, name, algo):
    """A, b, X and coefficients for the golden (case, heaving body, m0) set, m0 injected."""
    c = META1[name]
    nb = max(c["body_map"]) + 1
    geo = ofb.FlatGeometry.from_body_map(c["h"], c["a"], c["d"], c["NMK"], c["body_map"], list(range(nb)))
    plan = ofb.BatchPlan([geo], c["m0"], freq_kind=ofb.FREQ_M0, want_x=True,
                         algo=_algo(ofb, algo))
    res = plan.run()[0]
    assert np.all(res["status"] == 0)
    for mi, m0 in enumerate(c["m0"]):
        if algo == "dense":
            A, b, cv, phi = plan.assemble(0, mi)
        for body in range(nb):
            key = f"{name}/b{body}/m{mi}"
            if algo == "dense":
                assert relerr(b[body], Z1[key + "/b"]) <= 1e-12
                if body == 0 and key + "/A" in Z1:
                    Aref = Z1[key + "/A"]
                    assert np.array_equal(A != 0, Aref != 0)
                    assert relerr(A, Aref) <= 1e-12
            Xref = Z1[key + "/X"]
            X = res["X"][mi, body]
            assert np.max(np.abs(X - Xref)) <= 1e-9 * np.max(np.abs(Xref))
            # reference dicts: one entry per mode j for the solution of mode `body`
            assert relerr(res["added_mass"][mi, :, body], Z1[key + "/real"]) <= 1e-10
            assert relerr(res["damping"][mi, :, body], Z1[key + "/imag"]) <= 1e-10
            assert phase_err(res["excitation_phase"][mi, body], Z1[key + "/phase"][body]) <= 1e-10
            f_ref = Z1[key + "/force"][body]
            if np.isfinite(f_ref):
                assert relerr(res["excitation_force"][mi, body], f_ref) <= 1e-10
            else:
                assert not np.isfinite(res["excitation_force"][mi, body])


@pytest.mark.parametrize("algo", ["schur", "schur_lu", "dense"])
@pytest.mark.parametrize("name", sorted(META2))
def test_sweeps_vs_run_and_store_results(ofb, name, algo):
    s = META2[name]
    geo = ofb.FlatGeometry.from_body_map(s["h"], s["a"], s["d"], s["NMK"], s["body_map"], s["modes"])
    res = ofb.solve_batch([geo], s["omegas"], algo=_algo(ofb, algo), want_x=True)[0]
    assert relerr(res["m0"], Z2[f"{name}/m0"]) <= 2e-14
    ref = {v: Z2[f"{name}/{v}"] for v in ("added_mass", "damping", "excitation_force", "excitation_phase")}
    _check(res, ref)
    pot = Z2[f"{name}/potentials"]
    NMK = s["NMK"]
    col = 0
    for dom in range(len(NMK)):
        width = NMK[dom] if dom in (0, len(NMK) - 1) else 2 * NMK[dom]
        refp = pot[:, :, dom, :width]
        got = res["X"][:, :, col:col + width]
        assert np.max(np.abs(got - refp)) <= 1e-9 * np.max(np.abs(refp))
        col += width


def test_m0_infinite_limit(ofb):
    """solve_linear_system_multi(problem, np.inf) (test_high_frequency_convergence.py)."""
    geo_o = mo.FlatGeometry(1.001, [0.5, 1.0], [0.5, 0.25], [12, 12, 12])
    flags = [1, 1]
    X = mo.solve_single(geo_o, flags, np.inf)
    ref = mo.hydro_coefficients(geo_o, X, np.inf, [flags])[0]
    for algo in (ofb.ALGO_SCHUR, ofb.ALGO_SCHUR_LU, ofb.ALGO_DENSE):
        geo = ofb.FlatGeometry(1.001, [0.5, 1.0], [0.5, 0.25], [12, 12, 12], [flags])
        res = ofb.solve_batch([geo], [np.inf], freq_kind=ofb.FREQ_M0, algo=algo, want_x=True)[0]
        assert relerr(res["added_mass"][0, 0, 0], ref["real"]) <= 1e-10
        assert res["damping"][0, 0, 0] == 0.0
        assert abs(res["excitation_phase"][0, 0] + np.pi / 2) < 1e-15
        assert np.max(np.abs(res["X"][0, 0] - X)) <= 1e-9 * np.max(np.abs(X))


def test_overflow_gives_nan_and_status(ofb):
    """Extreme mode counts overflow exp() -> inf in A -> NaN outputs, batch continues (eng.py:611-614)."""
    bad = ofb.FlatGeometry(1.001, [0.5, 1.0], [0.5, 0.25], [400] * 3)
    good = ofb.FlatGeometry(1.001, [0.5, 1.0], [0.5, 0.25], [20] * 3)
    res = ofb.solve_batch([bad, good], [1.0, 2.0])
    assert np.all(res[0]["status"] != 0) and not np.isfinite(res[0]["added_mass"]).any()
    assert np.all(res[1]["status"] == 0) and np.isfinite(res[1]["added_mass"]).all()


def test_heterogeneous_batch_vs_oracle(ofb):
    """Random N-cylinder geometries (SURVEY §8d C4 generator, small): mixed M, overhangs, shared omega grid."""
    rng = np.random.default_rng(20260101)
    geos, ogeos = [], []
    for _ in range(24):
        M = int(rng.integers(2, 9))
        h = 50.0
        a = np.cumsum(rng.uniform(1.0, 4.0, M))
        while True:
            d = np.sort(rng.uniform(0.04, 0.6, M) * h)[::-1]
            if M == 1 or np.min(-np.diff(d)) >= 0.01 * h:
                break
        if rng.uniform() < 0.1:
            d = rng.permutation(d)
        NMK = [8] * (M + 1)
        geos.append(ofb.FlatGeometry(h, a, d, NMK))
        ogeos.append(mo.FlatGeometry(h, a, d, NMK))
    om = np.linspace(0.3, 2.5, 5)
    for algo in (ofb.ALGO_SCHUR, ofb.ALGO_SCHUR_LU, ofb.ALGO_DENSE):
        res = ofb.solve_batch(geos, om, algo=algo)
        for g, og in zip(res, ogeos):
            ref = mo.run_sweep(og, om, [0] * og.M, [0])
            _check(g, ref)


@pytest.mark.parametrize("name", sorted(META2))
def test_cooperative_geometry_stage(ofb, name, monkeypatch):
    """The grid-cooperative LU of A11 (k_geom_coop, normally chosen only for few large geometries) forced on
    the golden sweeps: same parity bar as the one-CTA-per-geometry kernel."""
    monkeypatch.setenv("MEEMB200_FORCE_COOP", "1")
    s = META2[name]
    geo = ofb.FlatGeometry.from_body_map(s["h"], s["a"], s["d"], s["NMK"], s["body_map"], s["modes"])
    res = ofb.solve_batch([geo, geo], s["omegas"], want_x=True)
    ref = {v: Z2[f"{name}/{v}"] for v in ("added_mass", "damping", "excitation_force", "excitation_phase")}
    for r in res:
        assert np.all(r["status"] == 0)
        _check(r, ref)


def test_cooperative_large_system_vs_oracle(ofb):
    """small_tricylinder at 50 modes/region (p = 250, odd-sized panels at the end) picks the cooperative path
    by itself; compare with the oracle's dense solve."""
    og = mo.FlatGeometry(20.0, [5, 10, 15], [15, 10, 5], [50] * 4)
    om = [0.4, 1.3, 2.9]
    ref = mo.run_sweep(og, om, [0, 1, 1], [0, 1])
    geo = ofb.FlatGeometry.from_body_map(20.0, [5, 10, 15], [15, 10, 5], [50] * 4, [0, 1, 1], [0, 1])
    res = ofb.solve_batch([geo], om)[0]
    assert np.all(res["status"] == 0)
    _check(res, ref)


def test_big_tricylinder_full_size(ofb):
    """BASELINE.json configs[2] at full size (N_T = 1200, 200 modes/region): the product path (cooperative LU,
    DMMA GEMMs, Cholesky) against the oracle at three frequencies, and against the independent reference-shaped
    dense complex LU on the device over a sweep that crosses the m0 h = 14 branch."""
    h, a, d, NMK = 25.0, [10, 15, 20], [20, 15, 10], [200] * 4
    og = mo.FlatGeometry(h, a, d, NMK)
    om_o = [0.1, 1.37, 3.0]
    ref = mo.run_sweep(og, om_o, [0, 0, 0], [0])
    geo = ofb.FlatGeometry.from_body_map(h, a, d, NMK, [0, 0, 0], [0])
    res = ofb.solve_batch([geo], om_o)[0]
    assert np.all(res["status"] == 0)
    _check(res, ref)
    om = np.linspace(0.1, 3.0, 24)
    r_s = ofb.solve_batch([geo], om, algo=ofb.ALGO_SCHUR)[0]
    r_d = ofb.solve_batch([geo], om, algo=ofb.ALGO_DENSE)[0]
    assert np.all(r_s["status"] == 0) and np.all(r_d["status"] == 0)
    _check(r_s, r_d)


def test_sweep_is_independent_of_batching(ofb):
    """Size-independent property: a system's result does not depend on what else is in the batch (work queues,
    cost sorting, waves): every frequency of a 4096-point sweep solved alone in small batches reproduces the
    big batch bit for bit (same kernels, same operation order)."""
    geo = ofb.FlatGeometry.from_body_map(20.0, [5, 10, 15], [15, 10, 5], [50] * 4, [0, 1, 1], [0, 1])
    om = np.linspace(0.1, 3.0, 4096)
    full = ofb.solve_batch([geo], om)[0]
    assert np.all(full["status"] == 0)
    idx = np.array([0, 1, 777, 2048, 4095])
    part = ofb.solve_batch([geo], om[idx])[0]
    for v in ("added_mass", "damping", "excitation_force", "excitation_phase"):
        assert np.array_equal(full[v][idx], part[v]), v
    # physics sanity over the whole sweep: radiation damping is positive and the 2 x 2 added mass is finite
    assert np.all(np.diagonal(full["damping"], axis1=1, axis2=2) > 0)
    assert np.isfinite(full["added_mass"]).all()


def test_geometry_sweep_c4_vs_oracle(ofb):
    """BASELINE.json configs[3] generator at its own mode count (16/region), first 12 geometries x 4 frequencies."""
    import bench
    gs = bench.c4_geometries(12)
    om = np.linspace(0.3, 2.5, 4)
    geos = [ofb.FlatGeometry.from_body_map(g["h"], g["a"], g["d"], g["NMK"], g["body_map"], g["modes"]) for g in gs]
    res = ofb.solve_batch(geos, om)
    for g, r in zip(gs, res):
        og = mo.FlatGeometry(g["h"], g["a"], g["d"], g["NMK"])
        _check(r, mo.run_sweep(og, om, g["body_map"], g["modes"]))


@pytest.mark.parametrize("env", [{"MEEMB200_SOLVE_VARIANT": "0"}, {"MEEMB200_SOLVE_VARIANT": "1"}, {"MEEMB200_SOLVE_VARIANT": "2"},
                                 {"MEEMB200_SOLVE_VARIANT": "3"}, {"MEEMB200_NO_HCHOL": "1"}, {"MEEMB200_NO_SMALL": "1"},
                                 {"MEEMB200_NO_SMALL": "1", "MEEMB200_NO_SYRK": "1"}, {"MEEMB200_NO_SMALL": "1", "MEEMB200_NO_BULK": "1"},
                                 {"MEEMB200_NO_SMALL": "1", "MEEMB200_NO_LL": "1"}],
                         ids=lambda e: "-".join(f"{k[9:]}={v}" for k, v in e.items()))
@pytest.mark.parametrize("name", sorted(META2))
def test_schur_kernel_variants_vs_reference(ofb, monkeypatch, name, env):
    """Every k_freq_solve variant (P in its L2 buffer with the left- or right-looking Cholesky / resident in shared memory
    at 128, 256, 512 threads), the unfactorised-H form of k_freq_gemm, its rank update with rectangular tiles, cp.async or TMA
    bulk staging, and the CTA-per-system path for small systems meet the same bar as the default path, X included (the
    overrides are read at plan creation)."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    s = META2[name]
    geo = ofb.FlatGeometry.from_body_map(s["h"], s["a"], s["d"], s["NMK"], s["body_map"], s["modes"])
    res = ofb.solve_batch([geo], s["omegas"], algo=ofb.ALGO_SCHUR, want_x=True)[0]
    assert np.all(res["status"] == 0)
    ref = {v: Z2[f"{name}/{v}"] for v in ("added_mass", "damping", "excitation_force", "excitation_phase")}
    _check(res, ref)
    pot = Z2[f"{name}/potentials"]
    NMK, col = s["NMK"], 0
    for dom in range(len(NMK)):
        width = NMK[dom] if dom in (0, len(NMK) - 1) else 2 * NMK[dom]
        refp = pot[:, :, dom, :width]
        assert np.max(np.abs(res["X"][:, :, col:col + width] - refp)) <= 1e-9 * np.max(np.abs(refp))
        col += width
    res2 = ofb.solve_batch([geo], s["omegas"], algo=ofb.ALGO_SCHUR)[0]          # coefficient-only path (no back-substitution)
    _check(res2, ref)


@pytest.mark.parametrize("small", [False, True], ids=["k_freq_solve", "k_freq_small"])
def test_failed_cholesky_systems_are_resolved_by_lu(ofb, monkeypatch, small):
    """A system whose Cholesky reports a non-positive pivot (status bit 2) is re-solved by the pivoted LU of the unreduced
    K x K system (k_freq_lu, only_flagged pass) and leaves with status 0.  meemb200_plan_set_test_hook(plan, 3) makes the solve kernels
    report every third system as failed: those systems must come back within the parity bar (and not bit-identical to the
    Cholesky answer - the re-solve really ran), the others bit-identical to the unhooked run.  The scan for flagged systems
    covers more systems than one CTA pass (blockDim x gridDim), so its chunked loop wraps."""
    if small:
        h, a, d, NMK, bm, modes = 1.001, [0.125, 0.25], [0.25, 0.125], [10, 10, 10], [0, 0], [0]
        om = np.linspace(0.1, 3.0, 600)
    else:
        h, a, d, NMK, bm, modes = 20.0, [5.0, 10.0, 15.0], [15.0, 10.0, 5.0], [20, 20, 20, 20], [0, 1, 2], [0, 1, 2]
        om = np.linspace(0.2, 3.0, 60)
    geo = ofb.FlatGeometry.from_body_map(h, a, d, NMK, bm, modes)
    # many copies of the geometry: > 296 CTAs x 256 threads status words for the k_freq_solve case would be slow, so the
    # wrap-around of the scan is exercised with the small systems (600 omega x 160 geometries = 96 000 systems)
    geos = [geo] * (160 if small else 2)
    plain = ofb.solve_batch(geos, om)
    hplan = ofb.BatchPlan(geos, om)
    hplan.set_test_hook(3)
    hooked = hplan.run()
    ref = mo.run_sweep(mo.FlatGeometry(h, a, d, NMK), om, bm, modes)
    F = len(om)
    differs = 0
    for gi, (rp, rh) in enumerate(zip(plain, hooked)):
        assert np.all(rh["status"] == 0) and np.all(rp["status"] == 0)
        if gi < 2 or gi == len(geos) - 1:
            _check(rh, ref)
        flagged = (gi * F + np.arange(F)) % 3 == 0
        for v in ("added_mass", "damping", "excitation_force"):
            assert np.array_equal(rp[v][~flagged], rh[v][~flagged])
            differs += int(np.any(rp[v][flagged] != rh[v][flagged]))
            np.testing.assert_allclose(rh[v][flagged], rp[v][flagged], rtol=1e-8, atol=1e-10 * np.max(np.abs(rp[v])))
    assert differs > 0


@pytest.mark.parametrize("geom", ["small_bicylinder", "small_tricylinder"])
def test_mode_count_convergence_sweep_c5(ofb, geom):
    """BASELINE.json configs[4]: mode-count convergence sweep, equal modes per region, outputs incl. the excitation
    force / phase (the reference's only incident-wave quantity).  All mode counts go through ONE heterogeneous batch
    (factorised and unfactorised k_freq_gemm, SYRK queue and rectangular tiles, k_freq_solve at K = 10 .. 300 all occur); where the
    reference overflows to NaN at high mode counts the device must produce NaN + status, not garbage."""
    h, a, d, bm = {"small_bicylinder": (1.001, [0.5, 1.0], [0.5, 0.25], [0, 0]),
                   "small_tricylinder": (20.0, [5.0, 10.0, 15.0], [15.0, 10.0, 5.0], [0, 0, 0])}[geom]
    ns = [10, 15, 20, 30, 40, 50, 75, 100, 150, 200] + ([300] if geom == "small_bicylinder" else [])
    om = np.array([0.5, 1.75, 3.0])
    geos = [ofb.FlatGeometry.from_body_map(h, a, d, [n] * (len(a) + 1), bm, [0]) for n in ns]
    res = ofb.solve_batch(geos, om)
    n_finite = 0
    for n, r in zip(ns, res):
        ref = mo.run_sweep(mo.FlatGeometry(h, a, d, [n] * (len(a) + 1)), om, bm, [0])
        _check(r, ref)
        ok = np.isfinite(ref["added_mass"]).reshape(len(om), -1).all(axis=1)
        assert np.array_equal(r["status"] == 0, ok), n
        n_finite += int(ok.sum())
    assert n_finite >= 3 * 8          # the sweep is not vacuous: at least 8 mode counts solve at every frequency
    # convergence itself (size-independent property): the finest two finite levels agree to 3 digits
    fin = [r["added_mass"][:, 0, 0] for r in res if np.all(r["status"] == 0)]
    assert np.max(np.abs(fin[-1] - fin[-2]) / np.abs(fin[-1])) < 1e-3


@pytest.mark.parametrize("M", [12, 24])
def test_many_region_staircase_vs_oracle(ofb, M):
    """SURVEY 8f rank 4 (slant-discretised bodies, dev/python/slants/helpers.py:80-125): a cone approximated by a staircase of
    M regions - one body, all regions heaving - plus an overhanging variant (depth increasing outwards over the last third).
    The device solves it with the same kernels (dense pivoted LU of the p x p block, p up to 376), not yet with a
    block-bidiagonal elimination; parity bar unchanged."""
    h, a_out, n = 10.0, 6.0, 8
    a = [a_out * (i + 1) / M for i in range(M)]
    d_cone = [6.0 - 5.0 * (i + 0.5) / M for i in range(M)]                       # depth decreasing outwards (case 1 everywhere)
    d_over = [x if i < 2 * M // 3 else d_cone[2 * M // 3 - 1] + 0.05 * (i - 2 * M // 3 + 1) for i, x in enumerate(d_cone)]
    om = np.array([0.6, 1.4, 2.2])
    for d in (d_cone, d_over):
        og = mo.FlatGeometry(h, a, d, [n] * (M + 1))
        ref = mo.run_sweep(og, om, [0] * M, [0])
        geo = ofb.FlatGeometry.from_body_map(h, a, d, [n] * (M + 1), [0] * M, [0])
        for algo in (ofb.ALGO_SCHUR, ofb.ALGO_DENSE):
            res = ofb.solve_batch([geo], om, algo=algo)[0]
            assert np.all(res["status"] == 0)
            _check(res, ref)
