"""GPU tests of the reference-named API (MEEMEngine / MEEMProblem / Geometry) against the golden
outputs of the untouched reference's same calls."""
import numpy as np
import pytest

from helpers import load_npz, relerr, phase_err

pytestmark = pytest.mark.gpu

Z1, META1 = load_npz("ref_single.npz")
Z2, META2 = load_npz("ref_sweeps.npz")


def _problem(ofb, c, heaving_map):
    geo = ofb.BasicRegionGeometry.from_vectors(a=np.asarray(c["a"], float), d=np.asarray(c["d"], float), h=c["h"],
                                               NMK=list(c["NMK"]), body_map=list(c["body_map"]), heaving_map=list(heaving_map))
    return ofb.MEEMProblem(geo)


@pytest.mark.parametrize("name", ["mini_bicylinder", "overhang_bicylinder", "config3_mixed5", "single_cylinder"])
def test_engine_pieces_match_reference_calls(name):
    """assemble_A_multi / assemble_b_multi / solve_linear_system_multi / compute_hydrodynamic_coefficients."""
    import openflash_b200 as ofb
    c = META1[name]
    nb = max(c["body_map"]) + 1
    for body in range(nb):
        hm = [False] * nb
        hm[body] = True
        prob = _problem(ofb, c, hm)
        eng = ofb.MEEMEngine([prob])
        for mi, m0 in list(enumerate(c["m0"]))[1:4:2]:
            key = f"{name}/b{body}/m{mi}"
            A = eng.assemble_A_multi(prob, m0)
            b = eng.assemble_b_multi(prob, m0)
            cache = eng.cache_list[prob]
            if body == 0:
                assert relerr(cache.m_k_arr, Z1[key + "/mk"]) <= 2e-14
                assert relerr(cache.N_k_arr, Z1[key + "/Nk"]) <= 1e-12
                if key + "/A" in Z1:
                    assert relerr(A, Z1[key + "/A"]) <= 1e-12
            assert relerr(b, Z1[key + "/b"]) <= 1e-12
            X = eng.solve_linear_system_multi(prob, m0)
            Xref = Z1[key + "/X"]
            assert np.max(np.abs(X - Xref)) <= 1e-9 * np.max(np.abs(Xref))
            res = eng.compute_hydrodynamic_coefficients(prob, Xref, m0)
            assert [r["mode"] for r in res] == list(range(nb))
            assert relerr([r["real"] for r in res], Z1[key + "/real"]) <= 1e-11
            assert relerr([r["imag"] for r in res], Z1[key + "/imag"]) <= 1e-11
            assert phase_err([r["excitation_phase"] for r in res], Z1[key + "/phase"]) <= 1e-11
            f_ref = Z1[key + "/force"]
            f = np.array([r["excitation_force"] for r in res])
            ok = np.isfinite(f_ref)
            assert np.array_equal(np.isfinite(f), ok) and relerr(f[ok], f_ref[ok]) <= 1e-11
            assert res[0]["c_vector"].shape == (A.shape[0] - c["NMK"][-1],)


@pytest.mark.parametrize("name", ["tutorial_2body", "C2_small_tricylinder_R2", "config3_all5", "single_cylinder"])
def test_run_and_store_results_dataset(name):
    import openflash_b200 as ofb
    s = META2[name]
    nb = max(s["body_map"]) + 1
    prob = _problem(ofb, s, [False] * nb)
    for bi, body in enumerate(prob.geometry.body_arrangement.bodies):
        body.heaving = bi in s["modes"]
    prob.set_frequencies(np.asarray(s["omegas"], float))
    eng = ofb.MEEMEngine([prob])
    ds = eng.run_and_store_results(0).dataset
    for var in ("added_mass", "damping", "excitation_force"):
        ref = Z2[f"{name}/{var}"]
        got = np.asarray(ds[var].values)
        assert got.shape == ref.shape
        ok = np.isfinite(ref)
        assert np.array_equal(np.isfinite(got), ok) and relerr(got[ok], ref[ok]) <= 1e-10
    assert phase_err(np.asarray(ds["excitation_phase"].values), Z2[f"{name}/excitation_phase"]) <= 1e-10
    pot = np.asarray(ds["potentials_real"].values) + 1j * np.asarray(ds["potentials_imag"].values)
    ref = Z2[f"{name}/potentials"]
    assert pot.shape == ref.shape
    assert np.array_equal(np.isnan(pot.real), np.isnan(ref.real))
    m = ~np.isnan(ref.real)
    assert np.max(np.abs(pot[m] - ref[m])) <= 1e-9 * np.max(np.abs(ref[m]))
    assert ds["c_vector"].values.shape[:2] == (len(s["omegas"]), len(s["modes"]))


def test_run_batch_many_problems_one_launch():
    import openflash_b200 as ofb
    probs = []
    for name in ("tutorial_2body", "overhang_2body"):
        s = META2[name]
        nb = max(s["body_map"]) + 1
        prob = _problem(ofb, s, [False] * nb)
        for bi, body in enumerate(prob.geometry.body_arrangement.bodies):
            body.heaving = bi in s["modes"]
        prob.set_frequencies(np.array([0.4, 1.3, 4.0]))
        probs.append(prob)
    eng = ofb.MEEMEngine(probs)
    out = eng.run_batch(probs)
    ref = Z2["overhang_2body/added_mass"]
    assert relerr(np.asarray(out[1].dataset["added_mass"].values), ref) <= 1e-10


def test_plan_rejects_bad_geometry():
    import openflash_b200 as ofb
    with pytest.raises(ValueError):
        ofb.BatchPlan([ofb.FlatGeometry(1.0, [1.0, 0.5], [0.2, 0.1], [3, 3, 3])], [1.0])
    with pytest.raises(ValueError):
        ofb.BatchPlan([ofb.FlatGeometry(1.0, [0.5], [1.2], [3, 3])], [1.0])
    with pytest.raises(ValueError):
        ofb.BatchPlan([ofb.FlatGeometry(1.0, [0.5], [0.2], [3, 3])], [-1.0])


ZF, METAF = load_npz("ref_fields.npz")


@pytest.mark.parametrize("name", sorted(METAF))
def test_fields_vs_reference_golden(name):
    """MEEMEngine.calculate_potentials / calculate_velocities (device kernel k_fields through meemb200_fields) against
    the untouched reference's outputs for the reference's own solution vector: same grid, same NaN masks, values
    within 1e-11 of the field's scale (different summation order over the harmonics)."""
    import openflash_b200 as ofb
    c = METAF[name]
    nb = max(c["body_map"]) + 1
    prob = _problem(ofb, c, [False] * nb)
    for bi, body in enumerate(prob.geometry.body_arrangement.bodies):
        body.heaving = bool(c["heaving_bodies"][bi])
    eng = ofb.MEEMEngine([prob])
    X = ZF[name + "/X"]
    pot = eng.calculate_potentials(prob, X, c["m0"], c["res"], c["sharp"])
    vel = eng.calculate_velocities(prob, X, c["m0"], c["res"], c["sharp"])
    assert np.array_equal(pot["R"], ZF[name + "/R"]) and np.array_equal(pot["Z"], ZF[name + "/Z"])
    for src, keys in ((pot, ("phiH", "phiP", "phi")), (vel, ("vrH", "vzH", "vrP", "vzP", "vr", "vz"))):
        for k in keys:
            ref, got = ZF[f"{name}/{k}"], src[k]
            assert np.array_equal(np.isnan(ref), np.isnan(got)), (name, k)
            ok = ~np.isnan(ref)
            scale = max(np.max(np.abs(ref[ok])), 1e-300)
            assert np.max(np.abs(got[ok] - ref[ok])) <= 1e-11 * scale, (name, k, np.max(np.abs(got[ok] - ref[ok])) / scale)


def test_fields_of_device_solution_vs_oracle():
    """End to end on the device: solve on the GPU, evaluate the fields of that X on the GPU, compare with the oracle's
    fields of the oracle's X (tolerance of X itself: 1e-9 of the field scale)."""
    import openflash_b200 as ofb
    from oracle import meem_oracle as mo
    h, a, d, NMK = 20.0, [5.0, 10.0, 15.0], [15.0, 10.0, 5.0], [20] * 4
    m0 = 0.35
    flags = [1, 1, 0]
    og = mo.FlatGeometry(h, a, d, NMK)
    Xo = mo.solve_single(og, flags, m0)
    rng = np.random.default_rng(5)
    r = np.concatenate([rng.uniform(0.0, 30.0, 500), np.asarray(a), [0.0]])
    z = np.concatenate([rng.uniform(-h, 0.0, 500), [-16.0, -12.0, -7.0], [-18.0]])
    fo = mo.fields(og, Xo, m0, flags, r, z)
    geo = ofb.FlatGeometry(h, a, d, NMK, [flags])
    plan = ofb.BatchPlan([geo], [m0], freq_kind=ofb.FREQ_M0, want_x=True)
    Xd = plan.run()[0]["X"][0, 0]
    raw = plan.fields(Xd, r, z)
    for k in ("vrH", "vzH"):
        scale = np.max(np.abs(fo[k]))
        assert np.max(np.abs(raw[k] - fo[k])) <= 1e-9 * scale, k
    ok = ~np.isnan(fo["phiH"])
    assert np.max(np.abs(raw["phiH"][ok] - fo["phiH"][ok])) <= 1e-9 * np.max(np.abs(fo["phiH"][ok]))
    assert np.max(np.abs(raw["phiP"][ok] - fo["phiP"][ok].real)) <= 1e-13 * np.max(np.abs(fo["phiP"][ok]))
