"""Generate golden fixtures by running the UNTOUCHED reference (build container only).

    python tests/golden/make_golden.py

Writes (small, committed):
  tests/golden/ref_single.npz   per-case A, b, X, m_k, N_k and per-mode coefficients from
                                MEEMEngine.assemble_A_multi / assemble_b_multi /
                                solve_linear_system_multi / compute_hydrodynamic_coefficients
  tests/golden/ref_sweeps.npz   MEEMEngine.run_and_store_results outputs (added_mass, damping,
                                excitation_force, excitation_phase, potentials) incl. R=2 cases
  tests/golden/debug_nondim_added_mass.json   values of the reference's own stored known-answer
                                CSVs debug_results/DEBUG_*_regenerated_v2.csv (column of_nondim)
  tests/golden/gold_standard_A_2region.npy    copy of package/test/data/gold_standard_A_2region.npy

The GPU box has no /root/reference; tests there read only these files.
"""
import contextlib
import io
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_loader  # noqa: E402

of = ref_loader.load()

CASES = {
    # name: h, a, d, NMK, body_map           (SURVEY §8 configs + edge layouts)
    "mini_bicylinder": dict(h=1.001, a=[0.125, 0.25], d=[0.25, 0.125], NMK=[10] * 3, body_map=[0, 0]),
    "small_bicylinder_2body": dict(h=1.001, a=[0.5, 1.0], d=[0.5, 0.25], NMK=[30] * 3, body_map=[0, 1]),
    "overhang_bicylinder": dict(h=1.001, a=[0.5, 1.0], d=[0.25, 0.5], NMK=[6, 7, 8], body_map=[0, 1]),
    "single_cylinder": dict(h=3.0, a=[1.0], d=[0.7], NMK=[9, 11], body_map=[0]),
    "config3_mixed5": dict(h=1.9, a=[0.3, 0.5, 1.0, 1.2, 1.6], d=[0.5, 0.7, 0.8, 0.2, 0.5],
                           NMK=[8, 9, 10, 11, 12, 13], body_map=[0, 1, 2, 3, 4]),
    "small_tricylinder_n20": dict(h=20.0, a=[5, 10, 15], d=[15, 10, 5], NMK=[20] * 4, body_map=[0, 1, 1]),
    "equal_depth_bicylinder": dict(h=2.0, a=[0.4, 0.9], d=[0.5, 0.5], NMK=[5, 5, 6], body_map=[0, 1]),
}
M0_LIST = [0.05, 0.3, 1.0, 5.0, 20.0]


def make_problem(c, heaving_map):
    geo = of.BasicRegionGeometry.from_vectors(
        a=np.asarray(c["a"], float), d=np.asarray(c["d"], float), h=c["h"], NMK=list(c["NMK"]),
        body_map=list(c["body_map"]), heaving_map=list(heaving_map))
    return of.MEEMProblem(geo)


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


def gen_single():
    out = {}
    meta = {}
    for name, c in CASES.items():
        nb = max(c["body_map"]) + 1
        meta[name] = dict(c, m0=M0_LIST)
        for body in range(nb):
            hm = [False] * nb
            hm[body] = True
            prob = make_problem(c, hm)
            eng = of.MEEMEngine([prob])
            for mi, m0 in enumerate(M0_LIST):
                key = f"{name}/b{body}/m{mi}"
                A = eng.assemble_A_multi(prob, m0)
                b = eng.assemble_b_multi(prob, m0)
                cache = eng.cache_list[prob]
                X = eng.solve_linear_system_multi(prob, m0)
                res = quiet(eng.compute_hydrodynamic_coefficients, prob, X, m0)
                if body == 0:
                    if A.shape[0] <= 110:
                        out[key + "/A"] = A
                    out[key + "/mk"] = cache.m_k_arr.copy()
                    out[key + "/Nk"] = cache.N_k_arr.copy()
                out[key + "/b"] = b
                out[key + "/X"] = X
                out[key + "/real"] = np.array([r["real"] for r in res])
                out[key + "/imag"] = np.array([r["imag"] for r in res])
                out[key + "/phase"] = np.array([r["excitation_phase"] for r in res])
                out[key + "/force"] = np.array([r["excitation_force"] for r in res])
                if mi == 0:          # the c vectors (one per mode j) do not depend on m0
                    out[key + "/c"] = np.array([r["c_vector"] for r in res])
    out["__meta__"] = np.array(json.dumps(meta))
    np.savez_compressed(os.path.join(HERE, "ref_single.npz"), **out)
    print("ref_single.npz:", len(out), "arrays")


SWEEPS = {
    # name: case, NMK override, omegas, heaving bodies (set after construction for R>1)
    "C1_mini_bicylinder": dict(case="mini_bicylinder", NMK=[10] * 3,
                               omegas=[float(np.sqrt(np.tanh(1.001) * 9.81)), 0.5, 3.0, 7.0], modes=[0]),
    "C2_small_tricylinder_R2": dict(h=20.0, a=[5, 10, 15], d=[15, 10, 5], NMK=[50] * 4, body_map=[0, 1, 1],
                                    omegas=list(np.linspace(0.1, 3.0, 256)[[0, 1, 37, 100, 180, 221, 222, 223, 255]]),
                                    modes=[0, 1]),
    "C3_big_tricylinder": dict(h=25.0, a=[10, 15, 20], d=[20, 15, 10], NMK=[200] * 4, body_map=[0, 0, 0],
                               omegas=list(np.linspace(0.1, 3.0, 4096)[[0, 2047, 4095]]), modes=[0]),
    "tutorial_2body": dict(case="small_bicylinder_2body", NMK=[30] * 3,
                           omegas=[float(np.sqrt(1.0 * np.tanh(1.0 * 1.001) * 9.81))], modes=[0, 1]),
    "config3_all5": dict(case="config3_mixed5", NMK=[8, 9, 10, 11, 12, 13], omegas=[0.7, 2.0, 4.5], modes=[0, 1, 2, 3, 4]),
    "single_cylinder": dict(case="single_cylinder", NMK=[9, 11], omegas=[0.4, 1.5, 6.0], modes=[0]),
    "overhang_2body": dict(case="overhang_bicylinder", NMK=[6, 7, 8], omegas=[0.4, 1.3, 4.0], modes=[0, 1]),
    "C5_small_bicylinder_n75": dict(h=1.001, a=[0.5, 1.0], d=[0.5, 0.25], NMK=[75] * 3, body_map=[0, 0],
                                    omegas=list(np.linspace(0.5, 3.0, 8)[[0, 7]]), modes=[0]),
}


def gen_sweeps():
    out = {}
    meta = {}
    for name, s in SWEEPS.items():
        c = dict(CASES[s["case"]]) if "case" in s else {k: s[k] for k in ("h", "a", "d", "body_map")}
        c["NMK"] = s["NMK"]
        nb = max(c["body_map"]) + 1
        prob = make_problem(c, [False] * nb)
        for bidx, body in enumerate(prob.geometry.body_arrangement.bodies):
            body.heaving = bidx in s["modes"]
        prob.set_frequencies(np.asarray(s["omegas"], float))
        eng = of.MEEMEngine([prob])
        res = quiet(eng.run_and_store_results, 0)
        ds = res.dataset
        for var in ("added_mass", "damping", "excitation_force", "excitation_phase"):
            out[f"{name}/{var}"] = np.asarray(ds[var].values)
        pot = np.asarray(ds["potentials_real"].values) + 1j * np.asarray(ds["potentials_imag"].values)
        out[f"{name}/potentials"] = pot
        out[f"{name}/c_vector"] = np.asarray(ds["c_vector"].values)        # (frequency, mode_i, p) complex, eng.py:646-652
        out[f"{name}/m0"] = np.array([of.wavenumber(w, c["h"]) for w in s["omegas"]])
        meta[name] = dict(h=c["h"], a=list(map(float, c["a"])), d=list(map(float, c["d"])), NMK=list(c["NMK"]),
                          body_map=list(c["body_map"]), omegas=list(map(float, s["omegas"])), modes=list(s["modes"]))
        print(name, "added_mass", out[f"{name}/added_mass"].ravel()[:2])
    out["__meta__"] = np.array(json.dumps(meta))
    np.savez_compressed(os.path.join(HERE, "ref_sweeps.npz"), **out)


def gen_debug_csv():
    import pandas as pd
    root = "/root/reference/debug_results"
    specs = {   # package/test/test_hydrocoeffs.py:30-35
        "mini_bicylinder": dict(h=1.001, d=[0.25, 0.125], a=[0.125, 0.25]),
        "small_bicylinder": dict(h=1.001, d=[0.5, 0.25], a=[0.5, 1.0]),
        "big_bicylinder": dict(h=1.001, d=[0.75, 0.5], a=[0.5, 0.75]),
        "mini_tricylinder": dict(h=2.001, d=[1.0, 0.5, 0.25], a=[0.25, 0.5, 1.0]),
        "small_tricylinder": dict(h=20.0, d=[15, 10, 5], a=[5, 10, 15]),
        "big_tricylinder": dict(h=25.0, d=[20, 15, 10], a=[10, 15, 20]),
    }
    out = {}
    for name, sp in specs.items():
        df = pd.read_csv(os.path.join(root, f"DEBUG_{name}_regenerated_v2.csv"))
        out[name] = dict(sp, NMK=50, m0=[float(x) for x in df["m0"]], of_nondim=[float(x) for x in df["of_nondim"]])
    with open(os.path.join(HERE, "debug_nondim_added_mass.json"), "w") as f:
        json.dump(out, f, indent=1)
    gold = np.load("/root/reference/package/test/data/gold_standard_A_2region.npy")
    np.save(os.path.join(HERE, "gold_standard_A_2region.npy"), gold)


if __name__ == "__main__":
    gen_debug_csv()
    gen_single()
    gen_sweeps()
