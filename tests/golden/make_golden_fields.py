"""Golden fixtures for the field evaluation, from the UNTOUCHED reference (build container only).

    python tests/golden/make_golden_fields.py   ->   tests/golden/ref_fields.npz

For each case: the reference solves for X (solve_linear_system_multi), then evaluates
MEEMEngine.calculate_potentials (meem_engine.py:395-467) and calculate_velocities (:478-553) on its own (R, Z) grid.
Stored: X, the grid, phiH / phiP / phi, vrH / vzH / vrP / vzP / vr / vz (NaN where the reference leaves NaN).
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
    "small_bicylinder": dict(h=1.001, a=[0.5, 1.0], d=[0.5, 0.25], NMK=[12, 13, 14], body_map=[0, 1], heaving=[1, 0],
                             m0=1.0, res=9, sharp=True),
    "tricylinder_deep": dict(h=20.0, a=[5, 10, 15], d=[15, 10, 5], NMK=[10, 11, 12, 13], body_map=[0, 1, 1], heaving=[0, 1],
                             m0=1.3, res=8, sharp=False),     # m0 h = 26 >= 14: asymptotic exterior branch
    "overhang_bicylinder": dict(h=1.001, a=[0.5, 1.0], d=[0.25, 0.5], NMK=[6, 7, 8], body_map=[0, 1], heaving=[1, 1],
                                m0=2.0, res=7, sharp=True),
    "single_cylinder": dict(h=3.0, a=[1.0], d=[0.7], NMK=[9, 11], body_map=[0], heaving=[1], m0=0.4, res=8, sharp=True),
}


def main():
    out, meta = {}, {}
    for name, c in CASES.items():
        nb = max(c["body_map"]) + 1
        geo = of.BasicRegionGeometry.from_vectors(a=np.asarray(c["a"], float), d=np.asarray(c["d"], float), h=c["h"],
                                                  NMK=list(c["NMK"]), body_map=list(c["body_map"]),
                                                  heaving_map=[False] * nb)
        prob = of.MEEMProblem(geo)
        for bi, body in enumerate(prob.geometry.body_arrangement.bodies):
            body.heaving = bool(c["heaving"][bi])
        # domains carry their own flags (meem_engine.py:110-116 reads them)
        region_body = [bi for bi, b in enumerate(prob.geometry.body_arrangement.bodies) for _ in range(len(b.a))]
        for k in sorted(prob.domain_list.keys())[:-1]:
            prob.domain_list[k].heaving = bool(c["heaving"][region_body[k]])
        eng = of.MEEMEngine([prob])
        with contextlib.redirect_stdout(io.StringIO()):
            X = eng.solve_linear_system_multi(prob, c["m0"])
            pot = eng.calculate_potentials(prob, X, c["m0"], c["res"], c["sharp"])
            vel = eng.calculate_velocities(prob, X, c["m0"], c["res"], c["sharp"])
        out[f"{name}/X"] = X
        out[f"{name}/R"] = pot["R"]
        out[f"{name}/Z"] = pot["Z"]
        for k in ("phiH", "phiP", "phi"):
            out[f"{name}/{k}"] = pot[k]
        assert np.array_equal(vel["R"], pot["R"]) and np.array_equal(vel["Z"], pot["Z"])
        for k in ("vrH", "vzH", "vrP", "vzP", "vr", "vz"):
            out[f"{name}/{k}"] = vel[k]
        heav_regions = [int(prob.domain_list[k].heaving) for k in sorted(prob.domain_list.keys())[:-1]]
        meta[name] = dict(h=c["h"], a=list(map(float, c["a"])), d=list(map(float, c["d"])), NMK=list(c["NMK"]),
                          body_map=list(c["body_map"]), heaving_bodies=list(c["heaving"]), heaving_regions=heav_regions,
                          body_heaving=[int(x) for x in prob.geometry.body_arrangement.heaving],
                          m0=c["m0"], res=c["res"], sharp=c["sharp"])
        print(name, pot["R"].shape, "nan in phi:", int(np.isnan(pot["phi"]).sum()), "max|phi|", np.nanmax(np.abs(pot["phi"])))
    out["__meta__"] = np.array(json.dumps(meta))
    np.savez_compressed(os.path.join(HERE, "ref_fields.npz"), **out)


if __name__ == "__main__":
    main()
