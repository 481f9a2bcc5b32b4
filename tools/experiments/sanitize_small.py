"""Tiny invocation of every kernel family for compute-sanitizer (memcheck / racecheck); results are checked against each
other loosely - the parity tests do the real checking.
    compute-sanitizer --tool memcheck python tools/experiments/sanitize_small.py"""
import os
import sys
import numpy as np
sys.path.insert(0, '.')
import bench
import openflash_b200 as ofb

om = np.array([0.5, 1.7])


def run(tag, geos, env=None, **kw):
    env = env or {}
    for k, v in env.items():
        os.environ[k] = v
    try:
        plan = ofb.BatchPlan(geos, om, **kw)
        res = plan.run()
    finally:
        for k in env:
            del os.environ[k]
    bad = sum(int(np.count_nonzero(r["status"])) for r in res)
    print(f"{tag:34s} {plan.geom_kernel:24s} {plan.freq_kernel:24s} am0 {res[0]['added_mass'].ravel()[0]:.6e} bad {bad}", flush=True)
    return res


c4 = [ofb.FlatGeometry.from_body_map(g["h"], g["a"], g["d"], g["NMK"], g["body_map"], g["modes"]) for g in bench.c4_geometries(8)]
tri = ofb.FlatGeometry.from_body_map(20.0, [5.0, 10.0, 15.0], [15.0, 10.0, 5.0], [20] * 4, [0, 1, 1], [0, 1])
over = ofb.FlatGeometry.from_body_map(1.001, [0.5, 1.0], [0.25, 0.5], [6, 7, 8], [0, 1], [0, 1])
run("c4 band<smem,128> + small", c4)
run("c4 band<smem,256>", c4, {"MEEMB200_BAND_VARIANT": "1"})
run("c4 band<global>", c4, {"MEEMB200_BAND_VARIANT": "2"})
run("c4 dense k_geom", c4, {"MEEMB200_NO_BAND": "1"})
run("c4 no small", c4, {"MEEMB200_NO_SMALL": "1"})
run("tri coop2", [tri], {"MEEMB200_FORCE_COOP": "1"})
run("tri coop (dense)", [tri], {"MEEMB200_FORCE_COOP": "1", "MEEMB200_NO_COOP2": "1"})
run("overhang coop fallback", [over], {"MEEMB200_FORCE_COOP": "1"})
run("tri want_x + c", [tri, over], want_x=True, want_c=True)
run("tri schur_lu", [tri], algo=ofb.ALGO_SCHUR_LU)
run("tri dense algo", [tri], algo=ofb.ALGO_DENSE)
run("tri no hchol", [tri], {"MEEMB200_NO_HCHOL": "1"})
p = ofb.BatchPlan([tri], om)
p.set_test_hook(2)
print("hook", p.run()[0]["status"], flush=True)
print("imk", p.imk_table(0, 0).shape, "coef", p.coefficients(np.ones(tri.size, dtype=complex), 0, 0).shape, flush=True)
# chunk-stream k_freq_gemm (v19): 72 modes -> 9 column blocks (one extra block per pass), 104 -> 13 (five extras), 16- and 8-row chunks
for nm in (72, 104):
    big = ofb.FlatGeometry.from_body_map(25.0, [10.0, 15.0, 20.0], [20.0, 15.0, 10.0], [nm] * 4, [0, 1, 2], [0, 2])
    run(f"tri {nm} modes chunk stream", [big])
    run(f"tri {nm} modes classic gemm", [big], {"MEEMB200_NO_STREAM": "1", "MEEMB200_NO_STREAM_SYRK": "1", "MEEMB200_NO_BATCH_INV": "1"})
