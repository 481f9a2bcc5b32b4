"""meemb200_solve_static_ws: the allocation-free variant of meemb200_solve whose workspace is the library's file-scope
static __device__ array (include/meemb200.h; no counterpart in the reference).

Called through the C-ABI (raw ctypes on the loaded library) and checked against the existing entry point meemb200_solve
with a caller-owned workspace, on consecutive calls that share the static array.  Collected last on purpose (file name):
a failure here must not hide the parity tests under `pytest -x`.  Passes alone in a fresh process and after the others.
"""
import ctypes as C

import numpy as np
import pytest

from helpers import load_npz, relerr

pytestmark = pytest.mark.gpu

Z2, META2 = load_npz("ref_sweeps.npz")
TOL = 1e-12      # two runs of the same kernels on the same inputs; the parity tests proper use 1e-10 against the oracle


@pytest.fixture(scope="module")
def ofb():
    import openflash_b200 as ofb
    return ofb


def _plans(ofb, which):
    if which == "golden_R2_x_c":        # golden sweep with two radiating bodies, X and c vector wanted
        name = [k for k in sorted(META2) if len(META2[k]["modes"]) == 2][0]
        s = META2[name]
        geo = ofb.FlatGeometry.from_body_map(s["h"], s["a"], s["d"], s["NMK"], s["body_map"], s["modes"])
        return ofb.BatchPlan([geo, geo], s["omegas"], want_x=True, want_c=True)
    if which == "c4_mixed":             # heterogeneous design-sweep batch: window elimination + warp-per-system solve
        import bench
        gs = bench.c4_geometries(60)
        geos = [ofb.FlatGeometry.from_body_map(g["h"], g["a"], g["d"], g["NMK"], g["body_map"], g["modes"]) for g in gs]
        return ofb.BatchPlan(geos, np.array([0.3, 1.1, 2.5]))
    if which == "small_tricylinder":    # 50 modes / region: CTA-per-system frequency kernels
        geo = ofb.FlatGeometry.from_body_map(20.0, [5.0, 10.0, 15.0], [15.0, 10.0, 5.0], [50] * 4, [0, 0, 0], [0])
        return ofb.BatchPlan([geo], np.linspace(0.2, 3.0, 12))
    raise KeyError(which)


def _outputs_on(plan, d_out):
    from openflash_b200 import _lib
    base = d_out.data_ptr()
    ptr = lambda name: (base + plan._seg[name][0]) if name in plan._seg else None  # noqa: E731
    return _lib.MeemOutputs(ptr("added_mass"), ptr("damping"), ptr("excitation_force"), ptr("excitation_phase"),
                            ptr("m0"), ptr("status"), ptr("X"), ptr("c_vector"))


def _solve_static(plan, stream=None):
    """One call of the new entry point into a fresh, poisoned output buffer; returns (rc, flat numpy outputs)."""
    from openflash_b200.batch import unpack_outputs
    t = plan.torch
    if not plan._uploaded:
        plan.upload()
        t.cuda.synchronize()
    d_out = t.full((plan._out_bytes,), 0xA5, dtype=t.uint8, device=plan.device)
    out = _outputs_on(plan, d_out)
    t.cuda.synchronize()
    s = t.cuda.current_stream(plan.device) if stream is None else stream
    rc = plan.lib.meemb200_solve_static_ws(plan.plan, plan.d_blob.data_ptr(), C.byref(out), C.c_void_p(s.cuda_stream))
    t.cuda.synchronize()
    return rc, unpack_outputs(d_out.cpu().numpy(), plan._seg, plan.G, plan.F, plan.want_c, plan.want_x)


def _solve_existing(plan):
    plan.upload()
    plan.solve()
    return plan.download()


def _same(got, ref):
    assert set(got) == set(ref)
    assert np.array_equal(got["status"], ref["status"])
    for k in ref:
        if k == "status":
            continue
        g, r = np.asarray(got[k]), np.asarray(ref[k])
        assert g.shape == r.shape, k
        ok = np.isfinite(r)
        assert np.array_equal(np.isfinite(g), ok), k
        if ok.any():
            assert relerr(g[ok], r[ok]) <= TOL, (k, relerr(g[ok], r[ok]))


@pytest.mark.parametrize("which", ["golden_R2_x_c", "c4_mixed", "small_tricylinder"])
def test_static_workspace_solve_matches_solve_two_calls_in_a_row(ofb, which):
    from openflash_b200 import _lib
    plan = _plans(ofb, which)
    assert int(plan.info.workspace_bytes) <= _lib.STATIC_WS_BYTES
    ref = _solve_existing(plan)
    assert np.all(ref["status"] == 0)
    for call in range(2):
        rc, got = _solve_static(plan)
        assert rc == 0, plan.lib.meemb200_last_error()
        _same(got, ref)
    # a third call on a side stream (the array is shared: the calls above have finished, so no overlap)
    side = plan.torch.cuda.Stream(device=plan.device)
    rc, got = _solve_static(plan, stream=side)
    assert rc == 0, plan.lib.meemb200_last_error()
    _same(got, ref)


def test_static_workspace_shared_between_plans_keeps_no_state(ofb):
    """Plans of different layouts alternate on the one array: each call must be independent of what the last one left."""
    names = ["golden_R2_x_c", "c4_mixed", "small_tricylinder"]
    plans = [_plans(ofb, n) for n in names]
    refs = [_solve_existing(p) for p in plans]
    for rnd in range(2):
        for p, ref in zip(plans, refs):
            rc, got = _solve_static(p)
            assert rc == 0, p.lib.meemb200_last_error()
            _same(got, ref)


def test_static_workspace_golden_values(ofb):
    """The new entry point against the untouched reference's golden sweep directly (not only against meemb200_solve)."""
    name = [k for k in sorted(META2) if len(META2[k]["modes"]) == 2][0]
    plan = _plans(ofb, "golden_R2_x_c")
    rc, got = _solve_static(plan)
    assert rc == 0, plan.lib.meemb200_last_error()
    res = plan.split(got)
    for r in res:
        for v in ("added_mass", "damping", "excitation_force"):
            assert relerr(r[v], Z2[f"{name}/{v}"]) <= 1e-10, v


def test_static_workspace_rejects_plans_beyond_the_maximum(ofb):
    from openflash_b200 import _lib
    lib = _lib.load()
    geo = ofb.FlatGeometry.from_body_map(25.0, [10.0, 15.0, 20.0], [20.0, 15.0, 10.0], [200] * 4, [0, 0, 0], [0])
    plan = ofb.BatchPlan([geo], np.linspace(0.1, 3.0, 300))
    assert int(plan.info.workspace_bytes) > _lib.STATIC_WS_BYTES
    rc, got = _solve_static(plan)
    assert rc == -3                                            # MEEMB200_ENOMEM, nothing enqueued
    msg = lib.meemb200_last_error().decode()
    assert str(int(plan.info.workspace_bytes)) in msg and str(_lib.STATIC_WS_BYTES) in msg and "MEEMB200_STATIC_WS_BYTES" in msg
    assert np.all(got["status"].view(np.uint8) == 0xA5)       # outputs untouched
    small = _plans(ofb, "small_tricylinder")
    out = _outputs_on(small, small.d_out)
    assert lib.meemb200_solve_static_ws(None, small.d_blob.data_ptr(), C.byref(out), None) == -1
    assert lib.meemb200_solve_static_ws(small.plan, None, C.byref(out), None) == -1
    assert lib.meemb200_solve_static_ws(small.plan, small.d_blob.data_ptr(), None, None) == -1
    # and the array still serves the next caller
    ref = _solve_existing(small)
    rc, got = _solve_static(small)
    assert rc == 0, lib.meemb200_last_error()
    _same(got, ref)
