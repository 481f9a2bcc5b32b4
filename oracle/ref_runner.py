"""Drive the staged reference (oracle/_ref, see build_ref.py) or, when it is absent, the oracle port on a bounded sample
of a benchmark workload.  Test / benchmark infrastructure: only bench.py's CPU legs and tests/ may import this."""
import contextlib
import io
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def reference_available():
    return os.path.isfile(os.path.join(REF_DIR, "openflash", "meem_engine.py"))


def _load_reference():
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    import openflash
    return openflash


def run_reference(g, omegas):
    """One geometry dict x omegas through the reference's public API: BasicRegionGeometry.from_vectors -> MEEMProblem ->
    MEEMEngine.run_and_store_results (meem_engine.py:555-690).  Returns the added-mass array (frequency, mode, mode)."""
    of = _load_reference()
    nb = max(g["body_map"]) + 1
    with contextlib.redirect_stdout(io.StringIO()), warnings.catch_warnings():
        warnings.simplefilter("ignore")
        geo = of.BasicRegionGeometry.from_vectors(a=np.asarray(g["a"], float), d=np.asarray(g["d"], float), h=g["h"],
                                                  NMK=list(g["NMK"]), body_map=list(g["body_map"]), heaving_map=[False] * nb)
        prob = of.MEEMProblem(geo)
        for bi, body in enumerate(prob.geometry.body_arrangement.bodies):
            body.heaving = bi in g["modes"]
        prob.set_frequencies(np.asarray(omegas, float))
        res = of.MEEMEngine([prob]).run_and_store_results(0)
        return np.asarray(res.dataset["added_mass"].values)


def run_port(g, omegas):
    from oracle import meem_oracle as mo
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        og = mo.FlatGeometry(g["h"], g["a"], g["d"], g["NMK"])
        return mo.run_sweep(og, omegas, g["body_map"], g["modes"])["added_mass"]


def cpu_chunk(args):
    """Worker of the multiprocessing pool: (kind, geometry dict, omegas) with one BLAS thread per process."""
    kind, g, omegas = args
    try:
        from threadpoolctl import threadpool_limits
        ctx = threadpool_limits(limits=1)
    except Exception:
        ctx = contextlib.nullcontext()
    with ctx:
        am = run_reference(g, omegas) if kind == "reference" else run_port(g, omegas)
    return float(np.nansum(am))
