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


def run_compiled(g, omegas, nthreads=1):
    """oracle/meem_cpu.cpp: the reference's dense algorithm compiled (g++ build of the shared closed-form headers, LAPACK
    zgetrf/zgetrs out of scipy), `nthreads` std::threads over the frequencies.  Returns the added-mass array."""
    from oracle import meem_cpu
    return meem_cpu.run_sweep(g["h"], g["a"], g["d"], g["NMK"], g["body_map"], g["modes"], omegas, nthreads=nthreads)["added_mass"]


def time_legs(geoms, omegas, cores, budget_s=4.0):
    """The other CPU legs of BASELINE.md section 4, each on a small bounded sample (~budget_s seconds):
    (1) the untouched reference in ONE process with the BLAS library's default thread count,
    (3) the compiled restatement with 1 thread and with `cores` threads (BLAS threads = 1).  systems/s each."""
    import time
    legs = {}

    def sample(n_sys):
        if len(geoms) == 1:
            idx = np.linspace(0, len(omegas) - 1, min(n_sys, len(omegas))).astype(int)
            return [(geoms[0], [float(omegas[i]) for i in idx])]
        ng = max(1, min(len(geoms), n_sys // len(omegas)))
        return [(g, [float(w) for w in omegas]) for g in geoms[:ng]]

    def timed(fn, n0):
        n, done, t0 = n0, 0, time.perf_counter()
        while True:
            for g, om in sample(n):
                fn(g, om)
                done += len(om)
            el = time.perf_counter() - t0
            if el >= budget_s or n >= 65536:
                return done / el, done, el
            n *= 4

    try:
        from threadpoolctl import threadpool_info, threadpool_limits
        blas_threads = max([int(i.get("num_threads", 1)) for i in threadpool_info() if i.get("user_api") == "blas"] or [1])
    except Exception:
        threadpool_limits, blas_threads = None, None
    if reference_available():
        _load_reference()
        v, n, t = timed(run_reference, 2)
        legs["reference_one_process"] = {"value": v, "blas_threads": blas_threads, "sample": f"{n} systems, {t:.1f} s, MEEMEngine.run_and_store_results"}
    from oracle import meem_cpu
    if meem_cpu.available():
        ctx = threadpool_limits(limits=1) if threadpool_limits else contextlib.nullcontext()
        with ctx:
            v, n, t = timed(lambda g, om: run_compiled(g, om, 1), 2)
            legs["compiled_1_thread"] = {"value": v, "sample": f"{n} systems, {t:.1f} s, oracle/meem_cpu.cpp (dense assembly + LAPACK zgetrf/zgetrs, the reference's algorithm)"}
            v, n, t = timed(lambda g, om: run_compiled(g, om, cores), 2 * cores)
            legs["compiled_all_cores"] = {"value": v, "threads": cores, "sample": f"{n} systems, {t:.1f} s, same, {cores} threads over the frequencies"}
    return legs


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
