"""ctypes driver of oracle/meem_cpu.cpp, the compiled CPU restatement of the reference's dense algorithm.

TEST / BENCHMARK INFRASTRUCTURE, not product: only tests/ and bench.py's cpu_baseline legs import this (nothing under
openflash_b200/ does).  LAPACK is scipy's own: zgetrf / zgetrs are taken out of scipy.linalg.cython_lapack's capsules,
the same routines scipy.linalg.lu_factor / lu_solve reach (reference call sites: meem_engine.py:609, :633).
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "meem_cpu.cpp")
LIB = os.path.join(HERE, "_build", "libmeem_cpu.so")
RHO = 1023.0

_lib = None
_lapack = None


def build(force=False):
    """g++ build into oracle/_build/ (git-ignored, travels to the GPU box with the snapshot)."""
    deps = [SRC, os.path.join(HERE, "..", "openflash_b200", "csrc", "meem_forms.cuh"), os.path.join(HERE, "..", "openflash_b200", "csrc", "meem_math.cuh")]
    if not force and os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(d) for d in deps if os.path.exists(d)):
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-pthread", "-o", LIB, SRC])
    return LIB


def available():
    return os.path.exists(LIB)


def _capsule_pointer(name):
    from scipy.linalg import cython_lapack
    cap = cython_lapack.__pyx_capi__[name]
    C.pythonapi.PyCapsule_GetName.restype = C.c_char_p
    C.pythonapi.PyCapsule_GetName.argtypes = [C.py_object]
    C.pythonapi.PyCapsule_GetPointer.restype = C.c_void_p
    C.pythonapi.PyCapsule_GetPointer.argtypes = [C.py_object, C.c_char_p]
    return C.pythonapi.PyCapsule_GetPointer(cap, C.pythonapi.PyCapsule_GetName(cap))


def load():
    global _lib, _lapack
    if _lib is None:
        if not available():
            build()
        lib = C.CDLL(LIB)
        vp = C.c_void_p
        lib.meem_cpu_sweep.argtypes = [C.c_int, C.c_double, vp, vp, vp, C.c_int, vp, C.c_int, vp, C.c_double, C.c_int, vp, vp,
                                       vp, vp, vp, vp, vp, vp]
        lib.meem_cpu_sweep.restype = C.c_int
        _lapack = (_capsule_pointer("zgetrf"), _capsule_pointer("zgetrs"))
        _lib = lib
    return _lib


def run_sweep(h, a, d, NMK, region_body, modes, omegas, rho=RHO, nthreads=1):
    """Same arguments and outputs as oracle.meem_oracle.run_sweep (reference index order [f, j, i])."""
    lib = load()
    M = len(a)
    region_body = list(region_body)
    flags = np.zeros((len(modes), M + 1), dtype=np.uint8)
    for j, m in enumerate(modes):
        flags[j, :M] = [1 if region_body[i] == m else 0 for i in range(M)]
    R, om = len(modes), np.ascontiguousarray(np.asarray(omegas, dtype=np.float64))
    F = len(om)
    aa = np.array(list(a) + [np.inf], dtype=np.float64)
    dd = np.array(list(d) + [0.0], dtype=np.float64)
    nmk = np.array(NMK, dtype=np.int32)
    out = dict(added_mass=np.empty((F, R, R)), damping=np.empty((F, R, R)), excitation_force=np.empty((F, R)),
               excitation_phase=np.empty((F, R)), m0=np.empty(F), status=np.empty(F, dtype=np.int32))
    rc = lib.meem_cpu_sweep(M, float(h), aa.ctypes.data, dd.ctypes.data, nmk.ctypes.data, R, flags.ctypes.data, F, om.ctypes.data,
                            float(rho), int(nthreads), _lapack[0], _lapack[1], out["added_mass"].ctypes.data, out["damping"].ctypes.data,
                            out["excitation_force"].ctypes.data, out["excitation_phase"].ctypes.data, out["m0"].ctypes.data,
                            out["status"].ctypes.data)
    if rc < 0:
        raise ValueError("meem_cpu_sweep: bad argument")
    return out
