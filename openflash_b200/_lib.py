"""ctypes binding of libmeemb200.so (include/meemb200.h).

There is no CPU fallback: importing this module without the built library, or creating a context
without a B200, raises.  Build with ``python -c "import __graft_entry__ as g; g.build()"`` or
``python -m openflash_b200.build``.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libmeemb200.so")

ALGO_SCHUR, ALGO_DENSE, ALGO_SCHUR_LU = 0, 1, 2
ROOTS_REFERENCE, ROOTS_CONVERGED = 0, 1
FREQ_OMEGA, FREQ_M0 = 0, 1
STATIC_WS_BYTES = 32 << 20      # MEEMB200_STATIC_WS_BYTES (include/meemb200.h)


class MeemBatch(C.Structure):
    _fields_ = [
        ("n_geom", C.c_int32), ("n_freq", C.c_int32),
        ("reg_off", C.c_void_p), ("h", C.c_void_p), ("a", C.c_void_p), ("d", C.c_void_p),
        ("nmk", C.c_void_p), ("n_modes", C.c_void_p), ("flag_off", C.c_void_p), ("heaving", C.c_void_p),
        ("freq", C.c_void_p),
        ("freq_kind", C.c_int32), ("algo", C.c_int32), ("root_mode", C.c_int32), ("want_x", C.c_int32),
        ("rho", C.c_double),
    ]


class MeemPlanInfo(C.Structure):
    _fields_ = [
        ("input_bytes", C.c_size_t), ("workspace_bytes", C.c_size_t),
        ("n_rr", C.c_int64), ("n_r", C.c_int64), ("n_sys", C.c_int64), ("n_x", C.c_int64),
        ("n_launches", C.c_int64), ("flops", C.c_double), ("n_c", C.c_int64),
        ("geom_kernel", C.c_int32), ("freq_kernel", C.c_int32),
    ]


class MeemOutputs(C.Structure):
    _fields_ = [
        ("added_mass", C.c_void_p), ("damping", C.c_void_p), ("exc_force", C.c_void_p), ("exc_phase", C.c_void_p),
        ("m0", C.c_void_p), ("status", C.c_void_p), ("x", C.c_void_p), ("c_vector", C.c_void_p),
    ]


class MeemError(RuntimeError):
    pass


_lib = None

EXPORTS = [
    "meemb200_version", "meemb200_last_error", "meemb200_create", "meemb200_destroy", "meemb200_sm_count",
    "meemb200_plan_create", "meemb200_plan_destroy", "meemb200_plan_info", "meemb200_plan_blob",
    "meemb200_plan_offsets", "meemb200_plan_set_freq", "meemb200_plan_set_test_hook", "meemb200_solve", "meemb200_solve_host", "meemb200_solve_static_ws",
    "meemb200_assemble", "meemb200_imk_table", "meemb200_coefficients", "meemb200_fields", "meemb200_roots", "meemb200_bessel", "meemb200_fp64_peak", "meemb200_last_timing", "meemb200_phase_profile",
    "meemb200_debug_layout",
]


def load():
    """Load the shared library (raises if it has not been built — no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MeemError(
            f"{LIB_PATH} is missing: build the CUDA library first (python -m openflash_b200.build). "
            "openflash_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    lib.meemb200_version.restype = C.c_int
    lib.meemb200_last_error.restype = C.c_char_p
    lib.meemb200_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.meemb200_destroy.argtypes = [vp]
    lib.meemb200_sm_count.argtypes = [vp]
    lib.meemb200_plan_create.argtypes = [vp, C.POINTER(MeemBatch), C.POINTER(vp)]
    lib.meemb200_plan_destroy.argtypes = [vp]
    lib.meemb200_plan_info.argtypes = [vp, C.POINTER(MeemPlanInfo)]
    lib.meemb200_plan_blob.argtypes = [vp]
    lib.meemb200_plan_blob.restype = vp
    lib.meemb200_plan_offsets.argtypes = [vp, C.c_int32, C.POINTER(C.c_int64)]
    lib.meemb200_plan_set_freq.argtypes = [vp, vp]
    lib.meemb200_plan_set_test_hook.argtypes = [vp, C.c_int32]
    lib.meemb200_solve.argtypes = [vp, vp, vp, C.c_size_t, C.POINTER(MeemOutputs), vp]
    lib.meemb200_solve_host.argtypes = [vp, C.POINTER(MeemOutputs), vp]
    lib.meemb200_solve_static_ws.argtypes = [vp, vp, C.POINTER(MeemOutputs), vp]
    lib.meemb200_assemble.argtypes = [vp, vp, vp, C.c_size_t, C.c_int32, C.c_int32, vp, vp, vp, vp]
    lib.meemb200_imk_table.argtypes = [vp, vp, vp, C.c_size_t, C.c_int32, C.c_int32, vp, vp]
    lib.meemb200_coefficients.argtypes = [vp, vp, vp, C.c_size_t, C.c_int32, C.c_int32, vp, vp, vp]
    lib.meemb200_fields.argtypes = [vp, vp, vp, C.c_size_t, C.c_int32, C.c_int32, C.c_int32, vp, C.c_int64, vp, vp, vp, vp]
    lib.meemb200_roots.argtypes = [vp, C.c_int32, C.c_int32, vp, vp, C.c_int32, C.c_int32, vp, vp, vp, vp]
    lib.meemb200_bessel.argtypes = [vp, C.c_int32, C.c_int64, vp, vp, vp]
    lib.meemb200_fp64_peak.argtypes = [vp, C.c_int32, C.POINTER(C.c_double), vp]
    lib.meemb200_last_timing.argtypes = [vp, C.POINTER(C.c_float)]
    lib.meemb200_phase_profile.argtypes = [vp, vp, C.POINTER(C.c_double), vp]
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise MeemError(f"meemb200 error {rc}: {load().meemb200_last_error().decode()}")
