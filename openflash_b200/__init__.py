"""openflash_b200 — B200-native (sm_100a) implementation of OpenFLASH's MEEM hot path.

Keeps the reference's Geometry / MEEMProblem / MEEMEngine API (package/src/openflash/__init__.py);
the numerics run in hand-written CUDA kernels behind the C ABI of include/meemb200.h.
"""
from .multi_constants import g, rho  # noqa: F401
from .model import (Body, SteppedBody, CoordinateBody, BodyArrangement, ConcentricBodyGroup, Domain,  # noqa: F401
                    Geometry, BasicRegionGeometry, MEEMProblem)
from .results import Results  # noqa: F401
from .batch import (BatchPlan, FlatGeometry, GeometryBatch, solve_batch, device_roots, device_bessel, fp64_peak,  # noqa: F401
                    ALGO_SCHUR, ALGO_DENSE, ALGO_SCHUR_LU, ROOTS_REFERENCE, ROOTS_CONVERGED, FREQ_OMEGA, FREQ_M0)
from .meem_engine import MEEMEngine, flatten_problem  # noqa: F401
from .distributed import ShardedPlan, solve_batch_distributed  # noqa: F401
