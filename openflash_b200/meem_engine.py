"""MEEMEngine with the reference's method names (meem_engine.py:23-690); numerics on the B200.

Every numeric call goes through ``BatchPlan`` -> libmeemb200.so.  There is no CPU solve path:
without the CUDA library or a B200 these methods raise.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np

from . import batch as _batch
from .batch import ALGO_DENSE, ALGO_SCHUR, FREQ_M0, FREQ_OMEGA, BatchPlan, FlatGeometry
from .model import MEEMProblem, SteppedBody
from .multi_constants import g as G_CONST, rho as RHO_DEFAULT
from .results import Results


class ProblemCacheView:
    """What tests read from the reference's ProblemCache (problem_cache.py:7-91): the most recent
    ``m_k_arr`` / ``N_k_arr`` and the wavenumber they belong to."""

    def __init__(self):
        self.m_k_arr = None
        self.N_k_arr = None
        self.cached_m0 = None


def flatten_problem(problem: MEEMProblem, modes: Optional[Sequence[int]] = None, from_domain_flags: bool = False) -> FlatGeometry:
    """Domains -> flat arrays (meem_engine.py:110-116).

    ``modes`` lists heaving-body indices, each becoming one radiating mode that moves all regions of
    that body (meem_engine.py:350-367, :618-627).  With ``from_domain_flags`` a single mode is built
    from the domains' current ``heaving`` flags, which is what assemble_b_multi reads.
    """
    doms = [problem.domain_list[k] for k in sorted(problem.domain_list.keys())]
    interior = doms[:-1]
    h = doms[0].h
    a = [dm.a for dm in interior]
    d = [dm.di for dm in interior]
    NMK = [dm.number_harmonics for dm in doms]
    if from_domain_flags:
        flags = [[1 if dm.heaving else 0 for dm in interior]]
    else:
        bodies = problem.geometry.body_arrangement.bodies
        for b in bodies:
            if not isinstance(b, SteppedBody):
                raise TypeError("run_and_store_results only supports SteppedBody.")
        region_body = [bi for bi, b in enumerate(bodies) for _ in range(len(b.a))]
        if modes is None:
            modes = list(range(len(bodies)))
        flags = [[1 if region_body[i] == m else 0 for i in range(len(interior))] for m in modes]
    return FlatGeometry(h, a, d, NMK, flags)


class MEEMEngine:
    def __init__(self, problem_list: List[MEEMProblem], device: int = 0, algo: int = ALGO_SCHUR):
        self.problem_list = problem_list
        self.device = device
        self.algo = algo
        self.cache_list: Dict[MEEMProblem, ProblemCacheView] = {p: ProblemCacheView() for p in problem_list}

    # ------------------------------------------------------------------ reference-named pieces
    def _cache(self, problem):
        if problem not in self.cache_list:
            self.cache_list[problem] = ProblemCacheView()
        return self.cache_list[problem]

    def _ensure_m_k_and_N_k_arrays(self, problem: MEEMProblem, m0):
        """meem_engine.py:33-42 — idempotent per (problem, m0)."""
        cache = self._cache(problem)
        if cache.m_k_arr is None or cache.cached_m0 != m0:
            h = problem.domain_list[0].h
            K = problem.domain_list[max(problem.domain_list.keys())].number_harmonics
            _, mk, nk = _batch.device_roots([m0], h, K, freq_kind=FREQ_M0, device=self.device)
            cache.m_k_arr, cache.N_k_arr, cache.cached_m0 = mk[0], nk[0], m0
        return cache

    def update_forcing(self, problem: MEEMProblem):
        """The reference refreshes cached b-vector pieces here (problem_cache.py:93-153); the CUDA path
        rebuilds b from the flags of every call, so nothing is cached on the host."""
        return None

    def _single_plan(self, problem, m0, want_x=False, algo=None):
        geo = flatten_problem(problem, from_domain_flags=True)
        return BatchPlan([geo], [m0], freq_kind=FREQ_M0, want_x=want_x, device=self.device,
                         algo=self.algo if algo is None else algo)

    def assemble_A_multi(self, problem: MEEMProblem, m0) -> np.ndarray:
        """(N_T, N_T) complex system matrix (meem_engine.py:44-55)."""
        self._ensure_m_k_and_N_k_arrays(problem, m0)
        return self._single_plan(problem, m0).assemble(0, 0)[0]

    def assemble_b_multi(self, problem: MEEMProblem, m0) -> np.ndarray:
        """(N_T,) right-hand side for the domains' current heaving flags (meem_engine.py:57-63)."""
        self._ensure_m_k_and_N_k_arrays(problem, m0)
        return self._single_plan(problem, m0).assemble(0, 0)[1][0]

    def assemble_c_multi(self, problem: MEEMProblem, heaving: Sequence[int]) -> np.ndarray:
        """c vector of length p for the given region flags (meem_engine.py:65-98)."""
        geo = flatten_problem(problem, from_domain_flags=True)
        M = geo.M
        geo.mode_flags = np.asarray([list(heaving)[:M]], dtype=np.uint8)
        plan = BatchPlan([geo], [1.0], freq_kind=FREQ_M0, device=self.device)
        return plan.assemble(0, 0)[2][0]

    def solve_linear_system_multi(self, problem: MEEMProblem, m0) -> np.ndarray:
        """X (N_T complex) for the domains' current heaving flags (meem_engine.py:307-313)."""
        self._ensure_m_k_and_N_k_arrays(problem, m0)
        res = self._single_plan(problem, m0, want_x=True).run()[0]
        return res["X"][0, 0].copy()

    def reformat_coeffs(self, x: np.ndarray, NMK, boundary_count):
        """Split X per domain: [C1^0 | C1^1 C2^1 | ... | C^e] (meem_engine.py:315-324)."""
        out, col = [], 0
        for i in range(boundary_count + 1):
            w = NMK[i] if i in (0, boundary_count) else 2 * NMK[i]
            out.append(x[col:col + w])
            col += w
        return out

    def compute_hydrodynamic_coefficients(self, problem, X, m0, modes_to_calculate=None, rho=None):
        """Reduce a solution X to one dict per mode j (meem_engine.py:326-393).  The c vectors and the
        particular-solution integrals, the dot products and the phase / force formulas all run in the device kernel
        k_reduce (meemb200_coefficients)."""
        rho = RHO_DEFAULT if rho is None else rho
        bodies = problem.geometry.body_arrangement.bodies
        if modes_to_calculate is None:
            modes_to_calculate = np.arange(len(bodies))
        modes_to_calculate = [int(m) for m in modes_to_calculate]
        geo = flatten_problem(problem, modes=modes_to_calculate)
        plan = BatchPlan([geo], [m0], freq_kind=FREQ_M0, device=self.device, rho=rho)
        vals = plan.coefficients(X, 0, 0)               # k_reduce: [mode j][real, imag, phase, force]
        _, _, c, _ = plan.assemble(0, 0)
        return [{"mode": mode, "real": vals[j, 0], "imag": (0 if np.isinf(m0) else vals[j, 1]), "excitation_phase": vals[j, 2],
                 "excitation_force": vals[j, 3], "c_vector": c[j]} for j, mode in enumerate(modes_to_calculate)]

    # ------------------------------------------------------------------ fields
    @staticmethod
    def make_R_Z(a, h, d, sharp, spatial_res, R_range=None, Z_range=None):
        """Evaluation grid of the reference (multi_equations.py:478-501): host-side bookkeeping."""
        a = np.asarray(a, dtype=float)
        d = np.asarray(d, dtype=float)
        if R_range is not None:
            r_vec = np.asarray(R_range, dtype=float)
        else:
            rmin = (2 * a[-1] / spatial_res) if sharp else 0.0
            r_vec = np.linspace(rmin, 2 * a[-1], spatial_res)
        z_vec = np.asarray(Z_range, dtype=float) if Z_range is not None else np.linspace(0, -h, spatial_res)
        if sharp:
            a_eps = 1.0e-4
            for i in range(len(a)):
                r_vec = np.append(r_vec, a[i] * (1 - a_eps))
                r_vec = np.append(r_vec, a[i] * (1 + a_eps))
            r_vec = np.unique(r_vec)
            for i in range(len(d)):
                z_vec = np.append(z_vec, -d[i])
            z_vec = np.unique(z_vec)
        return np.meshgrid(r_vec, z_vec, indexing="ij")

    def _raw_fields(self, problem, solution_vector, m0, R, Z):
        geo = flatten_problem(problem, from_domain_flags=True)
        # the particular solution follows the BODIES' heaving flags (meem_engine.py:405-407), per region
        geo.mode_flags = np.asarray([[1 if x else 0 for x in problem.geometry.body_arrangement.heaving]], dtype=np.uint8)
        plan = BatchPlan([geo], [m0], freq_kind=FREQ_M0, device=self.device)
        raw = plan.fields(solution_vector, R.ravel(), Z.ravel())
        return geo, {k: v.reshape(R.shape) for k, v in raw.items()}

    def calculate_potentials(self, problem, solution_vector, m0, spatial_res, sharp, R_range=None, Z_range=None):
        """phiH, phiP, phi on the reference's (R, Z) grid (meem_engine.py:395-467); arithmetic on the device, the
        region / inside-the-body masks applied here exactly as the reference does."""
        self._ensure_m_k_and_N_k_arrays(problem, m0)
        ba = problem.geometry.body_arrangement
        a, d, h = np.asarray(ba.a, float), np.asarray(ba.d, float), problem.geometry.h
        R, Z = self.make_R_Z(a, h, d, sharp, spatial_res)          # the reference ignores R_range / Z_range here
        _, raw = self._raw_fields(problem, solution_vector, m0, R, Z)
        M = len(a)
        inside = np.zeros(R.shape, dtype=bool)
        inside |= (R <= a[0]) & ~(Z < -d[0])
        for i in range(1, M):
            inside |= (R > a[i - 1]) & (R <= a[i]) & ~(Z < -d[i])
        nanc = np.nan + 1j * np.nan
        phiH = np.where(inside, nanc, raw["phiH"])
        phiP = np.where(inside, nanc, raw["phiP"].astype(complex))
        return {"R": R, "Z": Z, "phiH": phiH, "phiP": phiP, "phi": phiH + phiP}

    def calculate_velocities(self, problem, solution_vector, m0, spatial_res, sharp, R_range=None, Z_range=None):
        """vrH, vzH, vrP, vzP, vr, vz (meem_engine.py:478-553)."""
        self._ensure_m_k_and_N_k_arrays(problem, m0)
        ba = problem.geometry.body_arrangement
        a, d, h = np.asarray(ba.a, float), np.asarray(ba.d, float), problem.geometry.h
        R, Z = self.make_R_Z(a, h, d, sharp, spatial_res, R_range=R_range, Z_range=Z_range)
        _, raw = self._raw_fields(problem, solution_vector, m0, R, Z)
        vrH, vzH = raw["vrH"], raw["vzH"]
        vrP, vzP = raw["vrP"].astype(complex), raw["vzP"].astype(complex)
        vr, vz = vrH + vrP, vzH + vzP
        M = len(a)
        for i in range(M):
            reg = (R <= a[0]) if i == 0 else ((R > a[i - 1]) & (R <= a[i]))
            body = reg & (Z > -d[i])
            vr[body] = np.nan
            vz[body] = np.nan
        return {"R": R, "Z": Z, "vrH": vrH, "vzH": vzH, "vrP": vrP, "vzP": vzP, "vr": vr, "vz": vz}

    # ------------------------------------------------------------------ the hot path
    def run_and_store_results(self, problem_index: int) -> Results:
        """Frequency x mode sweep of one problem (meem_engine.py:555-690) in ONE device batch."""
        return self.run_batch([self.problem_list[problem_index]])[0]

    def run_batch(self, problems: Sequence[MEEMProblem], want_potentials: bool = True, group=None) -> List[Results]:
        """Many problems at once.  Problems sharing a frequency grid go into one device batch
        (geometry x frequency x mode); different grids are grouped.  With ``group`` (a torch.distributed process
        group, one process per GPU) every batch is sharded over the group's GPUs (openflash_b200.distributed.ShardedPlan)
        and every rank returns the complete results."""
        problems = list(problems)
        groups: Dict[bytes, List[int]] = {}
        for idx, prob in enumerate(problems):
            key = np.ascontiguousarray(np.asarray(prob.frequencies, dtype=np.float64)).tobytes()
            groups.setdefault(key, []).append(idx)
        results: List[Optional[Results]] = [None] * len(problems)
        for key, idxs in groups.items():
            omegas = np.frombuffer(key, dtype=np.float64)
            flats = []
            for i in idxs:
                prob = problems[i]
                modes = prob.modes
                if len(modes) == 0:
                    raise ValueError("problem has no heaving body: nothing to radiate")
                flats.append(flatten_problem(prob, modes=list(modes)))
            if group is not None:
                from .distributed import ShardedPlan
                out = ShardedPlan(flats, omegas, group=group, want_x=want_potentials, want_c=True, device=self.device,
                                  algo=self.algo).run()
            else:
                out = BatchPlan(flats, omegas, want_x=want_potentials, want_c=True, device=self.device, algo=self.algo).run()
            for i, geo, r in zip(idxs, flats, out):
                prob = problems[i]
                res = Results(prob)
                F, R = len(omegas), geo.R
                p = geo.size - int(geo.NMK[-1])
                # the c vector reported per (frequency, mode) is that of mode j = i (eng.py:646-652); it is frequency
                # independent and comes out of the same batch call (MeemOutputs.c_vector)
                cmat = np.broadcast_to(r["c_vector"].astype(complex)[None, :, :], (F, R, p)).copy()
                bad = r["status"] != 0
                cmat[bad] = np.nan + 0j
                res.store_hydrodynamic_coefficients(frequencies=prob.frequencies, added_mass_matrix=r["added_mass"],
                                                    damping_matrix=r["damping"], excitation_force=r["excitation_force"],
                                                    excitation_phase=r["excitation_phase"], c_vector_matrix=cmat)
                if want_potentials:
                    res.store_potential_coefficients(r["X"], [int(x) for x in geo.NMK])
                res.status = r["status"]
                res.m0 = r["m0"]
                results[i] = res
        return results
