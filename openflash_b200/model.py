"""Host-side object model kept from the reference API: bodies, arrangements, fluid domains and the
concentric-cylinder geometry (reference: body.py, geometry.py, domain.py, basic_region_geometry.py).

Only bookkeeping and validation live here; at the engine boundary a geometry is flattened to the
arrays ``h, a[], d[], NMK[], heaving[]`` (meem_engine.py:110-116) and handed to the CUDA library.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np


# ----------------------------------------------------------------------------------------- bodies
class Body:
    """Base class of physical bodies; every body carries a ``heaving`` flag."""
    heaving: bool = False


class SteppedBody(Body):
    """Concentric vertical-walled steps: outer radii ``a``, drafts ``d``, slant angles (body.py:25)."""

    def __init__(self, a, d, slant_angle, heaving: bool = False):
        if not (len(a) == len(d) == len(slant_angle)):
            raise AssertionError("Input arrays a, d, and slant_angle must have the same length.")
        self.a, self.d, self.slant_angle = a, d, slant_angle
        self.heaving = heaving


class CoordinateBody(Body):
    """Body outlined by (r, z) points; placeholder in the reference as well (body.py:42)."""

    def __init__(self, r_coords, z_coords, heaving: bool = False):
        if len(r_coords) != len(z_coords):
            raise AssertionError("r_coords and z_coords must be the same length.")
        self.r_coords, self.z_coords = r_coords, z_coords
        self.heaving = heaving

    def discretize(self):
        r, z = self.r_coords, self.z_coords
        return r, z, np.gradient(z, r)


class BodyArrangement:
    """A set of bodies of which at most one may be flagged heaving at construction (geometry.py:18-24)."""

    def __init__(self, bodies: Sequence[Body]):
        self.bodies = bodies
        if sum(1 for b in bodies if getattr(b, "heaving", False)) > 1:
            raise ValueError("Only 0 or 1 body can be marked as heaving")


class ConcentricBodyGroup(BodyArrangement):
    """Concentric SteppedBody objects; per-step properties are concatenated body by body."""

    def __init__(self, bodies: Sequence[Body]):
        super().__init__(bodies)
        if any(not isinstance(b, SteppedBody) for b in bodies):
            raise TypeError("ConcentricBodyGroup currently only supports SteppedBody objects.")

    def _cat(self, name):
        return np.concatenate([getattr(b, name) for b in self.bodies])

    @property
    def a(self):
        return self._cat("a")

    @property
    def d(self):
        return self._cat("d")

    @property
    def slant_angle(self):
        return self._cat("slant_angle")

    @property
    def heaving(self):
        return np.array([b.heaving for b in self.bodies for _ in range(len(b.a))], dtype=bool)

    @property
    def body_of_region(self):
        """region index -> body index (the expansion meem_engine.py:350-359 recomputes per call)."""
        return np.array([bi for bi, b in enumerate(self.bodies) for _ in range(len(b.a))], dtype=int)


# ----------------------------------------------------------------------------------------- domains
class Domain:
    """One annular fluid region (domain.py:14-58).  ``h``, ``di`` and ``a`` are the aliases the
    engine reads."""

    def __init__(self, index: int, NMK: int, a_inner: float, a_outer: float, d_lower: float, geometry_h: float,
                 heaving: Optional[bool] = None, slant: bool = False, category: str = "interior"):
        assert isinstance(NMK, (int, np.integer)) and NMK > 0, "NMK must be a positive integer."
        assert a_outer > a_inner >= 0, "Radii must be valid (a_outer > a_inner >= 0)."
        assert geometry_h >= d_lower >= 0, "Depths must be valid (d_upper >= d_lower >= 0)."
        self.index = index
        self.number_harmonics = int(NMK)
        self.a_inner, self.a_outer = a_inner, a_outer
        self.d_lower, self.d_upper = d_lower, geometry_h
        self.heaving, self.slant, self.category = heaving, slant, category

    @property
    def h(self):
        return self.d_upper

    @property
    def di(self):
        return self.d_lower

    @property
    def a(self):
        return self.a_outer

    @staticmethod
    def are_adjacent(d1: "Domain", d2: "Domain", atol: float = 1e-6) -> bool:
        for x, y in ((d1, d2), (d2, d1)):
            if np.isfinite(x.a_outer) and np.isclose(x.a_outer, y.a_inner, atol=atol):
                return True
        return False


# ----------------------------------------------------------------------------------------- geometry
class Geometry:
    """A body arrangement in water of depth ``h`` plus the fluid domains it induces."""

    def __init__(self, body_arrangement: BodyArrangement, h: float):
        self.body_arrangement = body_arrangement
        self.h = h
        self._fluid_domains: List[Domain] = []

    def make_fluid_domains(self) -> List[Domain]:
        raise NotImplementedError

    @property
    def fluid_domains(self):
        if not self._fluid_domains:
            self._fluid_domains = self.make_fluid_domains()
        return self._fluid_domains

    @property
    def domain_list(self) -> Dict[int, Domain]:
        fd = self.fluid_domains
        if isinstance(fd, dict):
            return fd
        return {dom.index: dom for dom in fd}


class BasicRegionGeometry(Geometry):
    """Radii increase outwards, so regions map 1:1 onto steps, plus one exterior region
    (basic_region_geometry.py:22-47, :118-163)."""

    def __init__(self, body_arrangement: ConcentricBodyGroup, h: float, NMK: List[int]):
        super().__init__(body_arrangement, h)
        self.NMK = NMK
        radii = np.asarray(body_arrangement.a, dtype=float)
        if not np.all(np.diff(radii) > 0):
            raise ValueError("Radii 'a' must be strictly increasing for BasicRegionGeometry. "
                             "Use AnyRegionGeometry for other cases.")
        if len(NMK) != len(radii) + 1:
            raise ValueError("Length of NMK must be one greater than the total number of body radii.")
        self._domain_dict = {dom.index: dom for dom in self.make_fluid_domains()}

    @property
    def domain_list(self) -> Dict[int, Domain]:
        return self._domain_dict

    def make_fluid_domains(self) -> List[Domain]:
        arr = self.body_arrangement
        out, inner = [], 0.0
        for i, (r, d, hv, sl) in enumerate(zip(arr.a, arr.d, arr.heaving, arr.slant_angle)):
            out.append(Domain(i, int(self.NMK[i]), inner, r, d, self.h, heaving=bool(hv), slant=bool(sl)))
            inner = r
        out.append(Domain(len(out), int(self.NMK[-1]), inner, np.inf, 0.0, self.h, category="exterior"))
        return out

    @classmethod
    def from_vectors(cls, a, d, h, NMK, slant_angle=None, body_map=None, heaving_map=None):
        """Build from flat vectors; ``body_map[i]`` assigns step i to a body (basic_region_geometry.py:49-116)."""
        a = np.asarray(a)
        d = np.asarray(d)
        slant_angle = np.zeros_like(a) if slant_angle is None else np.asarray(slant_angle)
        body_map = [0] * len(a) if body_map is None else list(body_map)
        n_bodies = max(body_map) + 1
        if heaving_map is None:
            heaving_map = [False] * n_bodies
        elif len(heaving_map) != n_bodies:
            raise ValueError(f"Length of heaving_map ({len(heaving_map)}) does not match inferred number of bodies ({n_bodies})")
        assert sum(bool(x) for x in heaving_map) <= 1, "Only 0 or 1 body can be marked as heaving"
        groups = [[j for j, b in enumerate(body_map) if b == k] for k in range(n_bodies)]
        last = -np.inf
        for k, idx in enumerate(groups):
            if not idx:
                raise ValueError(f"Body index {k} is declared in body_map but has no assigned radii.")
            for r in a[idx]:
                if r <= last:
                    raise ValueError("Radii 'a' must be strictly increasing after applying body_map. "
                                     "Detected non-monotonic or backtracking group arrangement.")
                last = r
        bodies = [SteppedBody(a=a[idx], d=d[idx], slant_angle=slant_angle[idx], heaving=heaving_map[k])
                  for k, idx in enumerate(groups)]
        return cls(ConcentricBodyGroup(bodies), h, NMK)


# ----------------------------------------------------------------------------------------- problem
class MEEMProblem:
    """Geometry + angular frequencies; ``modes`` are the indices of heaving bodies (meem_problem.py)."""

    def __init__(self, geometry: Geometry):
        self.geometry = geometry
        self.domain_list = geometry.domain_list
        self.frequencies = np.array([])

    def set_frequencies(self, frequencies: np.ndarray):
        assert isinstance(frequencies, np.ndarray), "frequencies must be a numpy array"
        assert np.all(frequencies > 0), "All frequencies must be positive"
        self.frequencies = frequencies

    @property
    def modes(self) -> np.ndarray:
        return np.array([i for i, b in enumerate(self.geometry.body_arrangement.bodies) if b.heaving])
