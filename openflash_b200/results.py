"""Results container with the reference's variable names and dimensions (results.py:118-250).

``Results.dataset`` is an ``xarray.Dataset`` when xarray is importable, otherwise a small stand-in
with the same access pattern (``ds["added_mass"].values``, ``.dims``, ``ds.coords[...]``).
"""
from __future__ import annotations

from typing import Optional

import numpy as np

try:  # pragma: no cover - depends on the environment
    import xarray as _xr
except Exception:  # xarray is optional
    _xr = None


class _Var:
    def __init__(self, dims, data, coords=None):
        self.dims = tuple(dims)
        self.data = np.asarray(data)
        self.coords = dict(coords or {})

    @property
    def values(self):
        return self.data

    @property
    def shape(self):
        return self.data.shape

    def __array__(self, dtype=None):
        return self.data if dtype is None else self.data.astype(dtype)


class DatasetLite:
    """Minimal named-array container used when xarray is not installed."""

    def __init__(self, coords=None, attrs=None):
        self.coords = dict(coords or {})
        self.attrs = dict(attrs or {})
        self.data_vars = {}

    def __setitem__(self, name, value):
        self.data_vars[name] = _Var(*value) if isinstance(value, tuple) else value

    def __getitem__(self, name):
        return self.data_vars[name]

    def __contains__(self, name):
        return name in self.data_vars

    def keys(self):
        return self.data_vars.keys()


class Results:
    def __init__(self, problem, modes=None):
        self.geometry = problem.geometry
        self.frequencies = problem.frequencies
        if modes is not None:
            self.modes = np.array(modes)
        else:
            self.modes = np.array([i for i, b in enumerate(self.geometry.body_arrangement.bodies) if b.heaving])
        coords = {"frequency": self.frequencies, "mode_i": self.modes, "mode_j": self.modes}
        self.dataset = _xr.Dataset(coords=coords) if _xr is not None else DatasetLite(coords=coords)

    def _put(self, name, dims, data, coords=None):
        if _xr is not None:
            self.dataset[name] = _xr.DataArray(data, dims=list(dims), coords=coords) if coords else (tuple(dims), data)
        else:
            self.dataset[name] = _Var(dims, data, coords)

    def store_hydrodynamic_coefficients(self, frequencies, added_mass_matrix, damping_matrix,
                                        excitation_force: Optional[np.ndarray] = None,
                                        excitation_phase: Optional[np.ndarray] = None,
                                        c_vector_matrix: Optional[np.ndarray] = None):
        """dims: added_mass/damping (frequency, mode_i, mode_j) with [f, j, i] = force on j from motion of i;
        excitation_* (frequency, mode_i); c_vector (frequency, mode_i, c_vector_index)."""
        expected = (len(frequencies), len(self.modes), len(self.modes))
        if added_mass_matrix.shape != expected or damping_matrix.shape != expected:
            raise ValueError(f"Matrices must have shape (num_frequencies, num_modes, num_modes). "
                             f"Expected {expected} (based on heaving flags), but got {added_mass_matrix.shape}.")
        self._put("added_mass", ("frequency", "mode_i", "mode_j"), added_mass_matrix)
        self._put("damping", ("frequency", "mode_i", "mode_j"), damping_matrix)
        if excitation_force is not None:
            self._put("excitation_force", ("frequency", "mode_i"), excitation_force)
        if excitation_phase is not None:
            self._put("excitation_phase", ("frequency", "mode_i"), excitation_phase)
        if c_vector_matrix is not None:
            if "c_vector_index" not in self.dataset.coords:
                self.dataset.coords["c_vector_index"] = np.arange(c_vector_matrix.shape[-1])
            self._put("c_vector", ("frequency", "mode_i", "c_vector_index"), c_vector_matrix)

    def store_potential_coefficients(self, X, NMK):
        """X[F, R, N_T] -> potentials_real/imag[F, R, domain, harmonic], NaN padded to the widest
        domain (2N for annular regions) — what store_all_potentials builds from the per-domain split
        (meem_engine.py:654-671, results.py:118-190)."""
        F, R, _ = X.shape
        nd = len(NMK)
        widths = [NMK[i] if i in (0, nd - 1) else 2 * NMK[i] for i in range(nd)]
        mx = max(widths)
        pot = np.full((F, R, nd, mx), np.nan + 1j * np.nan, dtype=complex)
        col = 0
        for i, w in enumerate(widths):
            pot[:, :, i, :w] = X[:, :, col:col + w]
            col += w
        coords = {"frequency": self.frequencies, "modes": self.modes, "domain_name": list(range(nd)), "harmonics": np.arange(mx)}
        for k, v in coords.items():
            if k not in self.dataset.coords:
                self.dataset.coords[k] = v
        dims = ("frequency", "modes", "domain_name", "harmonics")
        self._put("potentials_real", dims, pot.real, coords)
        self._put("potentials_imag", dims, pot.imag, coords)
        zc = np.full((nd, mx), np.nan)
        for i, w in enumerate(widths):
            zc[i, :w] = 0.0
        self._put("potential_r_coords", ("domain_name", "harmonics"), zc)
        self._put("potential_z_coords", ("domain_name", "harmonics"), zc.copy())

    def get_results(self):
        return self.dataset

    def export_to_netcdf(self, file_path: str):
        """NetCDF export with the reference's layout (results.py:252-269): every data variable under its name and
        dimensions, complex variables split into ``<name>_real`` / ``<name>_imag``, coordinates as coordinate variables.
        With xarray + h5netcdf installed this is the reference's own call; without them (this image) the same content is
        written as a NetCDF-3 (64-bit offset) file by ``scipy.io.netcdf_file``, which xarray / netCDF4 read back with the
        same variable names."""
        if _xr is not None:
            ds = self.dataset
            out = {}
            for name in ds.data_vars:
                data = ds[name].data
                if np.iscomplexobj(data):
                    out[name + "_real"] = (ds[name].dims, np.real(data))
                    out[name + "_imag"] = (ds[name].dims, np.imag(data))
                else:
                    out[name] = (ds[name].dims, data)
            _xr.Dataset(out, coords=ds.coords).to_netcdf(file_path, engine="h5netcdf")
            return
        from scipy.io import netcdf_file
        ds = self.dataset
        variables = {}
        for name, var in ds.data_vars.items():
            data = np.asarray(var.values)
            if np.iscomplexobj(data):
                variables[name + "_real"] = (var.dims, np.real(data))
                variables[name + "_imag"] = (var.dims, np.imag(data))
            else:
                variables[name] = (var.dims, data)
        sizes = {}
        for dims, data in variables.values():
            for d, n in zip(dims, data.shape):
                if sizes.setdefault(d, n) != n:
                    raise ValueError(f"dimension {d!r} has inconsistent lengths {sizes[d]} and {n}")
        with netcdf_file(file_path, "w", version=2) as nc:
            for d, n in sizes.items():
                nc.createDimension(d, int(n))
            for d in sizes:                                   # coordinate variables (numeric coordinates only)
                c = ds.coords.get(d)
                if c is None:
                    continue
                c = np.asarray(c)
                if c.shape == (sizes[d],) and c.dtype.kind in "iuf":
                    c = c.astype(np.float64 if c.dtype.kind == "f" else np.int32)
                    nc.createVariable(d, c.dtype, (d,))[:] = c
            for name, (dims, data) in variables.items():
                data = np.ascontiguousarray(data, dtype=np.float64 if data.dtype.kind == "f" else np.int32)
                v = nc.createVariable(name, data.dtype, tuple(dims))
                v[:] = data
            nc.history = "openflash_b200 Results.export_to_netcdf (NetCDF-3 fallback: xarray not installed)"
