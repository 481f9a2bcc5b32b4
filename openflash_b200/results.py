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

    def __str__(self):
        lines = ["<DatasetLite>", "Coordinates:"]
        lines += [f"    {k}: {np.shape(v)}" for k, v in self.coords.items()]
        lines.append("Data variables:")
        lines += [f"    {k} {v.dims} {v.shape}" for k, v in self.data_vars.items()]
        return "\n".join(lines)


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

    # -- reference storage API (results.py:40-218) -----------------------------------------------------------
    def store_results(self, domain_index: int, radial_data: np.ndarray, vertical_data: np.ndarray):
        """Eigenfunction samples of one domain, dims (frequency, modes, r) / (frequency, modes, z) (results.py:40-91).
        The sample coordinates are those of the arrangement's CoordinateBody objects, concatenated."""
        domain = self.geometry.domain_list.get(domain_index)
        if domain is None:
            raise ValueError(f"Domain index {domain_index} not found.")
        from .model import CoordinateBody
        rs = [np.asarray(b.r_coords) for b in self.geometry.body_arrangement.bodies if isinstance(b, CoordinateBody)]
        zs = [np.asarray(b.z_coords) for b in self.geometry.body_arrangement.bodies if isinstance(b, CoordinateBody)]
        r_coords = np.concatenate(rs) if rs else np.array([])
        z_coords = np.concatenate(zs) if zs else np.array([])
        nf, nm = len(self.frequencies), len(self.modes)
        radial_data, vertical_data = np.asarray(radial_data), np.asarray(vertical_data)
        if radial_data.shape != (nf, nm, len(r_coords)):
            raise ValueError(f"radial_data shape {radial_data.shape} does not match expected ({nf}, {nm}, {len(r_coords)})")
        if vertical_data.shape != (nf, nm, len(z_coords)):
            raise ValueError(f"vertical_data shape {vertical_data.shape} does not match expected ({nf}, {nm}, {len(z_coords)})")
        name = domain.category if getattr(domain, "category", None) else f"domain_{domain_index}"
        base = {"frequency": self.frequencies, "modes": self.modes}
        self._put(f"radial_eigenfunctions_{name}", ("frequency", "modes", "r"), radial_data, dict(base, r=r_coords))
        self._put(f"vertical_eigenfunctions_{name}", ("frequency", "modes", "z"), vertical_data, dict(base, z=z_coords))

    def store_single_potential_field(self, potential_data: dict, frequency_idx: int = 0, mode_idx: int = 0):
        """One computed potential field {R, Z, phi} (MEEMEngine.calculate_potentials) as four (z_coord, r_coord)
        variables named potential_{R,Z,phi_real,phi_imag}_<mode>_<frequency> (results.py:93-116)."""
        if not all(k in potential_data for k in ("R", "Z", "phi")):
            raise ValueError("potential_data must contain 'R', 'Z', and 'phi' keys.")
        tag, dims = f"{mode_idx}_{frequency_idx}", ("z_coord", "r_coord")
        phi = np.asarray(potential_data["phi"])
        self._put(f"potential_R_{tag}", dims, np.asarray(potential_data["R"]))
        self._put(f"potential_Z_{tag}", dims, np.asarray(potential_data["Z"]))
        self._put(f"potential_phi_real_{tag}", dims, phi.real)
        self._put(f"potential_phi_imag_{tag}", dims, phi.imag)

    def store_velocity_field(self, velocity_data: dict, frequency_idx: int = 0, mode_idx: int = 0):
        """vr / vz of MEEMEngine.calculate_velocities as velocity_{vr,vz}_{real,imag}_<mode>_<frequency> (results.py:205-218)."""
        tag, dims = f"{mode_idx}_{frequency_idx}", ("z_coord", "r_coord")
        for comp in ("vr", "vz"):
            v = np.asarray(velocity_data[comp])
            self._put(f"velocity_{comp}_real_{tag}", dims, v.real)
            self._put(f"velocity_{comp}_imag_{tag}", dims, v.imag)

    def store_all_potentials(self, all_potentials_batch):
        """The reference's batch format (results.py:118-203; built at meem_engine.py:654-671): a list of
        {"frequency_idx", "mode_idx", "data": {domain: {"potentials", "r_coords_dict", "z_coords_dict"}}} ->
        potentials_real/imag[frequency, modes, domain_name, harmonics] NaN-padded to the longest domain, plus the
        per-domain coordinate tables.  `store_potential_coefficients` is the array-in fast path the engine uses."""
        if not all_potentials_batch:
            return
        names = sorted({dn for item in all_potentials_batch for dn in item["data"].keys()})
        if not names:
            return
        width = max(len(dd["potentials"]) for item in all_potentials_batch for dd in item["data"].values())
        nd = len(names)
        pot = np.full((len(self.frequencies), len(self.modes), nd, width), np.nan + 1j * np.nan, dtype=complex)
        rc = np.full((nd, width), np.nan)
        zc = np.full((nd, width), np.nan)
        where = {dn: i for i, dn in enumerate(names)}
        for item in all_potentials_batch:
            fi, mi = item["frequency_idx"], item["mode_idx"]
            for dn, dd in item["data"].items():
                di = where[dn]
                vals = np.asarray(dd["potentials"], dtype=complex)
                pot[fi, mi, di, :len(vals)] = vals
                for table, key in ((rc, "r_coords_dict"), (zc, "z_coords_dict")):
                    flat = np.concatenate([np.atleast_1d(v) for _, v in sorted(dd[key].items())])
                    table[di, :len(flat)] = flat
        self._store_potential_arrays(pot, names, rc, zc)

    def _store_potential_arrays(self, pot, names, rc, zc):
        width = pot.shape[-1]
        coords = {"frequency": self.frequencies, "modes": self.modes, "domain_name": list(names), "harmonics": np.arange(width)}
        for k in ("harmonics", "domain_name", "modes"):
            if k not in self.dataset.coords:
                self.dataset.coords[k] = coords[k]
        dims = ("frequency", "modes", "domain_name", "harmonics")
        self._put("potentials_real", dims, pot.real, coords)
        self._put("potentials_imag", dims, pot.imag, coords)
        sub = {"domain_name": list(names), "harmonics": np.arange(width)}
        self._put("potential_r_coords", ("domain_name", "harmonics"), rc, sub)
        self._put("potential_z_coords", ("domain_name", "harmonics"), zc, sub)

    def display_results(self):
        """results.py:274-278."""
        return str(self.dataset) if self.dataset is not None else "No results stored."

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
        zc = np.full((nd, mx), np.nan)
        for i, w in enumerate(widths):
            zc[i, :w] = 0.0
        self._store_potential_arrays(pot, list(range(nd)), zc, zc.copy())

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
