"""CPU tests: object-model validation (mirrors the reference's test_geometry / test_basic_region_geometry /
test_meem_problem error cases), flattening at the engine boundary, and the C-ABI library exports."""
import ctypes
import os
import re

import numpy as np
import pytest

import openflash_b200 as ofb
from openflash_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _geo(**kw):
    args = dict(a=np.array([0.5, 1.0]), d=np.array([0.5, 0.25]), h=1.001, NMK=[4, 4, 4], body_map=[0, 1], heaving_map=[True, False])
    args.update(kw)
    return ofb.BasicRegionGeometry.from_vectors(**args)


def test_library_exports_every_declared_symbol():
    """include/meemb200.h <-> libmeemb200.so (no compute calls: works without a GPU)."""
    hdr = open(os.path.join(ROOT, "include", "meemb200.h")).read()
    declared = set(re.findall(r"\b(meemb200_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_lib.EXPORTS)
    lib = _lib.load()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.meemb200_version() >= 100


def test_no_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = _lib.load()
    ctx = ctypes.c_void_p()
    rc = lib.meemb200_create(0, ctypes.byref(ctx))
    assert rc == -4 and b"no CPU fallback" in lib.meemb200_last_error()
    with pytest.raises(Exception):
        ofb.solve_batch([ofb.FlatGeometry(1.0, [0.5], [0.2], [4, 4])], [1.0])


def test_geometry_validation_errors():
    with pytest.raises(ValueError):
        _geo(a=np.array([1.0, 0.5]))                       # radii must increase
    with pytest.raises(ValueError):
        _geo(NMK=[4, 4])                                   # len(NMK) = len(a) + 1
    with pytest.raises(ValueError):
        _geo(heaving_map=[True])                           # heaving_map length
    with pytest.raises(AssertionError):
        _geo(heaving_map=[True, True])                     # at most one heaving body
    with pytest.raises(ValueError):
        _geo(body_map=[0, 2])                              # body 1 has no radii
    with pytest.raises(ValueError):
        ofb.ConcentricBodyGroup([ofb.SteppedBody(np.array([1.0]), np.array([0.1]), np.array([0.0]), True),
                                 ofb.SteppedBody(np.array([2.0]), np.array([0.1]), np.array([0.0]), True)])
    with pytest.raises(TypeError):
        ofb.ConcentricBodyGroup([ofb.CoordinateBody(np.array([1.0, 2.0]), np.array([0.1, 0.2]))])
    with pytest.raises(AssertionError):
        ofb.SteppedBody(np.array([1.0, 2.0]), np.array([0.1]), np.array([0.0]))
    with pytest.raises(AssertionError):
        ofb.Domain(0, 0, 0.0, 1.0, 0.1, 1.0)
    with pytest.raises(AssertionError):
        ofb.Domain(0, 3, 1.0, 0.5, 0.1, 1.0)


def test_domains_and_problem():
    geo = _geo()
    doms = geo.domain_list
    assert sorted(doms) == [0, 1, 2]
    assert doms[2].category == "exterior" and np.isinf(doms[2].a) and doms[2].di == 0.0
    assert doms[1].a_inner == 0.5 and doms[1].h == 1.001 and doms[0].heaving and not doms[1].heaving
    prob = ofb.MEEMProblem(geo)
    assert list(prob.modes) == [0]
    with pytest.raises(AssertionError):
        prob.set_frequencies(np.array([1.0, -1.0]))
    with pytest.raises(AssertionError):
        prob.set_frequencies([1.0])
    prob.set_frequencies(np.array([0.5, 1.0]))
    for b in geo.body_arrangement.bodies:       # multi-mode by flipping flags afterwards
        b.heaving = True
    assert list(prob.modes) == [0, 1]
    flat = ofb.flatten_problem(prob, modes=list(prob.modes))
    assert flat.M == 2 and flat.R == 2 and flat.size == 16
    assert flat.mode_flags.tolist() == [[1, 0], [0, 1]]
    assert np.allclose(flat.a, [0.5, 1.0]) and np.allclose(flat.d, [0.5, 0.25]) and flat.h == 1.001


def test_flatten_multi_step_body():
    geo = ofb.BasicRegionGeometry.from_vectors(a=np.array([5.0, 10.0, 15.0]), d=np.array([15.0, 10.0, 5.0]), h=20.0,
                                               NMK=[5] * 4, body_map=[0, 1, 1], heaving_map=[False, True])
    prob = ofb.MEEMProblem(geo)
    flat = ofb.flatten_problem(prob, modes=[1])
    assert flat.mode_flags.tolist() == [[0, 1, 1]]
    flat2 = ofb.flatten_problem(prob, from_domain_flags=True)
    assert flat2.mode_flags.tolist() == [[0, 1, 1]]


def test_results_container_shapes():
    geo = _geo()
    prob = ofb.MEEMProblem(geo)
    prob.set_frequencies(np.array([0.5, 1.0, 2.0]))
    res = ofb.Results(prob)
    F, R = 3, 1
    res.store_hydrodynamic_coefficients(prob.frequencies, np.ones((F, R, R)), np.ones((F, R, R)), np.ones((F, R)),
                                        np.ones((F, R)), np.ones((F, R, 12), dtype=complex))
    ds = res.dataset
    assert ds["added_mass"].dims == ("frequency", "mode_i", "mode_j") and ds["added_mass"].values.shape == (F, R, R)
    assert ds["excitation_phase"].dims == ("frequency", "mode_i")
    assert ds["c_vector"].values.shape == (F, R, 12)
    with pytest.raises(ValueError):
        res.store_hydrodynamic_coefficients(prob.frequencies, np.ones((F, 2, 2)), np.ones((F, 2, 2)))
    X = np.arange(F * R * 16, dtype=complex).reshape(F, R, 16)
    res.store_potential_coefficients(X, [4, 4, 4])
    pr = ds["potentials_real"].values
    assert pr.shape == (F, R, 3, 8) and np.isnan(pr[0, 0, 0, 4]) and pr[0, 0, 1, 7] == 11.0 and pr[0, 0, 2, 0] == 12.0


def test_packed_triangle_and_syrk_tile_layouts():
    """Host view of two device layouts (meemb200_debug_layout; integer functions only, no GPU): the packed lower-triangular
    system matrix of k_freq_solve / k_hchol and the warp-tile list of the symmetric rank update in k_freq_gemm."""
    lib = _lib.load()
    lib.meemb200_debug_layout.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int), ctypes.c_int]
    for K in (1, 2, 7, 10, 16, 50, 199, 200, 400):
        out = (ctypes.c_int * (K + 2))()
        assert lib.meemb200_debug_layout(0, K, out, K + 2) == 0
        off = np.array(out[:K + 1])
        assert off[0] == 0 and np.all(off % 2 == 0)                       # 16-byte aligned rows (double2 accesses)
        assert np.all(np.diff(off) >= np.arange(K) + 1)                    # row r holds its r + 1 entries without overlap
        assert np.all(np.diff(off) <= np.arange(K) + 2)                    # ... with at most one padding slot
        assert out[K + 1] >= off[K] and out[K + 1] % 8 == 0 and out[K + 1] - off[K] < 8
    for Nc in (64, 100, 199, 200, 256):
        buf = (ctypes.c_int * (1 + 6 * 96))()
        assert lib.meemb200_debug_layout(1, Nc, buf, len(buf)) == 0
        nt, nb = buf[0], (Nc + 7) // 8
        tiles = np.array(buf[1:1 + 6 * nt]).reshape(nt, 6)
        cover = np.zeros((nb, nb), int)
        sizes = []
        for a, b, r0, nr, c0, nc in tiles:
            assert 1 <= nr <= 4 and 1 <= nc <= 4 and a >= b
            if a == b:
                assert (r0, nr) == (c0, nc)
                for i in range(nr):
                    cover[r0 + i, c0:c0 + i + 1] += 1                      # diagonal tiles compute their lower blocks only
                sizes.append(nr * (nr + 1) // 2)
            else:
                cover[r0:r0 + nr, c0:c0 + nc] += 1
                sizes.append(nr * nc)
        assert np.array_equal(cover, np.tril(np.ones((nb, nb), int)))      # every lower block exactly once, nothing above
        assert sizes == sorted(sizes, reverse=True)                        # largest first: balanced passes of 8 warps
    assert lib.meemb200_debug_layout(1, 1000, buf, len(buf)) != 0          # too wide for the staged SYRK: refused, not garbage


def test_chunk_stream_layouts_of_the_frequency_gemm():
    """Host view of the chunk stream of k_freq_gemm's triangular product (meemb200_debug_layout kinds 2 and 3; integer
    functions only, no GPU): every row group's blocks are dealt to the 8 warps exactly once and within one block of each
    other, a warp never holds more than 16 accumulator blocks, the ring has at least 3 stages of whatever chunk size is
    chosen, the chunk count matches the row groups, and the re-tiled factor's tiles are back to back with 36-double rows."""
    lib = _lib.load()
    lib.meemb200_debug_layout.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int), ctypes.c_int]
    out = (ctypes.c_int * 88)()
    seen_kc = set()
    for n in (56, 64, 72, 75, 90, 99, 100, 104, 128, 150, 184, 200, 224, 232, 248, 256, 264, 400):
        assert lib.meemb200_debug_layout(2, n, out, 88) == 0
        fits, kc, stages, ngr, chunks, nb, lat, lds = out[0:8]
        assert nb == (n + 7) // 8
        assert fits == int(8 <= nb <= 32), n                          # the stream's range: 8 .. 32 column blocks
        assert lds % 16 == 4 and lds >= nb * 8 + 4                      # slot rows: conflict-free fragment loads, room for every block
        if not fits:
            assert lat == 0                                             # no re-tiled factor where the classic path runs
            continue
        seen_kc.add(kc)
        assert kc in (8, 16) and stages >= 3 and stages * kc * (36 + lds) <= 12800
        mbw = (n - 1 + 7) // 8
        assert ngr == (mbw + 3) // 4
        assert chunks == -(-n // kc) + sum(-(-(n - 32 * a) // kc) for a in range(ngr))
        for a in range(ngr):
            NI, row0, *blocks = out[8 + 10 * a: 18 + 10 * a]
            assert NI == min(4, mbw - 4 * a) and row0 == 32 * a
            assert sum(blocks) == NI * nb                               # every block of the group exactly once
            assert max(blocks) - min(blocks) <= 1 and max(blocks) <= 16  # balanced over the warps, fits the accumulators
            assert blocks == sorted(blocks, reverse=True)               # the last warp (the producer) is the lightest
        np8 = (n + 7) // 8 * 8
        assert lat == 36 * (np8 + sum(np8 - 32 * a for a in range((n + 31) // 32)))
        nt = (n + 31) // 32
        tiles = (ctypes.c_int * (nt + 2))()
        assert lib.meemb200_debug_layout(3, n, tiles, nt + 2) == 0
        off = list(tiles)
        assert off[0] == 0 and off[1] == 36 * np8 and off[-1] == lat
        assert all(off[a + 2] - off[a + 1] == 36 * (np8 - 32 * a) for a in range(nt))
    assert seen_kc == {8, 16}


def test_results_netcdf_export_roundtrip(tmp_path):
    """Results.export_to_netcdf (results.py:252-269): variable names, dimensions and the real / imag split survive a round
    trip through the NetCDF file (scipy reader; xarray-free image)."""
    from scipy.io import netcdf_file
    geo = _geo(NMK=[3, 4, 5])
    prob = ofb.MEEMProblem(geo)
    om = np.array([0.5, 1.0, 2.0])
    prob.set_frequencies(om)
    res = ofb.Results(prob)
    R = len(res.modes)
    rng = np.random.default_rng(0)
    am, dp = rng.normal(size=(3, R, R)), rng.normal(size=(3, R, R))
    fo, ph = rng.normal(size=(3, R)), rng.normal(size=(3, R))
    res.store_hydrodynamic_coefficients(om, am, dp, fo, ph)
    X = rng.normal(size=(3, R, 3 + 2 * 4 + 5)) + 1j * rng.normal(size=(3, R, 3 + 2 * 4 + 5))
    res.store_potential_coefficients(X, [3, 4, 5])
    path = str(tmp_path / "out.nc")
    res.export_to_netcdf(path)
    if ofb.results._xr is not None:
        pytest.skip("xarray present: the reference's own h5netcdf writer was used")
    with netcdf_file(path, "r", mmap=False) as nc:
        assert nc.variables["added_mass"].dimensions == ("frequency", "mode_i", "mode_j")
        assert np.array_equal(nc.variables["added_mass"][:], am) and np.array_equal(nc.variables["damping"][:], dp)
        assert nc.variables["excitation_phase"].dimensions == ("frequency", "mode_i")
        assert np.array_equal(nc.variables["excitation_force"][:], fo)
        assert np.array_equal(nc.variables["frequency"][:], om)
        pr, pi = nc.variables["potentials_real"][:], nc.variables["potentials_imag"][:]
        assert nc.variables["potentials_real"].dimensions == ("frequency", "modes", "domain_name", "harmonics")
        assert pr.shape == (3, R, 3, 8)
        assert np.array_equal(pr[:, :, 1, :8], X.real[:, :, 3:11]) and np.array_equal(pi[:, :, 2, :5], X.imag[:, :, 11:16])
        assert np.isnan(pr[:, :, 0, 3:]).all()                # NaN padding to the widest domain, as in the reference
