"""CPU tests of the round-2 host pieces: the reference's Results storage API (mirrors package/test/test_results.py),
the packed output layout, and the sharded (multi-GPU) plan over a world-size-2 gloo group with a stand-in device plan."""
import os
from types import SimpleNamespace

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import openflash_b200 as ofb
from openflash_b200.batch import GeometryBatch, output_layout, split_outputs, unpack_outputs
from openflash_b200.distributed import ShardedPlan, geometry_cost


def _problem():
    """Three coordinate bodies (2 sample points in total), the last two heaving; two named domains."""
    bodies = [ofb.CoordinateBody(np.array([0.5]), np.array([-0.5]), heaving=False),
              ofb.CoordinateBody(np.array([1.0]), np.array([-1.0]), heaving=True),
              ofb.CoordinateBody(np.array([]), np.array([]), heaving=True)]
    geometry = SimpleNamespace(body_arrangement=SimpleNamespace(bodies=bodies),
                               domain_list={0: SimpleNamespace(category="inner", index=0), 1: SimpleNamespace(category="outer", index=1)})
    return SimpleNamespace(geometry=geometry, frequencies=np.array([0.5, 1.0, 1.5]))


def test_results_modes_and_store_results():
    res = ofb.Results(_problem())
    assert list(res.modes) == [1, 2]
    rad = np.random.default_rng(0).random((3, 2, 2))
    ver = np.random.default_rng(1).random((3, 2, 2))
    res.store_results(0, rad, ver)
    ds = res.get_results()
    assert ds["radial_eigenfunctions_inner"].dims == ("frequency", "modes", "r")
    assert np.array_equal(ds["vertical_eigenfunctions_inner"].values, ver)
    with pytest.raises(ValueError, match="not found"):
        res.store_results(7, rad, ver)
    with pytest.raises(ValueError, match="radial_data shape"):
        res.store_results(1, rad[:, :, :1], ver)
    with pytest.raises(ValueError, match="vertical_data shape"):
        res.store_results(1, rad, ver[:2])
    assert "radial_eigenfunctions_inner" in res.display_results()


def test_results_single_fields():
    res = ofb.Results(_problem())
    R, Z = np.meshgrid(np.linspace(0, 2, 4), np.linspace(0, -1, 3), indexing="ij")
    phi = (R + 1j * Z)
    res.store_single_potential_field({"R": R, "Z": Z, "phi": phi}, frequency_idx=2, mode_idx=1)
    ds = res.dataset
    assert ds["potential_phi_real_1_2"].dims == ("z_coord", "r_coord")
    assert np.array_equal(ds["potential_phi_imag_1_2"].values, Z) and np.array_equal(ds["potential_R_1_2"].values, R)
    with pytest.raises(ValueError, match="must contain"):
        res.store_single_potential_field({"R": R, "Z": Z})
    res.store_velocity_field({"vr": phi, "vz": 2 * phi}, frequency_idx=0, mode_idx=0)
    assert np.array_equal(ds["velocity_vz_imag_0_0"].values, 2 * Z) and np.array_equal(ds["velocity_vr_real_0_0"].values, R)


def test_results_store_all_potentials_batch_format():
    """The reference's batch dicts (meem_engine.py:654-671) and the engine's array fast path give the same variables."""
    res = ofb.Results(_problem())
    NMK = [3, 2, 4]
    widths = [3, 4, 4]
    rng = np.random.default_rng(5)
    X = rng.random((3, 2, 11)) + 1j * rng.random((3, 2, 11))
    batch = []
    for f in range(3):
        for m in range(2):
            data, col = {}, 0
            for dom, w in enumerate(widths):
                data[dom] = {"potentials": X[f, m, col:col + w], "r_coords_dict": {f"r_h{k}": 0.0 for k in range(w)},
                             "z_coords_dict": {f"z_h{k}": 0.0 for k in range(w)}}
                col += w
            batch.append({"frequency_idx": f, "mode_idx": m, "data": data})
    res.store_all_potentials(batch)
    fast = ofb.Results(_problem())
    fast.store_potential_coefficients(X, NMK)
    for name in ("potentials_real", "potentials_imag", "potential_r_coords", "potential_z_coords"):
        a, b = res.dataset[name], fast.dataset[name]
        assert a.dims == b.dims
        assert np.array_equal(np.asarray(a.values), np.asarray(b.values), equal_nan=True)
    assert res.dataset["potentials_real"].shape == (3, 2, 3, 4)
    assert np.isnan(res.dataset["potentials_real"].values[0, 0, 0, 3])          # first domain is padded
    ofb.Results(_problem()).store_all_potentials([])                             # empty batch: nothing stored, no error


def test_output_layout_roundtrip():
    F, R, NT, K = 3, [1, 2], [10, 14], [4, 5]
    seg, total = output_layout(F, R, NT, K, want_c=True, want_x=True)
    assert all(off % 256 == 0 for off, _ in seg.values()) and total % 256 == 0
    assert seg["c_vector"][1] == 8 * (1 * 6 + 2 * 9) and seg["X"][1] == 16 * 3 * (10 + 28)
    buf = np.zeros(total, dtype=np.uint8)
    off, n = seg["added_mass"]
    buf[off:off + n].view(np.float64)[:] = np.arange(n // 8)
    flat = unpack_outputs(buf, seg, 2, F, True, True)
    parts = split_outputs(flat, F, R, NT, K, True, True)
    assert parts[0]["added_mass"].shape == (3, 1, 1) and parts[1]["added_mass"].shape == (3, 2, 2)
    assert parts[1]["added_mass"][0, 0, 0] == 3.0 and parts[1]["c_vector"].shape == (2, 9) and parts[1]["X"].shape == (3, 2, 14)


class _FakePlan:
    """Stand-in for BatchPlan on CPU tensors: 'solves' by writing a value that encodes (geometry, frequency) identity."""

    def __init__(self, geometries, freqs, want_c=False, want_x=False, out_bytes_min=0, **kw):
        gb = geometries if isinstance(geometries, GeometryBatch) else GeometryBatch.from_list(geometries)
        self.gb, self.freqs = gb, np.asarray(freqs)
        self.F, self.G = len(self.freqs), len(gb)
        self.R, self.NT, self.K = gb.R, gb.NT, gb.K
        self.want_c, self.want_x = want_c, want_x
        self.seg, nbytes = output_layout(self.F, self.R, self.NT, self.K, want_c, want_x)
        self.d_out = torch.zeros(max(nbytes, out_bytes_min), dtype=torch.uint8)

    def upload(self):
        pass

    def solve(self):
        buf = self.d_out.numpy()
        am = []
        for h, R in zip(self.gb.h, self.gb.R):
            for w in self.freqs:
                am += [h * 1000 + w + 0.01 * (j * R + i) for j in range(R) for i in range(R)]
        off, n = self.seg["added_mass"]
        buf[off:off + n].view(np.float64)[:] = am
        if self.want_c:
            off, n = self.seg["c_vector"]
            buf[off:off + n].view(np.float64)[:] = np.concatenate([np.full(int(R * (NT - K)), h) for h, R, NT, K in
                                                                    zip(self.gb.h, self.R, self.NT, self.K)])

    def download(self):
        return unpack_outputs(self.d_out.numpy(), self.seg, self.G, self.F, self.want_c, self.want_x)

    def split(self, flat):
        return split_outputs(flat, self.F, self.R, self.NT, self.K, self.want_c, self.want_x)


def _sharded_worker(rank, world, port, tmp):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ok = True
    # geometry mode: 7 geometries of different size / mode count, each identified by its h
    geos = [ofb.FlatGeometry(float(i + 1), np.arange(1, 2 + i % 3, dtype=float), np.linspace(0.5, 0.2, 1 + i % 3), [4 + i] * (2 + i % 3),
                             [[1] * (1 + i % 3)] * (1 + i % 2)) for i in range(7)]
    om = np.array([0.25, 0.5, 0.75])
    sp = ShardedPlan(geos, om, plan_factory=_FakePlan, want_c=True)
    ok &= sp.mode == "geometry" and sorted(np.concatenate(sp.geom_parts).tolist()) == list(range(7))
    res = sp.run()
    for g, r in zip(geos, res):
        ok &= r["added_mass"].shape == (3, g.R, g.R)
        ok &= bool(np.allclose(r["added_mass"][:, 0, 0], g.h * 1000 + om))
        ok &= r["c_vector"].shape == (g.R, g.size - int(g.NMK[-1])) and bool(np.all(r["c_vector"] == g.h))
    # frequency mode: one geometry, 5 frequencies over 2 ranks, re-joined in order
    om2 = np.array([0.1, 0.2, 0.3, 0.4, 0.5])
    sp2 = ShardedPlan(geos[:1], om2, plan_factory=_FakePlan)
    res2 = sp2.run()
    ok &= sp2.mode == "frequency" and bool(np.allclose(res2[0]["added_mass"][:, 0, 0], 1000 + om2))
    # gather="root": only rank 1 receives
    sp3 = ShardedPlan(geos, om, plan_factory=_FakePlan, gather="root", root=1)
    res3 = sp3.run()
    if rank == 1:
        ok &= all(bool(np.allclose(r["added_mass"][:, 0, 0], g.h * 1000 + om)) for g, r in zip(geos, res3))
    else:
        ok &= res3 is None
    with open(os.path.join(tmp, f"ok{rank}"), "w") as f:
        f.write("1" if ok else "0")
    dist.destroy_process_group()


def test_sharded_plan_world2_gloo(tmp_path):
    world = 2
    port = 31500 + (os.getpid() % 2000)
    mp.spawn(_sharded_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert open(tmp_path / f"ok{r}").read() == "1"


def test_geometry_cost_orders_by_size():
    small = ofb.FlatGeometry(50.0, [1.0, 2.0], [5.0, 2.0], [16] * 3)
    big = ofb.FlatGeometry(50.0, np.arange(1.0, 9.0), np.linspace(20, 2, 8), [16] * 9)
    assert geometry_cost(big, 64) > geometry_cost(small, 64) > 0


def test_geometry_batch_packing_take_and_costs():
    """Packed batch == per-geometry flattening (meem_engine.py:110-116): exterior a = inf / d = 0 appended, flags padded,
    N_T = N_0 + 2 sum N_i + K; `take` re-packs a subset; costs grow with size."""
    geos = [ofb.FlatGeometry(float(i + 1), np.arange(1, 2 + i % 3, dtype=float), np.linspace(0.5, 0.2, 1 + i % 3), [4 + i] * (2 + i % 3),
                             [[1] * (1 + i % 3)] * (1 + i % 2)) for i in range(9)]
    gb = GeometryBatch.from_list(geos)
    assert len(gb) == 9 and list(gb.M) == [g.M for g in geos] and list(gb.NT) == [g.size for g in geos]
    assert list(gb.K) == [int(g.NMK[-1]) for g in geos] and list(gb.N) == [int(g.NMK[-2]) for g in geos]
    for gi, g in enumerate(geos):
        r0, r1 = gb.reg_off[gi], gb.reg_off[gi + 1]
        assert np.array_equal(gb.a[r0:r1 - 1], g.a) and np.isinf(gb.a[r1 - 1]) and gb.d[r1 - 1] == 0.0
        assert np.array_equal(gb.nmk[r0:r1], g.NMK)
        fl = gb.flags[gb.flag_off[gi]:gb.flag_off[gi + 1]].reshape(g.R, g.M + 1)
        assert np.array_equal(fl[:, :g.M], g.mode_flags) and np.all(fl[:, g.M] == 0)
    sub = gb.take([7, 2, 5])
    ref = GeometryBatch.from_list([geos[7], geos[2], geos[5]])
    for name in ("h", "reg_off", "a", "d", "nmk", "n_modes", "flag_off", "flags", "NT"):
        assert np.array_equal(getattr(sub, name), getattr(ref, name)), name
    c = gb.costs(64)
    assert np.all(c > 0) and np.allclose(c, [geometry_cost(g, 64) for g in geos])
    single = GeometryBatch.from_list([ofb.FlatGeometry(50.0, [1.0, 2.0], [5.0, 2.0], [16] * 3)] * 3)      # all R == 1: vectorised flag path
    assert np.array_equal(single.flags.reshape(3, 3), [[1, 1, 0]] * 3)
