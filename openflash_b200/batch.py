"""Batched (geometry x frequency x radiating mode) MEEM solves on one B200.

Host-side half of the C ABI in include/meemb200.h: flattens geometries the way the reference does
at its engine boundary (meem_engine.py:110-116), owns the device buffers (torch tensors used purely
as memory) and enqueues the CUDA pipeline.  All arithmetic happens in libmeemb200.so.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np

from . import _lib
from ._lib import ALGO_DENSE, ALGO_SCHUR, ALGO_SCHUR_LU, FREQ_M0, FREQ_OMEGA, ROOTS_CONVERGED, ROOTS_REFERENCE  # noqa: F401
from .multi_constants import rho as RHO_DEFAULT


@dataclass
class FlatGeometry:
    """One concentric-cylinder geometry as flat arrays.

    ``a`` and ``d`` list the M interior regions (the exterior's ``a=inf, d=0`` are appended when
    packing, as the reference does); ``NMK`` has M+1 entries; ``mode_flags[r][i]`` tells whether
    region i moves in radiating mode r (body -> region expansion of meem_engine.py:350-367).
    """
    h: float
    a: Sequence[float]
    d: Sequence[float]
    NMK: Sequence[int]
    mode_flags: Sequence[Sequence[int]] = field(default_factory=list)

    def __post_init__(self):
        self.a = np.asarray(self.a, dtype=np.float64).ravel()
        self.d = np.asarray(self.d, dtype=np.float64).ravel()
        self.NMK = np.asarray(self.NMK, dtype=np.int32).ravel()
        M = len(self.a)
        if len(self.d) != M or len(self.NMK) != M + 1:
            raise ValueError("need len(a) == len(d) == len(NMK) - 1")
        if len(self.mode_flags) == 0:
            self.mode_flags = [[1] * M]
        self.mode_flags = np.asarray(self.mode_flags, dtype=np.uint8).reshape(-1, M)

    @property
    def M(self):
        return len(self.a)

    @property
    def R(self):
        return self.mode_flags.shape[0]

    @property
    def size(self):
        N = self.NMK
        return int(N[0] + N[-1] + 2 * N[1:-1].sum())

    @classmethod
    def from_body_map(cls, h, a, d, NMK, body_map, modes=None):
        """Modes are heaving-body indices (meem_problem.py:37-49); body b moves all its regions."""
        body_map = np.asarray(body_map)
        if modes is None:
            modes = sorted(set(int(b) for b in body_map))
        flags = [[1 if body_map[i] == b else 0 for i in range(len(body_map))] for b in modes]
        return cls(h, a, d, NMK, flags)


class GeometryBatch:
    """Many geometries as packed arrays - exactly the CSR layout of `MeemBatch` (include/meemb200.h): per geometry M + 1
    entries of a / d / nmk (the exterior's a = inf, d = 0 included) from reg_off[g], and R_g * (M_g + 1) heaving flags from
    flag_off[g].  Design sweeps build this once with array arithmetic instead of one Python object per geometry;
    `BatchPlan` and `ShardedPlan` accept it wherever they accept a list of `FlatGeometry`."""

    def __init__(self, h, reg_off, a, d, nmk, n_modes, flag_off, flags):
        self.h = np.ascontiguousarray(h, dtype=np.float64)
        self.reg_off = np.ascontiguousarray(reg_off, dtype=np.int32)
        self.a = np.ascontiguousarray(a, dtype=np.float64)
        self.d = np.ascontiguousarray(d, dtype=np.float64)
        self.nmk = np.ascontiguousarray(nmk, dtype=np.int32)
        self.n_modes = np.ascontiguousarray(n_modes, dtype=np.int32)
        self.flag_off = np.ascontiguousarray(flag_off, dtype=np.int64)
        self.flags = np.ascontiguousarray(flags, dtype=np.uint8)
        G = len(self.h)
        if len(self.reg_off) != G + 1 or len(self.flag_off) != G + 1 or len(self.n_modes) != G:
            raise ValueError("inconsistent packed geometry arrays")
        self.M = (np.diff(self.reg_off) - 1).astype(np.int32)
        nmk64 = self.nmk.astype(np.int64)
        tot = np.add.reduceat(nmk64, self.reg_off[:-1])                      # sum of all M + 1 entries
        first, last = nmk64[self.reg_off[:-1]], nmk64[self.reg_off[1:] - 1]
        self.K = last
        self.N = nmk64[self.reg_off[1:] - 2]
        self.NT = 2 * tot - first - last                                       # N_0 + 2 sum N_1..M-1 + K (meem_engine.py:116)
        self.R = self.n_modes

    def __len__(self):
        return len(self.h)

    @classmethod
    def from_list(cls, geos):
        geos = list(geos)
        G = len(geos)
        M = np.fromiter((len(g.a) for g in geos), dtype=np.int64, count=G)
        R = np.fromiter((g.mode_flags.shape[0] for g in geos), dtype=np.int64, count=G)
        reg_off = np.zeros(G + 1, dtype=np.int64)
        np.cumsum(M + 1, out=reg_off[1:])
        total = int(reg_off[-1])
        interior = np.ones(total, dtype=bool)
        interior[reg_off[1:] - 1] = False
        a = np.full(total, np.inf)
        d = np.zeros(total)
        a[interior] = np.concatenate([g.a for g in geos])
        d[interior] = np.concatenate([g.d for g in geos])
        nmk = np.concatenate([g.NMK for g in geos])
        h = np.fromiter((g.h for g in geos), dtype=np.float64, count=G)
        flag_off = np.zeros(G + 1, dtype=np.int64)
        np.cumsum(R * (M + 1), out=flag_off[1:])
        flags = np.zeros(int(flag_off[-1]), dtype=np.uint8)
        if np.all(R == 1):
            flags[interior] = np.concatenate([g.mode_flags.ravel() for g in geos])
        else:
            for gi, g in enumerate(geos):
                fl = np.zeros((g.R, g.M + 1), dtype=np.uint8)
                fl[:, :g.M] = g.mode_flags
                flags[flag_off[gi]:flag_off[gi + 1]] = fl.ravel()
        return cls(h, reg_off, a, d, nmk, R, flag_off, flags)

    @classmethod
    def from_dicts(cls, dicts):
        """From dicts with keys h, a, d, NMK, body_map, modes (bench.py's workload generators): body b moves all its
        regions.  Single-mode dicts are packed with array arithmetic only."""
        dicts = list(dicts)
        if not all(len(g["modes"]) == 1 for g in dicts):
            out = []
            for g in dicts:
                bm = np.asarray(g["body_map"])
                fl = np.array([[1 if b == m else 0 for b in bm] for m in g["modes"]], dtype=np.uint8)
                out.append(FlatGeometry(g["h"], g["a"], g["d"], g["NMK"], fl))
            return cls.from_list(out)
        G = len(dicts)
        M = np.fromiter((len(g["a"]) for g in dicts), dtype=np.int64, count=G)
        reg_off = np.zeros(G + 1, dtype=np.int64)
        np.cumsum(M + 1, out=reg_off[1:])
        total = int(reg_off[-1])
        interior = np.ones(total, dtype=bool)
        interior[reg_off[1:] - 1] = False
        a = np.full(total, np.inf)
        d = np.zeros(total)
        a[interior] = np.concatenate([np.asarray(g["a"], dtype=np.float64) for g in dicts])
        d[interior] = np.concatenate([np.asarray(g["d"], dtype=np.float64) for g in dicts])
        nmk = np.concatenate([np.asarray(g["NMK"], dtype=np.int32) for g in dicts])
        if len(nmk) != total or len(a[interior]) != total - G:
            raise ValueError("need len(a) == len(d) == len(NMK) - 1 for every geometry")
        h = np.fromiter((g["h"] for g in dicts), dtype=np.float64, count=G)
        bm = np.concatenate([np.asarray(g["body_map"], dtype=np.int64) for g in dicts])
        mode = np.repeat(np.fromiter((g["modes"][0] for g in dicts), dtype=np.int64, count=G), M)
        flags = np.zeros(total, dtype=np.uint8)
        flags[interior] = (bm == mode).astype(np.uint8)
        return cls(h, reg_off, a, d, nmk, np.ones(G, dtype=np.int32), reg_off.copy(), flags)

    def take(self, idx):
        """Sub-batch of the geometries `idx` (array of indices), packed again - vectorised CSR gather."""
        idx = np.asarray(idx, dtype=np.int64)
        ln = (self.M[idx] + 1).astype(np.int64)
        new_off = np.zeros(len(idx) + 1, dtype=np.int64)
        np.cumsum(ln, out=new_off[1:])
        src = np.repeat(self.reg_off[idx].astype(np.int64) - new_off[:-1], ln) + np.arange(int(new_off[-1]))
        fln = self.n_modes[idx].astype(np.int64) * ln
        new_foff = np.zeros(len(idx) + 1, dtype=np.int64)
        np.cumsum(fln, out=new_foff[1:])
        fsrc = np.repeat(self.flag_off[idx] - new_foff[:-1], fln) + np.arange(int(new_foff[-1]))
        return GeometryBatch(self.h[idx], new_off, self.a[src], self.d[src], self.nmk[src], self.n_modes[idx], new_foff, self.flags[fsrc])

    def costs(self, n_freq):
        """Relative device cost per geometry: window elimination (~25 N^3 per interior boundary) + n_freq Schur stages."""
        n3 = self.nmk.astype(np.float64) ** 3
        geom = 25.0 * (np.add.reduceat(n3, self.reg_off[:-1]) - self.K.astype(np.float64) ** 3)
        N, K = self.N.astype(np.float64), self.K.astype(np.float64)
        return geom + n_freq * (N * N * K + N * K * K + K ** 3 / 3.0)


def output_layout(F, R, NT, K, want_c=False, want_x=False):
    """Byte layout of the packed output buffer of a batch: {name: (offset, nbytes)}, total bytes.  R, NT, K are the
    per-geometry mode counts, system sizes and exterior mode counts.  Pure host arithmetic: every rank of a sharded
    solve can compute every other rank's layout (openflash_b200.distributed)."""
    R = np.asarray(R, dtype=np.int64)
    NT = np.asarray(NT, dtype=np.int64)
    K = np.asarray(K, dtype=np.int64)
    n_rr, n_r, n_sys = int(F * (R * R).sum()), int(F * R.sum()), int(F * len(R))
    segs = [("added_mass", 8 * n_rr), ("damping", 8 * n_rr), ("excitation_force", 8 * n_r),
            ("excitation_phase", 8 * n_r), ("m0", 8 * n_sys), ("status", 4 * n_sys)]
    if want_c:
        segs.append(("c_vector", 8 * int((R * (NT - K)).sum())))
    if want_x:
        segs.append(("X", 16 * int(F * (R * NT).sum())))
    seg, off = {}, 0
    for name, nbytes in segs:
        seg[name] = (off, nbytes)
        off = (off + nbytes + 255) & ~255
    return seg, max(off, 256)


def unpack_outputs(buf, seg, G, F, want_c=False, want_x=False, copy=True):
    """Flat numpy arrays out of a packed output buffer (uint8 numpy array) with layout `seg`."""
    def arr(name, dtype):
        off, nbytes = seg[name]
        a = buf[off:off + nbytes].view(dtype)
        return a.copy() if copy else a
    out = dict(added_mass=arr("added_mass", np.float64), damping=arr("damping", np.float64),
               excitation_force=arr("excitation_force", np.float64), excitation_phase=arr("excitation_phase", np.float64),
               m0=arr("m0", np.float64).reshape(G, F), status=arr("status", np.int32).reshape(G, F))
    if want_c:
        out["c_vector"] = arr("c_vector", np.float64)
    if want_x:
        out["X"] = arr("X", np.complex128)
    return out


def split_outputs(flat, F, R, NT, K, want_c=False, want_x=False):
    """Per-geometry views with the reference's shapes: added_mass[F,R,R] (force on j, motion of i), ..."""
    R64 = np.asarray(R, dtype=np.int64)
    NT = np.asarray(NT, dtype=np.int64)
    K = np.asarray(K, dtype=np.int64)
    off_rr = np.concatenate([[0], np.cumsum(F * R64 * R64)])
    off_r = np.concatenate([[0], np.cumsum(F * R64)])
    off_x = np.concatenate([[0], np.cumsum(F * R64 * NT)])
    off_c = np.concatenate([[0], np.cumsum(R64 * (NT - K))])
    res = []
    for g in range(len(R64)):
        Rg = int(R64[g])
        r = dict(added_mass=flat["added_mass"][off_rr[g]:off_rr[g + 1]].reshape(F, Rg, Rg),
                 damping=flat["damping"][off_rr[g]:off_rr[g + 1]].reshape(F, Rg, Rg),
                 excitation_force=flat["excitation_force"][off_r[g]:off_r[g + 1]].reshape(F, Rg),
                 excitation_phase=flat["excitation_phase"][off_r[g]:off_r[g + 1]].reshape(F, Rg),
                 m0=flat["m0"][g], status=flat["status"][g])
        if want_x:
            r["X"] = flat["X"][off_x[g]:off_x[g + 1]].reshape(F, Rg, int(NT[g]))
        if want_c:      # [R, p]: c vector of every radiating mode (frequency independent, meem_engine.py:65-98)
            r["c_vector"] = flat["c_vector"][off_c[g]:off_c[g + 1]].reshape(Rg, -1)
        res.append(r)
    return res


_contexts = {}


def get_context(device: int = 0):
    """One meemb200 context per (process, GPU)."""
    if device not in _contexts:
        lib = _lib.load()
        ctx = C.c_void_p()
        _lib.check(lib.meemb200_create(int(device), C.byref(ctx)))
        _contexts[device] = ctx
    return _contexts[device]


class BatchPlan:
    """A validated, packed batch plus its device buffers; ``solve()`` may be called repeatedly."""

    def __init__(self, geometries: Sequence[FlatGeometry], freqs, *, freq_kind=FREQ_OMEGA, algo=ALGO_SCHUR,
                 root_mode=ROOTS_REFERENCE, want_x=False, want_c=False, rho: Optional[float] = None, device: int = 0,
                 out_bytes_min: int = 0):
        import torch
        self.torch = torch
        self.lib = _lib.load()
        self.device_index = int(device)
        self.ctx = get_context(self.device_index)
        self.device = torch.device("cuda", self.device_index)
        gb = geometries if isinstance(geometries, GeometryBatch) else None
        if gb is None:
            geometries = list(geometries)
            if not geometries:
                raise ValueError("empty batch")
            gb = GeometryBatch.from_list(geometries)
        if len(gb) == 0:
            raise ValueError("empty batch")
        self.batch = gb
        self.G = len(gb)
        self.freqs = np.ascontiguousarray(np.asarray(freqs, dtype=np.float64).ravel())
        self.F = len(self.freqs)
        self.R = gb.R
        self.M = gb.M
        self.NT = gb.NT
        self.want_x = bool(want_x)
        self.want_c = bool(want_c)
        self.algo = int(algo)
        reg_off, h, a, d, nmk, flag_off, flags = gb.reg_off, gb.h, gb.a, gb.d, gb.nmk, gb.flag_off, gb.flags
        self._keep = (reg_off, h, a, d, nmk, self.R, flag_off, flags)
        b = _lib.MeemBatch()
        b.n_geom, b.n_freq = self.G, self.F
        b.reg_off, b.h, b.a, b.d, b.nmk = (x.ctypes.data for x in (reg_off, h, a, d, nmk))
        b.n_modes, b.flag_off, b.heaving = self.R.ctypes.data, flag_off.ctypes.data, flags.ctypes.data
        b.freq = self.freqs.ctypes.data
        b.freq_kind, b.algo, b.root_mode, b.want_x = int(freq_kind), int(algo), int(root_mode), int(self.want_x)
        b.rho = float(RHO_DEFAULT if rho is None else rho)
        plan = C.c_void_p()
        rc = self.lib.meemb200_plan_create(self.ctx, C.byref(b), C.byref(plan))
        if rc == -1:
            raise ValueError(self.lib.meemb200_last_error().decode())
        _lib.check(rc)
        self.plan = plan
        self.info = _lib.MeemPlanInfo()
        _lib.check(self.lib.meemb200_plan_info(self.plan, C.byref(self.info)))
        # device buffers (torch tensors as plain memory) and the pinned staging copy of the inputs.  Every output lives
        # in ONE device buffer (segments 256-byte aligned) mirrored by ONE pinned host buffer, so download() is a single
        # device->host copy at full link speed instead of one pageable copy per array.
        nb = int(self.info.input_bytes)
        self.h_blob = torch.empty(nb, dtype=torch.uint8, pin_memory=True)
        self._refresh_host_blob()
        self.K = gb.K
        self._seg, self._out_bytes = output_layout(self.F, self.R, self.NT, self.K, self.want_c, self.want_x)
        self._out_bytes = max(self._out_bytes, int(out_bytes_min))
        with torch.cuda.device(self.device):
            self.d_blob = torch.empty(nb, dtype=torch.uint8, device=self.device)
            self.d_ws = torch.empty(int(self.info.workspace_bytes), dtype=torch.uint8, device=self.device)
            self.d_out = torch.empty(self._out_bytes, dtype=torch.uint8, device=self.device)
        self.h_out = torch.empty(self._out_bytes, dtype=torch.uint8, pin_memory=True)
        base = self.d_out.data_ptr()
        ptr = lambda name: (base + self._seg[name][0]) if name in self._seg else None  # noqa: E731
        self.out = _lib.MeemOutputs(ptr("added_mass"), ptr("damping"), ptr("excitation_force"), ptr("excitation_phase"),
                                    ptr("m0"), ptr("status"), ptr("X"), ptr("c_vector"))
        self._uploaded = False

    def _refresh_host_blob(self):
        nb = int(self.info.input_bytes)
        src = (C.c_ubyte * nb).from_address(self.lib.meemb200_plan_blob(self.plan))
        self.h_blob.numpy()[:] = np.frombuffer(src, dtype=np.uint8)

    def __del__(self):
        try:
            if getattr(self, "plan", None):
                self.lib.meemb200_plan_destroy(self.plan)
                self.plan = None
        except Exception:
            pass

    # -- data movement -------------------------------------------------------------------------
    @property
    def h2d_bytes(self):
        return int(self.info.input_bytes)

    @property
    def d2h_bytes(self):
        return int(16 * self.info.n_rr + 16 * self.info.n_r + 12 * self.info.n_sys + 16 * self.info.n_x
                   + (8 * self.info.n_c if self.want_c else 0))

    def set_frequencies(self, freqs):
        freqs = np.ascontiguousarray(np.asarray(freqs, dtype=np.float64).ravel())
        if len(freqs) != self.F:
            raise ValueError("frequency grid length is fixed by the plan")
        rc = self.lib.meemb200_plan_set_freq(self.plan, freqs.ctypes.data)
        if rc == -1:
            raise ValueError(self.lib.meemb200_last_error().decode())
        _lib.check(rc)
        self.freqs = freqs
        self._refresh_host_blob()
        self._uploaded = False

    def set_test_hook(self, flag_every: int):
        """Test support: every n-th system is reported as a failed Cholesky and re-solved by the pivoted LU (0 = off)."""
        _lib.check(self.lib.meemb200_plan_set_test_hook(self.plan, int(flag_every)))

    def upload(self):
        """ONE host->device copy of the packed inputs (pinned memory, current stream)."""
        self.d_blob.copy_(self.h_blob, non_blocking=True)
        self._uploaded = True

    def solve(self):
        """Enqueue the kernels on torch's current stream (asynchronous)."""
        if not self._uploaded:
            self.upload()
        stream = self.torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self.lib.meemb200_solve(self.plan, self.d_blob.data_ptr(), self.d_ws.data_ptr(),
                                           int(self.info.workspace_bytes), C.byref(self.out), C.c_void_p(stream)))

    def stage_ms(self):
        ms = (C.c_float * 5)()
        _lib.check(self.lib.meemb200_last_timing(self.plan, ms))
        return dict(roots=ms[0], geometry=ms[1], frequency=ms[2], freq_gemm=ms[3], freq_solve=ms[4])

    PHASES = ("imk", "dots", "gemm_H_Imk", "gemm_ImkT_T", "cholesky", "backsolve", "coefficients", "x")

    def phase_profile(self):
        """Share of SM cycles per phase of the frequency-stage kernel of the last solve (measurement only)."""
        cyc = (C.c_double * 16)()
        stream = self.torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self.lib.meemb200_phase_profile(self.plan, self.d_ws.data_ptr(), cyc, C.c_void_p(stream)))
        cyc[1] = 0.0                      # slot 1 counts factorised geometries (k_hchol), not cycles
        tot = sum(cyc[:8]) or 1.0
        out = {k: cyc[i] / tot for i, k in enumerate(self.PHASES)}
        out["coop_geom_cycles"] = [cyc[i] for i in range(8, 16)]
        return out

    def factorised_geometries(self):
        """Number of geometries of the last solve whose coupling matrix H was Cholesky-factorised (k_hchol), i.e. that ran
        the W = L^T Imk form of the frequency GEMM (measurement only)."""
        cyc = (C.c_double * 16)()
        stream = self.torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self.lib.meemb200_phase_profile(self.plan, self.d_ws.data_ptr(), cyc, C.c_void_p(stream)))
        return int(cyc[1])

    def _view(self, buf, name, dtype):
        off, nbytes = self._seg[name]
        return buf[off:off + nbytes].view(dtype)

    def device_output(self, name):
        """Device-resident view (torch tensor) of one output array of the last solve: 'added_mass', 'damping',
        'excitation_force', 'excitation_phase', 'm0' (float64), 'status' (int32), 'c_vector', 'X' when requested."""
        t = self.torch
        return self._view(self.d_out, name, t.int32 if name == "status" else t.float64)

    def download(self, copy=True):
        """ONE device->host copy of all outputs into the plan's pinned buffer; returns flat numpy arrays
        (copies by default; ``copy=False`` returns views of the pinned buffer, valid until the next download)."""
        t = self.torch
        self.h_out.copy_(self.d_out, non_blocking=True)
        t.cuda.current_stream(self.device).synchronize()
        return unpack_outputs(self.h_out.numpy(), self._seg, self.G, self.F, self.want_c, self.want_x, copy)

    def download_stacked(self, copy=False):
        """For batches whose geometries all have the same number of modes R: the outputs as arrays with a leading geometry
        axis - added_mass / damping [G, F, R, R], excitation_force / excitation_phase [G, F, R], m0 / status [G, F] - without
        a Python loop over geometries (design sweeps).  Views of the pinned download buffer unless copy=True."""
        R = int(self.R[0])
        if np.any(self.R != R):
            raise ValueError("download_stacked needs the same number of modes for every geometry")
        flat = self.download(copy=copy)
        G, F = self.G, self.F
        return dict(added_mass=flat["added_mass"].reshape(G, F, R, R), damping=flat["damping"].reshape(G, F, R, R),
                    excitation_force=flat["excitation_force"].reshape(G, F, R), excitation_phase=flat["excitation_phase"].reshape(G, F, R),
                    m0=flat["m0"], status=flat["status"])

    def run(self):
        self.upload()
        self.solve()
        return self.split(self.download())

    def split(self, flat):
        """Per-geometry views with the reference's shapes: added_mass[F,R,R] (force on j, motion of i), ..."""
        return split_outputs(flat, self.F, self.R, self.NT, self.K, self.want_c, self.want_x)

    # -- parity / debugging exports ---------------------------------------------------------------
    def assemble(self, g=0, f=0):
        """Dense A [NT,NT], b [R,NT], c [R,p], phi [R] of system (g, f) (MEEMEngine.assemble_A_multi / b / c)."""
        t = self.torch
        if not self._uploaded:
            self.upload()
        NT, R = int(self.NT[g]), int(self.R[g])
        p = NT - int(self._keep[4][self._keep[0][g + 1] - 1])
        A = t.empty(2 * NT * NT, dtype=t.float64, device=self.device)
        b = t.empty(2 * R * NT, dtype=t.float64, device=self.device)
        c = t.empty(2 * R * (p + 1), dtype=t.float64, device=self.device)
        stream = t.cuda.current_stream(self.device).cuda_stream
        _lib.check(self.lib.meemb200_assemble(self.plan, self.d_blob.data_ptr(), self.d_ws.data_ptr(),
                                              int(self.info.workspace_bytes), g, f, A.data_ptr(), b.data_ptr(),
                                              c.data_ptr(), C.c_void_p(stream)))
        cc = c.cpu().numpy().view(np.complex128).reshape(R, p + 1)
        return (A.cpu().numpy().view(np.complex128).reshape(NT, NT),
                b.cpu().numpy().view(np.complex128).reshape(R, NT),
                np.ascontiguousarray(cc[:, :p]), cc[:, p].real.copy())


    def imk_table(self, g=0, f=0):
        """The product frequency stage's coupling table I_mk [N, K] of system (g, f) (parity export)."""
        t = self.torch
        if not self._uploaded:
            self.upload()
        nmk = self._keep[4][self._keep[0][g]:self._keep[0][g + 1]]
        N, K = int(nmk[-2]), int(nmk[-1])
        out = t.empty(N * K, dtype=t.float64, device=self.device)
        stream = t.cuda.current_stream(self.device).cuda_stream
        rc = self.lib.meemb200_imk_table(self.plan, self.d_blob.data_ptr(), self.d_ws.data_ptr(), int(self.info.workspace_bytes),
                                         int(g), int(f), out.data_ptr(), C.c_void_p(stream))
        if rc == -1:
            raise ValueError(self.lib.meemb200_last_error().decode())
        _lib.check(rc)
        return out.cpu().numpy().reshape(N, K)

    GEOM_KERNELS = {-1: "none", 0: "k_geom", 1: "k_geom_coop", 2: "k_geom_band<smem,128>", 3: "k_geom_band<smem,256>",
                    4: "k_geom_band<global,512>", 5: "k_geom_coop2"}

    @property
    def geom_kernel(self):
        return self.GEOM_KERNELS[int(self.info.geom_kernel)]

    @property
    def freq_kernel(self):
        return ("k_freq_gemm+k_freq_solve", "k_freq_small", "k_freq_lu", "k_dense")[int(self.info.freq_kernel)]

    def coefficients(self, X, g=0, f=0):
        """compute_hydrodynamic_coefficients for a caller-supplied solution X ([NT] complex) of system (g, f): array
        [R, 4] = real, imag, excitation_phase, excitation_force per mode j of the plan (meem_engine.py:326-393)."""
        t = self.torch
        if not self._uploaded:
            self.upload()
        Xc = np.ascontiguousarray(np.asarray(X, dtype=np.complex128).ravel())
        if Xc.size != int(self.NT[g]):
            raise ValueError("X must have N_T entries")
        d_X = t.from_numpy(Xc.view(np.float64).copy()).to(self.device)
        R = int(self.R[g])
        out = t.empty(4 * R, dtype=t.float64, device=self.device)
        stream = t.cuda.current_stream(self.device).cuda_stream
        _lib.check(self.lib.meemb200_coefficients(self.plan, self.d_blob.data_ptr(), self.d_ws.data_ptr(),
                                                  int(self.info.workspace_bytes), int(g), int(f), d_X.data_ptr(),
                                                  out.data_ptr(), C.c_void_p(stream)))
        return out.cpu().numpy().reshape(R, 4)

    def fields(self, X, r, z, g=0, f=0, mode=0):
        """Raw field values of solution X ([NT] complex) of system (g, f), radiating mode `mode`, at points (r, z):
        dict(phiH, vrH, vzH complex; phiP, vrP, vzP real), region chosen by r alone (no body masks)."""
        t = self.torch
        if not self._uploaded:
            self.upload()
        r = np.ascontiguousarray(np.asarray(r, dtype=np.float64).ravel())
        z = np.ascontiguousarray(np.asarray(z, dtype=np.float64).ravel())
        n = r.size
        if z.size != n or n == 0:
            raise ValueError("r and z must be non-empty and of equal size")
        Xc = np.ascontiguousarray(np.asarray(X, dtype=np.complex128).ravel())
        if Xc.size != int(self.NT[g]):
            raise ValueError("X must have N_T entries")
        d_X = t.from_numpy(Xc.view(np.float64).copy()).to(self.device)
        d_r = t.from_numpy(r.copy()).to(self.device)
        d_z = t.from_numpy(z.copy()).to(self.device)
        out = t.empty(9 * n, dtype=t.float64, device=self.device)
        stream = t.cuda.current_stream(self.device).cuda_stream
        rc = self.lib.meemb200_fields(self.plan, self.d_blob.data_ptr(), self.d_ws.data_ptr(), int(self.info.workspace_bytes),
                                      int(g), int(f), int(mode), d_X.data_ptr(), n, d_r.data_ptr(), d_z.data_ptr(),
                                      out.data_ptr(), C.c_void_p(stream))
        if rc == -1:
            raise ValueError(self.lib.meemb200_last_error().decode())
        _lib.check(rc)
        o = out.cpu().numpy().reshape(9, n)
        return dict(phiH=o[0] + 1j * o[1], vrH=o[2] + 1j * o[3], vzH=o[4] + 1j * o[5], phiP=o[6].copy(), vrP=o[7].copy(), vzP=o[8].copy())


def solve_batch(geometries, freqs, **kw):
    """Convenience: plan, upload, solve, download; returns one result dict per geometry."""
    return BatchPlan(geometries, freqs, **kw).run()


def device_roots(freq, h, K, *, freq_kind=FREQ_OMEGA, root_mode=ROOTS_REFERENCE, device=0):
    """m0[n], m_k[n,K], N_k[n,K] on the GPU (wavenumber + _ensure_m_k_and_N_k_arrays)."""
    import torch
    lib = _lib.load()
    ctx = get_context(device)
    dev = torch.device("cuda", device)
    freq = np.atleast_1d(np.asarray(freq, dtype=np.float64))
    hh = np.array(np.broadcast_to(np.asarray(h, dtype=np.float64), freq.shape), dtype=np.float64, order="C")
    n = len(freq)
    d_f = torch.from_numpy(np.ascontiguousarray(freq)).to(dev)
    d_h = torch.from_numpy(hh).to(dev)
    m0 = torch.empty(n, dtype=torch.float64, device=dev)
    mk = torch.empty(n * K, dtype=torch.float64, device=dev)
    nk = torch.empty(n * K, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    _lib.check(lib.meemb200_roots(ctx, n, int(K), d_f.data_ptr(), d_h.data_ptr(), int(freq_kind), int(root_mode),
                                  m0.data_ptr(), mk.data_ptr(), nk.data_ptr(), C.c_void_p(stream)))
    return m0.cpu().numpy(), mk.cpu().numpy().reshape(n, K), nk.cpu().numpy().reshape(n, K)


def device_bessel(kind, x, device=0):
    """kind 0..7 = I0e, I1e, K0e, K1e, J0, Y0, J1, Y1 evaluated by the device kernels."""
    import torch
    lib = _lib.load()
    ctx = get_context(device)
    dev = torch.device("cuda", device)
    d_x = torch.from_numpy(np.ascontiguousarray(np.asarray(x, dtype=np.float64))).to(dev)
    out = torch.empty_like(d_x)
    stream = torch.cuda.current_stream(dev).cuda_stream
    _lib.check(lib.meemb200_bessel(ctx, int(kind), d_x.numel(), d_x.data_ptr(), out.data_ptr(), C.c_void_p(stream)))
    return out.cpu().numpy()


def fp64_peak(kind=0, device=0):
    """Measured FP64 throughput in FLOP/s (0 = DFMA loop, 1 = DMMA m8n8k4) — roofline denominator."""
    import torch
    lib = _lib.load()
    ctx = get_context(device)
    val = C.c_double()
    stream = torch.cuda.current_stream(torch.device("cuda", device)).cuda_stream
    _lib.check(lib.meemb200_fp64_peak(ctx, int(kind), C.byref(val), C.c_void_p(stream)))
    return val.value
