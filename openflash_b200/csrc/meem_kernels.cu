// meemb200: CUDA kernels (sm_100a) and C ABI of the batched MEEM hot path.  See include/meemb200.h.
//
// Pipeline of one meemb200_solve (all on the caller's stream):
//   k_m0, k_mk              dispersion roots per (root set = unique (h, K), frequency)
//   k_tables                radial Bessel tables + c-vector ingredients per geometry
//   ALGO_SCHUR: k_geom      per geometry: assemble A11 = A[:p,:p] (real, frequency independent),
//                           Gaussian elimination of [A11 | E | b1_i], reduce to H^T, u_i, w_j, s_ji
//               k_freq      per (geometry, frequency): I_mk, T = H I_mk, S = D - I_mk^T T (two FP64
//                           GEMMs), K x K solve with the single complex column carried last,
//                           coefficients (+ X on request)
//   ALGO_DENSE: k_dense     per (geometry, frequency): the reference's N_T x N_T complex system,
//                           assembled in place and solved by blocked partial-pivot LU
// Persistent CTAs pull work items from an atomic queue (items are cost-sorted by the planner) and
// own one scratch slot each, so scratch stays L2-resident.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <cstdlib>
#include <cmath>
#include <string>
#include <vector>

#include "../../include/meemb200.h"
#include "meem_forms.cuh"
#include "meem_lu.cuh"
#include "meem_chol.cuh"
#include "meem_geom_coop.cuh"
#include "meem_geom_band.cuh"

using namespace meem;

// ------------------------------------------------------------------------------------------------
// packed input blob (host-built by the planner, uploaded with one copy)
struct GeomDesc {
    int32_t M, R, K, N;          // N = harmonics of the last interior region
    int32_t NT, p;
    int32_t reg_off;             // into a / d / nmk
    int32_t root_set;
    int64_t flag_off;            // into heaving bytes
    int64_t tab_off;             // doubles, into tables workspace
    int64_t gs_off;              // doubles, into geometry-stage store
    int64_t xg_off;              // doubles, into (G, y0) store (want_x) or -1
    int64_t out_rr, out_r, out_sys, out_x, out_c;
    double h;
};
struct RootSet {
    double h;
    int32_t K, pad;
    int64_t off;                 // doubles into roots workspace: m0[F], mk[F][K], Nk[F][K]
};
struct BlobHeader {
    int32_t n_geom, n_freq, n_sets, freq_kind, root_mode, algo, want_x, pad;
    double rho;
    int64_t off_geom, off_sets, off_a, off_d, off_nmk, off_flags, off_freq, off_order;
    // workspace byte offsets
    int64_t ws_roots, ws_tables, ws_gstore, ws_xstore, ws_slots, ws_counters, ws_sysbuf;
    int64_t sys_stride;          // doubles per system buffer (frequency stage, Cholesky path)
    int64_t slot_doubles_geom, slot_doubles_freq, slot_doubles_dense;
};

struct DevView {   // resolved pointers, built on device from blob + workspace
    const BlobHeader* hdr;
    const GeomDesc* geom;
    const RootSet* sets;
    const double* a;
    const double* d;
    const int32_t* nmk;
    const uint8_t* flags;
    const double* freq;
    const int32_t* order;
    double* roots;
    double* tables;
    double* gstore;
    double* xstore;
    double* slots;
    double* sysbuf;
    int* counters;
};

__device__ __forceinline__ DevView make_view(const void* blob, void* ws) {
    const char* b = static_cast<const char*>(blob);
    char* w = static_cast<char*>(ws);
    const BlobHeader* h = reinterpret_cast<const BlobHeader*>(b);
    DevView v;
    v.hdr = h;
    v.geom = reinterpret_cast<const GeomDesc*>(b + h->off_geom);
    v.sets = reinterpret_cast<const RootSet*>(b + h->off_sets);
    v.a = reinterpret_cast<const double*>(b + h->off_a);
    v.d = reinterpret_cast<const double*>(b + h->off_d);
    v.nmk = reinterpret_cast<const int32_t*>(b + h->off_nmk);
    v.flags = reinterpret_cast<const uint8_t*>(b + h->off_flags);
    v.freq = reinterpret_cast<const double*>(b + h->off_freq);
    v.order = reinterpret_cast<const int32_t*>(b + h->off_order);
    v.roots = reinterpret_cast<double*>(w + h->ws_roots);
    v.tables = reinterpret_cast<double*>(w + h->ws_tables);
    v.gstore = reinterpret_cast<double*>(w + h->ws_gstore);
    v.xstore = reinterpret_cast<double*>(w + h->ws_xstore);
    v.slots = reinterpret_cast<double*>(w + h->ws_slots);
    v.sysbuf = reinterpret_cast<double*>(w + h->ws_sysbuf);
    v.counters = reinterpret_cast<int*>(w + h->ws_counters);
    return v;
}

__device__ __forceinline__ Geom geom_of(const DevView& v, const GeomDesc& gd) {
    Geom g;
    g.M = gd.M; g.h = gd.h;
    g.a = v.a + gd.reg_off; g.d = v.d + gd.reg_off; g.nmk = v.nmk + gd.reg_off;
    return g;
}

// geometry-stage store layout (doubles): HT[N][ldn = pad8(N+2R)] | w[R*N] | u[R*N] | s0[R*R] | phi[R] | bc[R] | status
// ldn = N rounded up to 8 so that rows are 64-byte aligned for the cp.async operand staging.
__host__ __device__ inline int pad8(int n) { return (n + 7) & ~7; }
// Columns N .. N+R-1 of HT hold u'_i = u_i + bc_i e_0 and columns N+R .. N+2R-1 hold w_j, so that the frequency
// stage's T = H Imk GEMM also yields the right-hand sides and q_j = Imk^T w_j (rows N .. N+2R-1 of its output).
__host__ __device__ inline int gs_ldn(int N, int R) { return pad8(N + 2 * R); }
__host__ __device__ inline int64_t gs_base_doubles(int N, int R) {
    return (((int64_t)N * gs_ldn(N, R) + 2LL * R * N + (int64_t)R * R + 2 * R + 2) + 7) & ~7LL;
}
// ... followed by the factorised form of H (k_hchol): LA[N][ldn] | perm[N] (int32).  status is followed by hrank.
// H is symmetric positive semidefinite (tools/experiments/exp_hchol.py): with its zero rows permuted to the end,
// Pm^T H Pm = L L^T (L: N x hrank, lower trapezoidal) reproduces it to rounding.  Row k of LA belongs to harmonic
// perm[k]: columns 0 .. 2R-1 = [u'_i | w_j] of that harmonic, columns 2R .. 2R+hrank-1 = L[k][:].  With it
// P = Imk^T H Imk - Re D = W^T W - Re D, W = L^T (Pm^T Imk): a triangular product (half the flops of H Imk) followed by
// a symmetric rank-N update.  hrank <= 0: factorisation not available (H not PSD to rounding, or N too large for the
// shared-memory factor) and the frequency stage uses H itself.
// LAT: LA re-tiled for the chunk stream of the triangular product (cta_trmm_stream): tile -1 = columns 0..7 of LA (the
// aug columns), rows 0..N-1; tile a >= 0 = columns 2R + 32 a .. + 31, rows 32 a .. N-1; every tile row-major with a
// leading dimension of 36 doubles, so that kTrKC rows of a tile are ONE contiguous bulk copy into the padded stage.
__host__ __device__ inline int64_t lat_tile_off(int N, int a) {          // doubles, a = -1 .. ng; rows zero-padded to pad8(N)
    const int Np = pad8(N);
    int64_t rows = 0;
    if (a >= 0) rows = Np;
    for (int i = 0; i < a; ++i) rows += Np - 32 * i;
    return rows * 36;
}
__host__ __device__ inline int64_t lat_doubles(int N, int K, int R) { return trmm_fits(N, K, R) ? lat_tile_off(N, (N + 31) / 32) : 0; }   // only where the chunk stream runs
__host__ __device__ inline int64_t gs_lat_off(int N, int R) { return gs_base_doubles(N, R) + (int64_t)N * gs_ldn(N, R) + pad8((N + 1) / 2); }
__host__ __device__ inline int64_t gs_doubles(int N, int R, int K) {
    return gs_lat_off(N, R) + lat_doubles(N, K, R);
}
__host__ __device__ inline size_t hchol_smem_bytes(int N) {
    return chol_smem_bytes(N, 0) + sizeof(double) * (size_t)tri_doubles(N) + sizeof(int) * (size_t)(N + 8);
}

struct OutPtrs {
    double *am, *dp, *force, *phase, *m0;
    int32_t* status;
    double* x;
};

// c vectors (meem_engine.py:65-98) of every radiating mode of every geometry: out[g][j][p_g], frequency independent.
__global__ void k_cvec(const void* blob, void* ws, double* out) {
    DevView v = make_view(blob, ws);
    __shared__ Layout L;
    const GeomDesc gd = v.geom[blockIdx.x];
    const Geom g = geom_of(v, gd);
    if (threadIdx.x == 0) make_layout(g, L);
    __syncthreads();
    for (int j = 0; j < gd.R; ++j)
        fill_c<double>(g, L, v.tables + gd.tab_off, v.flags + gd.flag_off + (int64_t)j * (gd.M + 1),
                       out + gd.out_c + (int64_t)j * gd.p, threadIdx.x, blockDim.x);
}

__device__ __forceinline__ double warp_sum_all(double v) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}

// Reduction of ONE given solution X (N_T complex) of system (g, f) to the reference's per-mode dicts
// (compute_hydrodynamic_coefficients, meem_engine.py:326-393): out[j][0..3] = real (added mass), imag (damping),
// excitation phase, excitation force for every radiating mode j of the plan's geometry.  One warp per mode.
__global__ void k_reduce(const void* blob, void* ws, int gi, int f, const cplx* X, double* out) {
    DevView v = make_view(blob, ws);
    __shared__ Layout L;
    const GeomDesc gd = v.geom[gi];
    const Geom g = geom_of(v, gd);
    if (threadIdx.x == 0) make_layout(g, L);
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    const RootSet rs = v.sets[gd.root_set];
    const double m0 = v.roots[rs.off + f];
    const double* tab = v.tables + gd.tab_off;
    const int M = gd.M;
    for (int j = warp; j < gd.R; j += nwarp) {
        const uint8_t* hv = v.flags + gd.flag_off + (int64_t)j * (M + 1);
        double re = 0.0, im = 0.0;
        for (int i = 0; i < M; ++i) {
            if (!hv[i]) continue;
            const int N = g.nmk[i];
            for (int n = lane; n < (i == 0 ? 1 : 2) * N; n += 32) {
                const double c = tab[L.tab[i] + 8 * N + n];          // int_R1 * z_n | int_R2 * z_n
                const cplx x = X[L.col[i] + n];
                re = fma(c, x.re, re); im = fma(c, x.im, im);
            }
        }
        re = warp_sum_all(re); im = warp_sum_all(im);
        if (lane == 0) {
            const bool inf_m0 = isinf(m0);
            const double phi = phi_sum(g, L, tab, hv);
            const double cre = 2.0 * kPi * (re + phi), cim = 2.0 * kPi * im;
            const double dpv = inf_m0 ? 0.0 : cim * omega_of_m0(m0, g.h) * v.hdr->rho;
            const cplx xe = X[gd.p];
            out[4 * j + 0] = cre * v.hdr->rho;
            out[4 * j + 1] = dpv;
            out[4 * j + 2] = inf_m0 ? -(kPi / 2.0) : -(kPi / 2.0) + atan2(xe.im, xe.re) - hankel0_angle(m0 * g.a[M - 1]);
            out[4 * j + 3] = excitation_force(dpv, m0, g.h, kRhoDefault);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// roots
__global__ void k_m0(const void* blob, void* ws) {
    DevView v = make_view(blob, ws);
    const int F = v.hdr->n_freq;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= v.hdr->n_sets * F) return;
    const int s = t / F, f = t - s * F;
    const RootSet rs = v.sets[s];
    const double fr = v.freq[f];
    const double m0 = (v.hdr->freq_kind == MEEMB200_FREQ_M0) ? fr : wavenumber_mode(fr, rs.h, v.hdr->root_mode);
    v.roots[rs.off + f] = m0;
}

__global__ void k_mk(const void* blob, void* ws) {
    DevView v = make_view(blob, ws);
    const int F = v.hdr->n_freq;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    for (int s = blockIdx.y; s < v.hdr->n_sets; s += gridDim.y) {      // gridDim.y is capped at 65 535 root sets per pass
        const RootSet rs = v.sets[s];
        if (t >= F * rs.K) continue;
        const int f = t / rs.K, k = t - f * rs.K;
        const double m0 = v.roots[rs.off + f];
        const double mk = mk_entry(k, m0, rs.h, v.hdr->root_mode);
        double* mkp = v.roots + rs.off + F;
        double* nkp = mkp + (int64_t)F * rs.K;
        mkp[(int64_t)f * rs.K + k] = mk;
        nkp[(int64_t)f * rs.K + k] = nk_entry(k, m0, rs.h, mk);
    }
}

__global__ void k_roots_flat(int n, int K, const double* freq, const double* h, int freq_kind, int root_mode,
                             double* m0o, double* mko, double* nko) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * K) return;
    const int i = t / K, k = t - i * K;
    const double m0 = (freq_kind == MEEMB200_FREQ_M0) ? freq[i] : wavenumber_mode(freq[i], h[i], root_mode);
    if (k == 0) m0o[i] = m0;
    const double mk = mk_entry(k, m0, h[i], root_mode);
    mko[(int64_t)i * K + k] = mk;
    nko[(int64_t)i * K + k] = nk_entry(k, m0, h[i], mk);
}

__global__ void k_bessel(int kind, int64_t n, const double* x, double* out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double j0, y0, j1, y1;
    switch (kind) {
        case 0: out[i] = i0e(x[i]); break;
        case 1: out[i] = i1e(x[i]); break;
        case 2: out[i] = k0e(x[i]); break;
        case 3: out[i] = k1e(x[i]); break;
        default:
            bessel_jy01(x[i], &j0, &y0, &j1, &y1);
            out[i] = kind == 4 ? j0 : kind == 5 ? y0 : kind == 6 ? j1 : y1;
    }
}

// ------------------------------------------------------------------------------------------------
// per-geometry tables
__global__ void k_tables(const void* blob, void* ws) {
    DevView v = make_view(blob, ws);
    __shared__ Layout L;
    const GeomDesc gd = v.geom[blockIdx.x];
    const Geom g = geom_of(v, gd);
    if (threadIdx.x == 0) make_layout(g, L);
    __syncthreads();
    fill_tables(g, L, v.tables + gd.tab_off, threadIdx.x, blockDim.x);
}

// ------------------------------------------------------------------------------------------------
// shared helpers
__device__ __forceinline__ int next_item(int* counter, int* sh) {
    __syncthreads();
    if (threadIdx.x == 0) *sh = atomicAdd(counter, 1);
    __syncthreads();
    return *sh;
}

template <class T>
__device__ __forceinline__ void cta_zero(T* p, int64_t n) {
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) p[i] = zero_of(T());
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}

// ------------------------------------------------------------------------------------------------
// Schur path, geometry stage
template <int NB, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_geom(const void* blob, void* ws, int n_slots_unused) {
    constexpr bool SMALL = MINB > 1;
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ Layout L;
    __shared__ int sh_item, sh_status;
    DevView v = make_view(blob, ws);
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    double* W = v.slots + (int64_t)blockIdx.x * v.hdr->slot_doubles_geom;
    for (;;) {
        const int item = next_item(&v.counters[0], &sh_item);
        if (item >= v.hdr->n_geom) break;
        const int gi = v.order[item];
        const GeomDesc gd = v.geom[gi];
        const Geom g = geom_of(v, gd);
        if (tid == 0) { make_layout(g, L); sh_status = 0; }
        __syncthreads();
        const int p = gd.p, N = gd.N, R = gd.R, M = gd.M;
        const int ncols = p + N + R;
        const long ld = ncols;
        const double* tab = v.tables + gd.tab_off;
        const uint8_t* flags = v.flags + gd.flag_off;
        cta_zero(W, (int64_t)p * ld);
        __syncthreads();
        fill_A_static<double>(g, L, tab, W, ld, tid, nthr);
        for (int n = tid; n < N; n += nthr) W[(long)(L.pot_row[M - 1] + n) * ld + p + n] = 1.0;   // E
        for (int i = 0; i < R; ++i) fill_b_static<double>(g, L, flags + (int64_t)i * (M + 1), W + p + N + i, ld, tid, nthr);
        __syncthreads();
        lu_forward<double, NB, SMALL>(W, ld, p, p, ncols, smem, &sh_status);
        lu_backward<double, NB, SMALL>(W, ld, p, p, ncols, smem, p);
        __syncthreads();
        // reduce to what the frequency stage needs
        double* gs = v.gstore + gd.gs_off;
        double* HT = gs;
        const int ldn = gs_ldn(N, R);
        double* wv = HT + (int64_t)N * ldn;
        double* uv = wv + (int64_t)R * N;
        double* s0 = uv + (int64_t)R * N;
        double* phi = s0 + R * R;
        double* bc = phi + R;
        double* st = bc + R;
        const int bd = M - 1;
        const double* TL = tab + L.tab[bd];
        const int c1L = L.col[bd], c2L = c1L + N;
        // H[n][n'] = R1'(a)[n] G[c1L+n][n'] + R2'(a)[n] G[c2L+n][n'];  stored transposed
        for (int e = tid; e < N * N; e += nthr) {
            const int n = e / N, n2 = e - n * N;
            double hv = TL[2 * N + n] * W[(long)(c1L + n) * ld + p + n2];
            if (bd > 0) hv += TL[3 * N + n] * W[(long)(c2L + n) * ld + p + n2];
            HT[(int64_t)n2 * ldn + n] = hv;
        }
        for (int e = tid; e < R * N; e += nthr) {
            const int i = e / N, n = e - i * N;
            double u = TL[2 * N + n] * W[(long)(c1L + n) * ld + p + N + i];
            if (bd > 0) u += TL[3 * N + n] * W[(long)(c2L + n) * ld + p + N + i];
            uv[e] = u;
        }
        // w_j[n'] = sum_col c_j[col] G[col][n'],  s0[j][i] = sum_col c_j[col] y0_i[col]: one warp per (j, column)
        for (int t = warp; t < R * (N + R); t += nwarp) {
            const int j = t / (N + R), cidx = t - j * (N + R);
            const uint8_t* hvj = flags + (int64_t)j * (M + 1);
            double acc = 0.0;
            for (int reg = 0; reg < M; ++reg) {
                if (!hvj[reg]) continue;
                const int Nr = g.nmk[reg];
                const double* T8 = tab + L.tab[reg] + 8 * Nr;
                for (int n = lane; n < Nr; n += 32) {
                    acc += T8[n] * W[(long)(L.col[reg] + n) * ld + p + cidx];
                    if (reg > 0) acc += T8[Nr + n] * W[(long)(L.col[reg] + Nr + n) * ld + p + cidx];
                }
            }
            acc = warp_sum(acc);
            if (lane == 0) {
                if (cidx < N) wv[j * N + cidx] = acc;
                else s0[j * R + (cidx - N)] = acc;
            }
        }
        for (int j = tid; j < R; j += nthr) {
            phi[j] = phi_sum(g, L, tab, flags + (int64_t)j * (M + 1));
            bc[j] = b_end_constant(g, flags + (int64_t)j * (M + 1));
        }
        if (tid == 0) st[0] = (double)sh_status;
        __syncthreads();
        for (int e = tid; e < N * (ldn - N); e += nthr) {
            const int n = e / (ldn - N), c = e - n * (ldn - N);
            double val = 0.0;
            if (c < R) val = uv[c * N + n] + (n == 0 ? bc[c] : 0.0);
            else if (c < 2 * R) val = wv[(c - R) * N + n];
            HT[(int64_t)n * ldn + N + c] = val;
        }
        if (gd.xg_off >= 0) {     // keep G and y0 for X reconstruction
            double* xg = v.xstore + gd.xg_off;
            for (int64_t e = tid; e < (int64_t)p * (N + R); e += nthr) {
                const int r = (int)(e / (N + R)), c = (int)(e - (int64_t)r * (N + R));
                xg[e] = W[(long)r * ld + p + c];
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Schur path, geometry stage by window elimination (meem_geom_band.cuh): one CTA per geometry, the window in shared
// memory (SMEMWIN: dynamic smem = LU scratch | window | hand-over buffer) or in the CTA's L2-resident global slot.
// Used whenever X is not requested (the X reconstruction needs all of G, which only the dense elimination forms).
constexpr int kBandSmallChunk = 72;      // U12 staging chunk of the small variant (>= the 65 window columns of N = 16, R = 1)
__host__ __device__ inline int band_ld(int wcols, bool smem) { return smem ? (wcols | 1) : ((wcols + 1) & ~1); }
template <int NB, int THREADS, int MINB, bool SMALL, bool SMEMWIN, int CHUNK>
__global__ void __launch_bounds__(THREADS, MINB) k_geom_band(const void* blob, void* ws, int wrows_max, int wcols_max, int side_alias,
                                                             int use_static) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ Layout L;
    __shared__ int sh_item, sh_status;
    DevView v = make_view(blob, ws);
    const int tid = threadIdx.x, nthr = blockDim.x;
    const long ld = band_ld(wcols_max, SMEMWIN);
    const size_t lu_bytes = (lu_smem_bytes<double, NB, CHUNK>(wrows_max) + 15) & ~(size_t)15;
    double* W = SMEMWIN ? reinterpret_cast<double*>(smem + lu_bytes) : v.slots + (int64_t)blockIdx.x * v.hdr->slot_doubles_geom;
    // the hand-over buffer between two windows is only live while no elimination runs: it may share the LU scratch
    double* side = side_alias ? reinterpret_cast<double*>(smem) : W + (long)wrows_max * ld;
    for (;;) {
        const int item = next_item(&v.counters[0], &sh_item);
        if (item >= v.hdr->n_geom) break;
        const int gi = v.order[item];
        const GeomDesc gd = v.geom[gi];
        const Geom g = geom_of(v, gd);
        if (tid == 0) { make_layout(g, L); sh_status = 0; }
        __syncthreads();
        const int N = gd.N, R = gd.R, M = gd.M;
        const double* tab = v.tables + gd.tab_off;
        const uint8_t* flags = v.flags + gd.flag_off;
        band_eliminate<NB, SMALL, CHUNK>(g, L, tab, flags, R, W, ld, side, smem, &sh_status, use_static != 0);
        // reduce to what the frequency stage needs (same store layout as k_geom)
        double* gs = v.gstore + gd.gs_off;
        double* HT = gs;
        const int ldn = gs_ldn(N, R);
        double* wv = HT + (int64_t)N * ldn;
        double* uv = wv + (int64_t)R * N;
        double* s0 = uv + (int64_t)R * N;
        double* phi = s0 + R * R;
        double* bc = phi + R;
        double* st = bc + R;
        const int bd = M - 1;
        const int n0 = band_unknowns(g, bd);
        const double* TL = tab + L.tab[bd];
        // H[n][n'] = R1'(a)[n] G[C1 row n][n'] + R2'(a)[n] G[C2 row n][n'];  stored transposed
        for (int e = tid; e < N * N; e += nthr) {
            const int n2 = e / N, n = e - n2 * N;
            double hv = TL[2 * N + n] * W[(long)n * ld + n0 + n2];
            if (bd > 0) hv += TL[3 * N + n] * W[(long)(N + n) * ld + n0 + n2];
            HT[(int64_t)n2 * ldn + n] = hv;
        }
        for (int e = tid; e < R * N; e += nthr) {
            const int i = e / N, n = e - i * N;
            double u = TL[2 * N + n] * W[(long)n * ld + n0 + N + i];
            if (bd > 0) u += TL[3 * N + n] * W[(long)(N + n) * ld + n0 + N + i];
            uv[e] = u;
            wv[e] = -W[(long)(n0 + i) * ld + n0 + n];
        }
        for (int e = tid; e < R * R; e += nthr) {
            const int j = e / R, i = e - j * R;
            s0[e] = -W[(long)(n0 + j) * ld + n0 + N + i];
        }
        for (int j = tid; j < R; j += nthr) {
            phi[j] = phi_sum(g, L, tab, flags + (int64_t)j * (M + 1));
            bc[j] = b_end_constant(g, flags + (int64_t)j * (M + 1));
        }
        if (tid == 0) st[0] = (double)sh_status;
        __syncthreads();
        for (int e = tid; e < N * (ldn - N); e += nthr) {
            const int n = e / (ldn - N), c = e - n * (ldn - N);
            double val = 0.0;
            if (c < R) val = uv[c * N + n] + (n == 0 ? bc[c] : 0.0);
            else if (c < 2 * R) val = wv[(c - R) * N + n];
            HT[(int64_t)n * ldn + N + c] = val;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Schur path, geometry stage, grid-cooperative variant (few large geometries): see meem_geom_coop.cuh.
// Slot 0 holds W [p][ld] followed by the panel scratch (Lbuf [p][16], L11 [256], moves).
__host__ __device__ inline int64_t coop_ld(int p, int N, int R) { return (p + N + R + 1) & ~1; }
__host__ __device__ inline int64_t coop_slot_doubles(int p, int N, int R) {
    return (int64_t)p * coop_ld(p, N, R) + (int64_t)p * kCoopNB + kCoopNB * kCoopNB + 2 * kCoopMaxMoves + 16;
}
__host__ __device__ inline size_t coop_smem_bytes(int p) { return sizeof(double) * (size_t)kCoopNB * (size_t)((p + 31) & ~31); }

__global__ void __launch_bounds__(kCoopThreads, 1) k_geom_coop(const void* blob, void* ws) {
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ Layout L;
    __shared__ CoopShared S;
    DevView v = make_view(blob, ws);
    double* dsm = reinterpret_cast<double*>(smem);
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int lane = tid & 31;
    const long gtid = (long)blockIdx.x * nthr + tid, gthr = (long)gridDim.x * nthr;
    const long gwarp = gtid >> 5, gnwarp = gthr >> 5;
    unsigned long long* cprof = reinterpret_cast<unsigned long long*>(v.counters + 16) + 8;
    for (int item = 0; item < v.hdr->n_geom; ++item) {
        const int gi = v.order[item];
        const GeomDesc gd = v.geom[gi];
        const Geom g = geom_of(v, gd);
        if (tid == 0) { make_layout(g, L); S.status = 0; }
        __syncthreads();
        const int p = gd.p, N = gd.N, R = gd.R, M = gd.M;
        const int ncols = p + N + R;
        const long ld = coop_ld(p, N, R);
        double* W = v.slots;
        CoopScratch sc;
        sc.Lbuf = W + (int64_t)p * ld;
        sc.L11g = sc.Lbuf + (int64_t)p * kCoopNB;
        sc.moves = reinterpret_cast<int*>(sc.L11g + kCoopNB * kCoopNB);
        const double* tab = v.tables + gd.tab_off;
        const uint8_t* flags = v.flags + gd.flag_off;
        double* gs = v.gstore + gd.gs_off;
        const int ldn = gs_ldn(N, R);
        double* HT = gs;
        double* wv = HT + (int64_t)N * ldn;
        double* uv = wv + (int64_t)R * N;
        double* s0 = uv + (int64_t)R * N;
        double* phi = s0 + R * R;
        double* bc = phi + R;
        double* st = bc + R;
        for (int64_t i = gtid; i < (int64_t)p * ld; i += gthr) W[i] = 0.0;
        if (gtid == 0) st[0] = 0.0;
        grid.sync();
        fill_A_static<double>(g, L, tab, W, ld, (int)gtid, (int)gthr);
        for (int n = (int)gtid; n < N; n += (int)gthr) W[(long)(L.pot_row[M - 1] + n) * ld + p + n] = 1.0;   // E
        for (int i = 0; i < R; ++i) fill_b_static<double>(g, L, flags + (int64_t)i * (M + 1), W + p + N + i, ld, (int)gtid, (int)gthr);
        grid.sync();
        const int prow = (p + 31) & ~31;
        int pidx = 0;
        for (int k0 = 0; k0 < p; k0 += kCoopNB, ++pidx) {
            const int nb = min(kCoopNB, p - k0);
            const int owner = pidx % gridDim.x;
            long long tq0 = clock64();
            if ((int)blockIdx.x == owner) {
                const int rows = p - k0;
                if (rows <= kCoopThreads) coop_panel_factor_reg<1>(W, ld, k0, nb, rows, rows, S, sc);
                else if (rows <= 2 * kCoopThreads) coop_panel_factor_reg<2>(W, ld, k0, nb, rows, rows, S, sc);
                else coop_panel_factor(W, ld, k0, nb, rows, rows, dsm, prow, S, sc);
                __syncthreads();
                if (tid == 0 && S.status) { st[0] = (double)((int)st[0] | S.status); S.status = 0; }
                if (tid == 0) atomicAdd(&cprof[0], (unsigned long long)(clock64() - tq0));
            }
            grid.sync();
            long long tq1 = clock64();
            if (gtid == 32) cprof[1] += (unsigned long long)(tq1 - tq0);          // factor + barrier, seen by a non-owner warp of CTA 0
            coop_panel_update(W, ld, p, ncols, k0, nb, S, sc, cprof + 6);
            __syncthreads();
            long long tq2 = clock64();
            if (gtid == 32) cprof[2] += (unsigned long long)(tq2 - tq1);          // update (CTA 0's share)
            grid.sync();
            if (gtid == 32) cprof[3] += (unsigned long long)(clock64() - tq2);    // barrier after update
        }
        long long tq3 = clock64();
        coop_backsolve(W, ld, p, N + R, dsm, S);
        __syncthreads();
        if (gtid == 32) cprof[4] += (unsigned long long)(clock64() - tq3);
        grid.sync();
        if (gtid == 32) cprof[5] += (unsigned long long)(clock64() - tq3);
        // reduce to what the frequency stage needs (same quantities as k_geom, grid-strided)
        const int bd = M - 1;
        const double* TL = tab + L.tab[bd];
        const int c1L = L.col[bd], c2L = c1L + N;
        for (long e = gtid; e < (long)N * N; e += gthr) {
            const int n = (int)(e / N), n2 = (int)(e - (long)n * N);
            double hv = TL[2 * N + n] * W[(long)(c1L + n) * ld + p + n2];
            if (bd > 0) hv += TL[3 * N + n] * W[(long)(c2L + n) * ld + p + n2];
            HT[(int64_t)n2 * ldn + n] = hv;
        }
        for (long e = gtid; e < (long)R * N; e += gthr) {
            const int i = (int)(e / N), n = (int)(e - (long)i * N);
            double u = TL[2 * N + n] * W[(long)(c1L + n) * ld + p + N + i];
            if (bd > 0) u += TL[3 * N + n] * W[(long)(c2L + n) * ld + p + N + i];
            uv[e] = u;
        }
        for (long t = gwarp; t < (long)R * (N + R); t += gnwarp) {
            const int j = (int)(t / (N + R)), cidx = (int)(t - (long)j * (N + R));
            const uint8_t* hvj = flags + (int64_t)j * (M + 1);
            double acc = 0.0;
            for (int reg = 0; reg < M; ++reg) {
                if (!hvj[reg]) continue;
                const int Nr = g.nmk[reg];
                const double* T8 = tab + L.tab[reg] + 8 * Nr;
                for (int n = lane; n < Nr; n += 32) {
                    acc += T8[n] * W[(long)(L.col[reg] + n) * ld + p + cidx];
                    if (reg > 0) acc += T8[Nr + n] * W[(long)(L.col[reg] + Nr + n) * ld + p + cidx];
                }
            }
            acc = warp_sum(acc);
            if (lane == 0) {
                if (cidx < N) wv[j * N + cidx] = acc;
                else s0[j * R + (cidx - N)] = acc;
            }
        }
        for (long j = gtid; j < R; j += gthr) {
            phi[j] = phi_sum(g, L, tab, flags + (int64_t)j * (M + 1));
            bc[j] = b_end_constant(g, flags + (int64_t)j * (M + 1));
        }
        if (gd.xg_off >= 0) {
            double* xg = v.xstore + gd.xg_off;
            for (int64_t e = gtid; e < (int64_t)p * (N + R); e += gthr) {
                const int r = (int)(e / (N + R)), c = (int)(e - (int64_t)r * (N + R));
                xg[e] = W[(long)r * ld + p + c];
            }
        }
        grid.sync();
        for (long e = gtid; e < (long)N * (ldn - N); e += gthr) {
            const int n = (int)(e / (ldn - N)), c = (int)(e - (long)n * (ldn - N));
            double val = 0.0;
            if (c < R) val = uv[c * N + n] + (n == 0 ? bc[c] : 0.0);
            else if (c < 2 * R) val = wv[(c - R) * N + n];
            HT[(int64_t)n * ldn + N + c] = val;
        }
        grid.sync();
    }
}

// ------------------------------------------------------------------------------------------------
// Schur path, geometry stage, grid-cooperative WINDOW elimination with static pivots (few large geometries whose
// depths decrease outwards, i.e. every boundary is "case 1" of meem_engine.py:170-217 - big_tricylinder).
// Same window recursion as k_geom_band, but inside window bd the unknowns C1^bd are eliminated FIRST, with the potential
// rows of boundary bd as pivot rows: their C1 block is the diagonal matrix (h - d_bd) diag R_1n(a_bd) = (h - d_bd) {1/2, 1,
// 1, ...}, so that step needs no pivot search at all and is one grid-wide DMMA product (coop2_static_update).  Only the
// C2^bd block (N_bd columns among the N_bd left-over + N_bd+1 velocity rows) goes through the partial-pivot panel
// machinery of k_geom_coop: (M - 1) N sequential pivot columns instead of (2M - 1) N, on windows instead of the p x p
// matrix.  Numerical probe against LAPACK's partial pivoting: tools/experiments/exp_static.py (<= 3e-15 on H, u, w, s0).
// Window layout:  rows [left over (nc) | velocity bd (hv) | c rows (R) | potential bd (Nb)],
//                 cols [C2^bd (n2) | region bd+1: C1, C2 (or E in the last window) (n1) | b (R) | pad | C1^bd (Nb)].
struct Coop2Dims { int wrows, wcols; };
MEEM_HDX void coop2_dims(const Geom& g, const Layout& L, int R, Coop2Dims& D) {
    D.wrows = D.wcols = 0;
    int nc = 0;
    for (int bd = 0; bd < g.M; ++bd) {
        const bool last = bd == g.M - 1;
        const int Nb = g.nmk[bd], n2 = bd > 0 ? Nb : 0, n1 = last ? g.nmk[g.M - 1] : 2 * g.nmk[bd + 1];
        const int hv = last ? 0 : g.nmk[bd + 1];
        const int rows = nc + hv + R + Nb, cols = ((n2 + n1 + R + 1) & ~1) + Nb;
        if (rows > D.wrows) D.wrows = rows;
        if (cols > D.wcols) D.wcols = cols;
        nc = hv;
    }
}
__host__ __device__ inline int64_t coop2_ld(int wcols) { return (wcols + 1) & ~1; }
__host__ __device__ inline int64_t coop2_slot_doubles(int wrows, int wcols) {
    return 2 * (int64_t)wrows * coop2_ld(wcols) + (int64_t)wrows * kCoopNB + kCoopNB * kCoopNB + 2 * kCoopMaxMoves + 16;
}
__host__ __device__ inline size_t coop2_smem_bytes(int wrows, int maxN) {
    const size_t a = (size_t)kCoopNB * (size_t)((wrows + 31) & ~31), b = 2 * (size_t)((maxN + 15) & ~15) + 16, c = (size_t)maxN + 16;
    return sizeof(double) * (a > b ? (a > c ? a : c) : (b > c ? b : c));
}

__global__ void __launch_bounds__(kCoopThreads, 1) k_geom_coop2(const void* blob, void* ws, int wrows_max, int wcols_max) {
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ Layout L;
    __shared__ CoopShared S;
    DevView v = make_view(blob, ws);
    double* dsm = reinterpret_cast<double*>(smem);
    const int tid = threadIdx.x, nthr = blockDim.x;
    const long gtid = (long)blockIdx.x * nthr + tid, gthr = (long)gridDim.x * nthr;
    const long ld = coop2_ld(wcols_max);
    double* Wb[2] = {v.slots, v.slots + (int64_t)wrows_max * ld};
    CoopScratch sc;
    sc.Lbuf = v.slots + 2 * (int64_t)wrows_max * ld;
    sc.L11g = sc.Lbuf + (int64_t)wrows_max * kCoopNB;
    sc.moves = reinterpret_cast<int*>(sc.L11g + kCoopNB * kCoopNB);
    for (int item = 0; item < v.hdr->n_geom; ++item) {
        const int gi = v.order[item];
        const GeomDesc gd = v.geom[gi];
        const Geom g = geom_of(v, gd);
        if (tid == 0) { make_layout(g, L); S.status = 0; }
        __syncthreads();
        const int N = gd.N, R = gd.R, M = gd.M;
        const double* tab = v.tables + gd.tab_off;
        const uint8_t* flags = v.flags + gd.flag_off;
        double* gs = v.gstore + gd.gs_off;
        const int ldn = gs_ldn(N, R);
        double* HT = gs;
        double* wv = HT + (int64_t)N * ldn;
        double* uv = wv + (int64_t)R * N;
        double* s0 = uv + (int64_t)R * N;
        double* phi = s0 + R * R;
        double* bc = phi + R;
        double* st = bc + R;
        if (gtid == 0) st[0] = 0.0;
        int nc = 0, n2o = 0, srows_o = 0;
        for (int bd = 0; bd < M; ++bd) {
            const bool last = bd == M - 1;
            const int Nb = g.nmk[bd], n2 = bd > 0 ? Nb : 0, n1 = last ? N : 2 * g.nmk[bd + 1];
            const int hv = last ? 0 : g.nmk[bd + 1];
            const int srows = nc + hv, nsub = srows + R, ncs = n2 + n1 + R, c1L = (ncs + 1) & ~1, potr0 = nsub;
            const int ncols_w = c1L + Nb, nrows_w = nsub + Nb;
            double* W = Wb[bd & 1];
            const double* Wo = Wb[(bd & 1) ^ 1];
            for (long e = gtid; e < (long)nrows_w * ncols_w; e += gthr) W[(e / ncols_w) * ld + (e % ncols_w)] = 0.0;
            // reciprocal static pivots (every CTA keeps its own copy): D1_n = (h - d_bd) R_1n(a_bd)
            {
                const double hdL = g.h - g.d[bd];
                const double* TL = tab + L.tab[bd];
                for (int n = tid; n < Nb; n += nthr) dsm[n] = 1.0 / (hdL * TL[n]);
            }
            grid.sync();
            if (bd > 0) {      // left-over rows and c rows of the previous window: columns [C1^bd | C2^bd | b] there
                const int sw = 2 * Nb + R;
                for (long e = gtid; e < (long)(nc + R) * sw; e += gthr) {
                    const int r = (int)(e / sw), c = (int)(e % sw);
                    const int ro = r < nc ? n2o + r : srows_o + (r - nc), rn = r < nc ? r : srows + (r - nc);
                    const int cn = c < Nb ? c1L + c : (c < 2 * Nb ? c - Nb : n2 + n1 + (c - 2 * Nb));
                    W[(long)rn * ld + cn] = Wo[(long)ro * ld + n2o + c];
                }
            } else {
                for (long e = gtid; e < (long)R * Nb; e += gthr) {
                    const int j = (int)(e / Nb), n = (int)(e % Nb);
                    W[(long)(srows + j) * ld + c1L + n] = c_entry(g, L, tab, flags + (long)j * (M + 1), 0, 0, n);
                }
            }
            if (!last) {
                const int Nr = g.nmk[bd + 1];
                for (long e = gtid; e < (long)R * n1; e += gthr) {
                    const int j = (int)(e / n1), c = (int)(e % n1);
                    W[(long)(srows + j) * ld + n2 + c] =
                        c_entry(g, L, tab, flags + (long)j * (M + 1), bd + 1, c < Nr ? 0 : 1, c < Nr ? c : c - Nr);
                }
            } else {
                for (long n = gtid; n < n1; n += gthr) W[(long)(potr0 + n) * ld + n2 + n] = 1.0;     // E
            }
            fill_A_boundary_at<double>(g, L, tab, bd, W, ld, potr0, nc, c1L, 0, n2, (int)gtid, (int)gthr);
            for (long e = gtid; e < (long)R * (Nb + hv); e += gthr) {
                const int i = (int)(e / (Nb + hv)), n = (int)(e % (Nb + hv));
                const bool vel = n >= Nb;
                W[(long)(vel ? nc + (n - Nb) : potr0 + n) * ld + n2 + n1 + i] =
                    b_static_entry(g, L, flags + (long)i * (M + 1), bd, vel, vel ? n - Nb : n);
            }
            grid.sync();
            coop2_static_update(W, ld, nsub, ncs, c1L, potr0, Nb, dsm);
            grid.sync();
            int pidx = 0;
            for (int k0 = 0; k0 < n2; k0 += kCoopNB, ++pidx) {
                const int nb = min(kCoopNB, n2 - k0);
                const int owner = (pidx + 7 * bd) % gridDim.x;
                if ((int)blockIdx.x == owner) {
                    const int rows = nsub - k0, sr = srows - k0;
                    if (rows <= kCoopThreads) coop_panel_factor_reg<1>(W, ld, k0, nb, rows, sr, S, sc);
                    else if (rows <= 2 * kCoopThreads) coop_panel_factor_reg<2>(W, ld, k0, nb, rows, sr, S, sc);
                    else coop_panel_factor(W, ld, k0, nb, rows, sr, dsm, (wrows_max + 31) & ~31, S, sc);
                    __syncthreads();
                    if (tid == 0 && S.status) { st[0] = (double)((int)st[0] | S.status); S.status = 0; }
                }
                grid.sync();
                coop_panel_update(W, ld, nsub, ncs, k0, nb, S, sc, nullptr);
                __syncthreads();
                grid.sync();
            }
            if (last) {
                if (n2 > 0) {
                    coop_backsolve(W, ld, n2, N + R, dsm, S);
                    __syncthreads();
                    grid.sync();
                }
                // rows of the last region of [G | y0]: C2 rows = rows [0, N) of the window, C1 rows from their pivot rows:
                // C1_n = (rhs_n - D2_n C2_n) / D1_n.  H[n][n'] = R1'(a)[n] G[C1 n][n'] + R2'(a)[n] G[C2 n][n'], stored transposed
                const double* TL = tab + L.tab[bd];
                const double hdL = g.h - g.d[bd];
                for (long e = gtid; e < (long)N * (N + R); e += gthr) {
                    const int cidx = (int)(e / N), n = (int)(e % N);
                    const double d1 = hdL * TL[n];
                    double g2 = 0.0, rhs = W[(long)(potr0 + n) * ld + n2 + cidx];
                    if (bd > 0) { g2 = W[(long)n * ld + n2 + cidx]; rhs -= hdL * TL[N + n] * g2; }
                    double hv2 = TL[2 * N + n] * (rhs / d1);
                    if (bd > 0) hv2 += TL[3 * N + n] * g2;
                    if (cidx < N) HT[(int64_t)cidx * ldn + n] = hv2;
                    else uv[(cidx - N) * N + n] = hv2;
                }
                for (long e = gtid; e < (long)R * (N + R); e += gthr) {
                    const int j = (int)(e / (N + R)), cidx = (int)(e % (N + R));
                    const double val = -W[(long)(srows + j) * ld + n2 + cidx];
                    if (cidx < N) wv[j * N + cidx] = val;
                    else s0[j * R + (cidx - N)] = val;
                }
                for (long j = gtid; j < R; j += gthr) {
                    phi[j] = phi_sum(g, L, tab, flags + (int64_t)j * (M + 1));
                    bc[j] = b_end_constant(g, flags + (int64_t)j * (M + 1));
                }
                grid.sync();
                for (long e = gtid; e < (long)N * (ldn - N); e += gthr) {
                    const int n = (int)(e / (ldn - N)), c = (int)(e - (long)n * (ldn - N));
                    double val = 0.0;
                    if (c < R) val = uv[c * N + n] + (n == 0 ? bc[c] : 0.0);
                    else if (c < 2 * R) val = wv[(c - R) * N + n];
                    HT[(int64_t)n * ldn + N + c] = val;
                }
                grid.sync();
            }
            n2o = n2; srows_o = srows; nc = hv;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// CTA-level FP64 GEMM  C[m][n] = sum_k A[k][m] * B[k][n]  (A stored K-major: "TN"), 256 threads,
// 64x64 tile, 4x4 register micro-tile, operands staged through shared memory.
struct GemmSmem {
    double As[16][64];
    double Bs[16][64];
};

// ------------------------------------------------------------------------------------------------
// Cholesky of H without its numerically zero rows (one CTA per geometry, packed triangle in shared memory, the same
// blocked DMMA factorisation as the frequency systems); see gs_doubles.  The zero rows (structurally: harmonic 0, whose
// radial derivative vanishes) go to the end of the permutation; the remaining block is well conditioned (smallest /
// largest eigenvalue >= 4e-3 over every geometry probed), so no pivoting is needed.  Any non-positive pivot or negative
// diagonal marks the geometry as not factorised and the frequency stage uses H itself.
constexpr int kHcholThreads = 512;          // maximum; small geometries launch 128
__global__ void __launch_bounds__(kHcholThreads) k_hchol(const void* blob, void* ws, unsigned smem_limit) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ double red_v[kHcholThreads / 32];
    __shared__ int sh_fail, sh_rank;
    DevView v = make_view(blob, ws);
    const int tid = threadIdx.x, nthr = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    for (int gi = blockIdx.x; gi < v.hdr->n_geom; gi += gridDim.x) {
        const GeomDesc gd = v.geom[gi];
        const int N = gd.N, R = gd.R, ldn = gs_ldn(N, R);
        double* gs = v.gstore + gd.gs_off;
        const double* HT = gs;
        double* st = gs + (int64_t)N * ldn + 2LL * R * N + (int64_t)R * R + 2 * R;
        double* LA = gs + gs_base_doubles(N, R);
        int* permg = reinterpret_cast<int*>(LA + (int64_t)N * ldn);
        __syncthreads();
        if (hchol_smem_bytes(N) > smem_limit || st[0] != 0.0) {
            if (tid == 0) st[1] = -1.0;
            continue;
        }
        double* sm = reinterpret_cast<double*>(smem);
        double* tri = sm + chol_smem_bytes(N, 0) / sizeof(double);
        int* perm = reinterpret_cast<int*>(tri + tri_doubles(N));
        // largest diagonal entry
        double dmax = -INFINITY;
        bool nan = false;
        double* diagv = sm;                                // the diagonal, staged for thread 0's serial scan (the factorisation's scratch is free until then)
        for (int i = tid; i < N; i += nthr) { const double x = HT[(int64_t)i * ldn + i]; diagv[i] = x; dmax = fmax(dmax, x); nan |= !(x == x); }
        if (nan) dmax = INFINITY;
        for (int o = 16; o; o >>= 1) dmax = fmax(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
        if (lane == 0) red_v[warp] = dmax;
        if (tid == 0) sh_fail = 0;
        __syncthreads();
        dmax = red_v[0];
        for (int w = 1; w < nwarp; ++w) dmax = fmax(dmax, red_v[w]);
        const double tol = 64.0 * (double)N * 2.220446049250313e-16 * dmax;
        if (tid == 0) {
            int nz = 0, bad = !(dmax > 0.0) || !(dmax < INFINITY);
            for (int i = 0; i < N && !bad; ++i) {
                const double x = diagv[i];
                if (x > tol) perm[nz++] = i;
                else if (x < -tol) bad = 1;
            }
            if (!bad) {
                int z = nz;
                for (int i = 0; i < N; ++i) if (!(diagv[i] > tol)) perm[z++] = i;
            }
            sh_rank = bad ? -1 : nz;
        }
        __syncthreads();
        int rank = sh_rank;
        if (rank > 0) {
            if (rank >= 64) {
                for (int r = warp; r < rank; r += nwarp) {     // one row per warp, lanes over the columns: no index division, and
                    const double* src = HT + (int64_t)perm[r] * ldn;   // perm is the identity up to the zero rows moved to the end
                    double* dst = tri + tri_off(r);
#pragma unroll 4
                    for (int c = lane; c <= r; c += 32) dst[c] = src[perm[c]];
                }
            } else {                                           // small blocks (design sweeps, N = 16): rows shorter than a warp
                for (int e = tid; e < rank * rank; e += nthr) {
                    const int r = e / rank, c = e - r * rank;
                    if (c <= r) tri[tri_off(r) + c] = HT[(int64_t)perm[r] * ldn + perm[c]];
                }
            }
            TriMat P;
            P.p = tri; P.aug = tri; P.K = rank; P.lda = 0;
            __syncthreads();
            cta_cholesky<true>(P, 0, sm, &sh_fail);
            if (sh_fail) rank = -1;
        }
        if (tid == 0) {
            st[1] = (double)rank;
            if (rank > 0) atomicAdd(reinterpret_cast<unsigned long long*>(v.counters + 16) + 1, 1ULL);
        }
        if (rank > 0) {
            for (int e = tid; e < N * ldn; e += nthr) {
                const int k = e / ldn, c = e - k * ldn, m = c - 2 * R;
                double val = 0.0;
                if (c < 2 * R) val = HT[(int64_t)perm[k] * ldn + N + c];
                else if (k < rank && m <= k) val = (m == k) ? 1.0 / tri[tri_off(k) + k] : tri[tri_off(k) + m];
                LA[e] = val;
            }
            for (int i = tid; i < N; i += nthr) permg[i] = perm[i];
            if (trmm_fits(N, gd.K, R)) {
                double* LAT = gs + gs_lat_off(N, R);
                auto la_val = [&](int k, int c) -> double {
                    if (c >= ldn || k >= N) return 0.0;
                    const int m = c - 2 * R;
                    if (c < 2 * R) return HT[(int64_t)perm[k] * ldn + N + c];
                    if (k < rank && m <= k) return (m == k) ? 1.0 / tri[tri_off(k) + k] : tri[tri_off(k) + m];
                    return 0.0;
                };
                const int ngl = (N + 31) / 32;
                for (int a = -1; a < ngl; ++a) {
                    const int k0 = a < 0 ? 0 : 32 * a, c0 = a < 0 ? 0 : 2 * R + 32 * a, wc = a < 0 ? 8 : 32;
                    double* tile = LAT + lat_tile_off(N, a);
                    for (int e = tid; e < (pad8(N) - k0) * 36; e += nthr) {
                        const int kr = e / 36, c = e - kr * 36;
                        tile[e] = c < wc ? la_val(k0 + kr, c0 + c) : 0.0;
                    }
                }
            }
        }
    }
}

template <class Store>
__device__ void cta_gemm_tn(const double* A, long lda, const double* B, long ldb, int Mm, int Nn, int Kd,
                            GemmSmem& sm, Store store) {
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    for (int m0 = 0; m0 < Mm; m0 += 64) {
        for (int n0 = 0; n0 < Nn; n0 += 64) {
            double acc[4][4];
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) acc[r][c] = 0.0;
            for (int k0 = 0; k0 < Kd; k0 += 16) {
                __syncthreads();
                for (int idx = tid; idx < 16 * 64; idx += 256) {
                    const int kk = idx >> 6, x = idx & 63;
                    const int k = k0 + kk;
                    sm.As[kk][x] = (k < Kd && m0 + x < Mm) ? A[(long)k * lda + m0 + x] : 0.0;
                    sm.Bs[kk][x] = (k < Kd && n0 + x < Nn) ? B[(long)k * ldb + n0 + x] : 0.0;
                }
                __syncthreads();
#pragma unroll
                for (int kk = 0; kk < 16; ++kk) {
                    double av[4], bv[4];
#pragma unroll
                    for (int r = 0; r < 4; ++r) av[r] = sm.As[kk][ty * 4 + r];
#pragma unroll
                    for (int c = 0; c < 4; ++c) bv[c] = sm.Bs[kk][tx * 4 + c];
#pragma unroll
                    for (int r = 0; r < 4; ++r)
#pragma unroll
                        for (int c = 0; c < 4; ++c) acc[r][c] = fma(av[r], bv[c], acc[r][c]);
                }
            }
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int m = m0 + ty * 4 + r, n = n0 + tx * 4 + c;
                    if (m < Mm && n < Nn) store(m, n, acc[r][c]);
                }
        }
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------------
// Schur path, frequency stage: one CTA (256 threads) per (geometry, frequency)
// slot layout (doubles): Imk[N*K] | T[N*K] | Wf[K*ldf] | ext[2K] | xs[2*R*K] | vv[2*R*N]
__host__ __device__ inline int freq_ld(int K, int R) { return (K + 1 + 2 * R + 1) & ~1; }
__host__ __device__ inline int64_t freq_slot_doubles(int N, int K, int R) {
    return 2LL * N * K + (int64_t)K * freq_ld(K, R) + 2LL * K + 2LL * R * K + 2LL * R * N + 8;
}

// Pivoted-LU variant.  only_flagged = 1: re-solve just the systems whose Cholesky (k_freq_chol) hit a
// non-positive pivot (status bit 2); only_flagged = 0: solve everything (plan option).
template <int NB>
__global__ void __launch_bounds__(256, 2) k_freq_lu(const void* blob, void* ws, OutPtrs out, int only_flagged) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ Layout L;
    __shared__ Exterior ex;
    __shared__ int sh_item, sh_status;
    __shared__ int64_t sh_scan;
    __shared__ GemmSmem gsm;
    DevView v = make_view(blob, ws);
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    const int F = v.hdr->n_freq;
    const int64_t total = (int64_t)v.hdr->n_geom * F;
    double* slot = v.slots + (int64_t)blockIdx.x * v.hdr->slot_doubles_freq;
    if (tid == 0) sh_scan = blockIdx.x;
    __syncthreads();
    for (;;) {
        int item;
        if (only_flagged) {      // static striding: every CTA scans its share of the status words, blockDim.x of them at a time
            for (;;) {
                const int64_t base = sh_scan;
                const int64_t it = base + (int64_t)tid * gridDim.x;
                bool flagged = false;
                if (it < total) {
                    const int gi_ = v.order[it / F];
                    flagged = (out.status[v.geom[gi_].out_sys + it % F] & 4) != 0;
                }
                if (tid == 0) sh_item = 0x7fffffff;
                __syncthreads();
                if (flagged) atomicMin(&sh_item, (int)it);
                __syncthreads();
                item = sh_item;
                const int64_t next = (item != 0x7fffffff) ? (int64_t)item + gridDim.x : base + (int64_t)nthr * gridDim.x;
                __syncthreads();
                if (tid == 0) sh_scan = next;
                __syncthreads();
                if (item != 0x7fffffff || next >= total) break;
            }
        } else {
            item = next_item(&v.counters[2], &sh_item);
        }
        if (item >= total) break;
        const int gi = v.order[item / F], f = item % F;
        const GeomDesc gd = v.geom[gi];
        const Geom g = geom_of(v, gd);
        const RootSet rs = v.sets[gd.root_set];
        const double m0 = v.roots[rs.off + f];
        const double* mk = v.roots + rs.off + F + (int64_t)f * rs.K;
        const double* Nk = v.roots + rs.off + F + (int64_t)F * rs.K + (int64_t)f * rs.K;
        if (tid == 0) { make_layout(g, L); make_exterior(g, m0, ex); sh_status = 0; }
        __syncthreads();
        const int N = gd.N, K = gd.K, R = gd.R, M = gd.M, p = gd.p;
        const int ldf = freq_ld(K, R);
        double* Imk = slot;
        double* T = Imk + (int64_t)N * K;
        double* Wf = T + (int64_t)N * K;
        double* ext = Wf + (int64_t)K * ldf;
        double* xs = ext + 2 * K;                 // x2_i[k] complex
        double* vv = xs + 2 * R * K;              // v_i[n] complex
        const double* gs = v.gstore + gd.gs_off;
        const double* HT = gs;
        const int ldn = gs_ldn(N, R);
        const double* wv = HT + (int64_t)N * ldn;
        const double* uv = wv + (int64_t)R * N;
        const double* s0 = uv + (int64_t)R * N;
        const double* phi = s0 + R * R;
        const double* bc = phi + R;
        const int gstatus = (int)bc[R];
        // 1. coupling table and exterior diagonal
        for (int e = tid; e < N * K; e += nthr) {
            const int n = e / K, k = e - n * K;
            Imk[e] = imk(n, k, g, ex, mk, Nk);
        }
        for (int k = tid; k < K; k += nthr) {
            const cplx dg = ext_diag(k, g, ex, mk);
            ext[2 * k] = dg.re; ext[2 * k + 1] = dg.im;
        }
        __syncthreads();
        // 2. T = H Imk
        cta_gemm_tn(HT, ldn, Imk, K, N, K, N, gsm, [&](int m, int n, double val) { T[(long)m * K + n] = val; });
        // 3. S = D - Imk^T T, column k -> k-1 (k >= 1), the complex column k = 0 -> (K-1 | K)
        cta_gemm_tn(Imk, K, T, K, K, K, N, gsm, [&](int m, int n, double val) {
            double sv = -val;
            if (m == n) sv += ext[2 * n];
            const int col = (n == 0) ? K - 1 : n - 1;
            Wf[(long)m * ldf + col] = sv;
        });
        for (int k = tid; k < K; k += nthr) Wf[(long)k * ldf + K] = (k == 0) ? ext[1] : 0.0;
        // right-hand sides  y_i[k] = bc_i Imk[0][k] + sum_n Imk[n][k] u_i[n]
        for (int e = tid; e < R * K; e += nthr) {
            const int i = e / K, k = e - i * K;
            double acc = bc[i] * Imk[k];
            for (int n = 0; n < N; ++n) acc = fma(Imk[(long)n * K + k], uv[i * N + n], acc);
            Wf[(long)k * ldf + K + 1 + i] = acc;
        }
        __syncthreads();
        // 4. eliminate the K-1 real columns, then the complex pivot
        lu_forward<double, NB>(Wf, ldf, K, K - 1, K + 1 + R, smem, &sh_status);
        __syncthreads();
        {
            const double are = Wf[(long)(K - 1) * ldf + K - 1], aim = Wf[(long)(K - 1) * ldf + K];
            if (tid == 0) {
                if (are == 0.0 && aim == 0.0) atomicOr(&sh_status, 1);
                if (!isfinite(are) || !isfinite(aim)) atomicOr(&sh_status, 2);
            }
            for (int i = 0; i < R; ++i) {
                const cplx x0 = cmake(Wf[(long)(K - 1) * ldf + K + 1 + i], 0.0) / cmake(are, aim);
                if (tid == 0) { xs[2 * (i * K)] = x0.re; xs[2 * (i * K) + 1] = x0.im; }
                for (int r = tid; r < K - 1; r += nthr) {
                    const double cre = Wf[(long)r * ldf + K - 1], cim = Wf[(long)r * ldf + K];
                    const double y = Wf[(long)r * ldf + K + 1 + i];
                    Wf[(long)r * ldf + K + 1 + i] = y - (x0.re * cre - x0.im * cim);
                    Wf[(long)r * ldf + K + 1 + R + i] = -(x0.re * cim + x0.im * cre);
                }
            }
        }
        __syncthreads();
        lu_backward<double, NB>(Wf, ldf, K - 1, K + 1, K + 1 + 2 * R, smem, K);
        __syncthreads();
        for (int e = tid; e < R * (K - 1); e += nthr) {
            const int i = e / (K - 1), r = e - i * (K - 1);
            xs[2 * (i * K + r + 1)] = Wf[(long)r * ldf + K + 1 + i];
            xs[2 * (i * K + r + 1) + 1] = Wf[(long)r * ldf + K + 1 + R + i];
        }
        __syncthreads();
        // 5. v_i = Imk x2_i (complex), one warp per (i, n)
        for (int t = warp; t < R * N; t += nwarp) {
            const int i = t / N, n = t - i * N;
            double ar = 0.0, ai = 0.0;
            for (int k = lane; k < K; k += 32) {
                const double c = Imk[(long)n * K + k];
                ar = fma(c, xs[2 * (i * K + k)], ar);
                ai = fma(c, xs[2 * (i * K + k) + 1], ai);
            }
            ar = warp_sum(ar); ai = warp_sum(ai);
            if (lane == 0) { vv[2 * t] = ar; vv[2 * t + 1] = ai; }
        }
        __syncthreads();
        // 6. coefficients (eng.py:370-391)
        const int status = sh_status | gstatus;
        const int64_t sys = gd.out_sys + f;
        if (tid == 0) { out.m0[sys] = m0; out.status[sys] = status; }
        const double hk_angle = ex.inf_m0 ? 0.0 : hankel0_angle(m0 * g.a[M - 1]);
        for (int t = warp; t < R * R; t += nwarp) {
            const int j = t / R, i = t - j * R;
            double ar = 0.0, ai = 0.0;
            for (int n = lane; n < N; n += 32) {
                ar = fma(wv[j * N + n], vv[2 * (i * N + n)], ar);
                ai = fma(wv[j * N + n], vv[2 * (i * N + n) + 1], ai);
            }
            ar = warp_sum(ar); ai = warp_sum(ai);
            if (lane == 0) {
                const double cre = 2.0 * kPi * ((s0[j * R + i] + ar) + phi[j]);
                const double cim = 2.0 * kPi * ai;
                double am = cre * v.hdr->rho;
                double dpv = ex.inf_m0 ? 0.0 : cim * omega_of_m0(m0, g.h) * v.hdr->rho;
                if (status) { am = NAN; dpv = NAN; }
                const int64_t o = gd.out_rr + ((int64_t)f * R + j) * R + i;
                out.am[o] = am;
                out.dp[o] = dpv;
                if (i == j) {
                    const int64_t o1 = gd.out_r + (int64_t)f * R + j;
                    double ph = ex.inf_m0 ? -(kPi / 2.0)
                                          : -(kPi / 2.0) + atan2(xs[2 * (i * K) + 1], xs[2 * (i * K)]) - hk_angle;
                    double fo = excitation_force(dpv, m0, g.h, kRhoDefault);
                    if (status) { ph = NAN; fo = NAN; }
                    out.force[o1] = fo;
                    out.phase[o1] = ph;
                }
            }
        }
        // 7. optional X = [y0 + G v ; x2]
        if (out.x != nullptr && gd.xg_off >= 0) {
            const double* xg = v.xstore + gd.xg_off;
            const int NT = gd.NT;
            for (int t = warp; t < R * p; t += nwarp) {
                const int i = t / p, r = t - i * p;
                const double* grow = xg + (int64_t)r * (N + R);
                double ar = 0.0, ai = 0.0;
                for (int n = lane; n < N; n += 32) {
                    ar = fma(grow[n], vv[2 * (i * N + n)], ar);
                    ai = fma(grow[n], vv[2 * (i * N + n) + 1], ai);
                }
                ar = warp_sum(ar); ai = warp_sum(ai);
                if (lane == 0) {
                    double* xo = out.x + 2 * (gd.out_x + ((int64_t)f * R + i) * NT + r);
                    xo[0] = status ? NAN : grow[N + i] + ar;
                    xo[1] = status ? NAN : ai;
                }
            }
            for (int e = tid; e < R * K; e += nthr) {
                const int i = e / K, k = e - i * K;
                double* xo = out.x + 2 * (gd.out_x + ((int64_t)f * R + i) * NT + p + k);
                xo[0] = status ? NAN : xs[2 * e];
                xo[1] = status ? NAN : xs[2 * e + 1];
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Schur path, frequency stage (default), two kernels per wave of systems:
//   k_freq_gemm   256 threads, 2 CTAs/SM, throughput-bound: coupling table Imk, T = H Imk and the lower triangle of
//                 P = Imk^T T - Re(D) on the DMMA pipe; P, the right-hand sides and q_j = Imk^T w_j go to the
//                 system's buffer in HBM/L2
//   k_freq_solve  128 threads, many CTAs/SM, latency-bound: Cholesky of P, back-substitution, Sherman-Morrison
//                 closure of the single complex entry, hydrodynamic coefficients (+ X)
// Splitting lets each half run at the occupancy it needs (the solve has ~K sequential steps per system and hides
// its latency only through many systems in flight per SM).
// The exterior index k is stored permuted, kp = k - 1 (k >= 1), k = 0 -> kp = K - 1, so that the single complex
// diagonal entry sits last and the leading K x K real part is what gets factorised.
// gemm slot (doubles): Imk[N][ldk] | T[N][ldk]
// system buffer: Ptri[tri_doubles(K)] (packed lower triangle) | aug[2R][ldk] (rows r_i, then q_j) | ext[ldk] | xs[R+1][ldk] | misc[8]
// The q_j ride through the factorisation as right-hand sides like the r_i: with P = L L^T, q_j . P^{-1} r_i =
// (L^{-1} q_j) . (L^{-1} r_i), and everything the coefficients need from z = P^{-1} e_last is L_KK and (L^{-1} q_j)_K -
// so the coefficient path has no back-substitution at all (only the optional X output does).
// (rows padded to a multiple of 16 and kept zero beyond the live ones: the bulk-copied contraction chunks of the rank update
// read whole 16-row chunks)
__host__ __device__ inline int pad16(int n) { return (n + 15) & ~15; }
__host__ __device__ inline int64_t freq2_slot_doubles(int N, int K, int R) { return 2LL * pad16(N) * syrk_lds(K) + 16; }
__host__ __device__ inline int64_t freq2_sys_doubles(int K, int R) { return (int64_t)tri_doubles(K) + (int64_t)(3 * R + 2) * pad8(K) + 8; }
// k_freq_solve dynamic shared memory: Cholesky scratch | xc, vv (X path) | [resident variant: Ptri | aug[R][ldk] | xs[R+1][ldk]]
__host__ __device__ inline size_t freq2_solve_smem_bytes(int K, int R, int N, bool want_x, bool resident) {
    size_t b = chol_smem_bytes(K, 2 * R);
    if (want_x) b += sizeof(double) * (size_t)(2 * R * K + 2 * R * N);
    if (resident) b += sizeof(double) * ((size_t)tri_doubles(K) + (size_t)(3 * R + 1) * pad8(K));
    return b;
}

enum { PH_IMK = 0, PH_DOTS, PH_GEMM1, PH_GEMM2, PH_CHOL, PH_BACK, PH_OUT, PH_X, PH_COUNT };
#define PHASE_END(ph) do { __syncthreads(); const long long tn_ = clock64(); tph[ph] += tn_ - tprev; tprev = tn_; } while (0)
#ifdef MEEM_CHOL_PROF      // the solve kernel's four profile slots then carry the Cholesky's sub-phases instead
#define PHASE_END_S(ph) __syncthreads()
#else
#define PHASE_END_S(ph) PHASE_END(ph)
#endif

// Coupling table I_mk of one (geometry, frequency) system as the frequency stage consumes it: Imk[row][kp] with
// column kp = exterior mode kp + 1 (the propagating mode k = 0 is carried LAST, kp = K - 1) and row kk = harmonic
// perm[kk] of the last interior region (perm == nullptr: identity).  All threads of the CTA must call; sm needs
// 6K + 3N doubles for the angle-addition form (else every entry goes through the reference-ordered imk()).
// Interior entries (m >= 1, k >= 1) use the angle-addition form of me.py:119-124:
//   sin((m_k +- lambda) L) = sin(m_k L) cos(lambda L) +- cos(m_k L) sin(lambda L)
// with the 2K + 2N sines/cosines tabulated once per system, so an entry costs two divisions instead of two
// sines, four divisions and a square root.  Both forms carry ~1e-13 absolute argument-rounding noise in the
// sines (|theta| eps, theta up to ~1e3); the exact-equality branch |m_k| == lambda is kept.  Row m = 0 and
// column k = 0 go through the reference-ordered imk().  (meemb200_assemble keeps imk() for every entry.)
__device__ void cta_imk_table(const Geom& g, const Exterior& ex, const double* mk, const double* Nk, int N, int K, int ldk,
                              const int* perm, double* Imk, double* sm, int sm_doubles, int* sh_row0, bool batched = false) {
    const int tid = threadIdx.x, nthr = blockDim.x;
#define HARM(kk) (perm ? perm[kk] : (kk))
    const bool fast = !ex.inf_m0 && (6 * K + 3 * N) <= sm_doubles;
    if (fast) {
        double* tmk = sm;               // [K] m_k (permuted)
        double* tsk = tmk + K;          // sin(m_k L)
        double* tck = tsk + K;          // cos(m_k L)
        double* tnk = tck + K;          // sqrt(2 / N_k) / 2
        double* tpk = tnk + K;          // 2 tnk m_k sin(m_k L)
        double* tqk = tpk + K;          // 2 tnk cos(m_k L)
        double* tlm = tqk + K;          // [N] lambda_m
        double* tsl = tlm + N;          // sin(lambda_m L)
        double* tcl = tsl + N;          // cos(lambda_m L)
        const double Lh = g.h - g.d[g.M - 1];
        for (int kp = tid; kp < K; kp += nthr) {
            const int k = (kp == K - 1) ? 0 : kp + 1;
            const double mv = mk[k];
            double sv, cv;
            const double pr = mv * Lh, er = fma(mv, Lh, -pr);     // exact product = pr + er
            sincos(pr, &sv, &cv);
            const double sk = fma(er, cv, sv), ck = fma(-er, sv, cv), nk2 = sqrt(2.0 / Nk[k]) * 0.5;
            tmk[kp] = mv; tsk[kp] = sk; tck[kp] = ck; tnk[kp] = nk2;
            tpk[kp] = 2.0 * nk2 * mv * sk; tqk[kp] = 2.0 * nk2 * ck;
        }
        for (int n = tid; n < N; n += nthr) {
            const double lam = lambda_ni(HARM(n), g.M - 1, g);
            double sv, cv;
            const double pr = lam * Lh, er = fma(lam, Lh, -pr);
            sincos(pr, &sv, &cv);
            tlm[n] = lam; tsl[n] = fma(er, cv, sv); tcl[n] = fma(-er, sv, cv);
        }
        __syncthreads();
        // row m = 0 and column k = 0 (reference-ordered forms), kept out of the main loop so that it stays uniform
        for (int e = tid; e < N * 2; e += nthr) {                 // column k = 0 of every row; find the row of harmonic 0
            const int kk = e >> 1;
            if (e & 1) { if (HARM(kk) == 0) (*sh_row0) = kk; }
            else Imk[(long)kk * ldk + K - 1] = imk(HARM(kk), 0, g, ex, mk, Nk);
        }
        __syncthreads();
        for (int kp = tid; kp < K - 1; kp += nthr) Imk[(long)(*sh_row0) * ldk + kp] = imk(0, kp + 1, g, ex, mk, Nk);
        // interior: one row per warp, lanes over the columns.  With a = sin(m_k L) cos(lam L), b = cos(m_k L) sin(lam L),
        // sp = m_k + lam, sd = m_k - lam the entry is  nk [(a + b) / sp + sin(sd L) / sd].  Away from a resonance
        // (|sd L| >= 1) sin(sd L) = a - b and the sum over the common denominator collapses to the rank-2 form
        //   nk (2 m_k a - 2 lam b) / (sp sd) = (P_k cos(lam L) - Q_k lam sin(lam L)) / (sp sd)
        // with P_k, Q_k tabulated per column: 8 FP64 instructions and one division per entry.  Near a resonance a - b
        // would lose digits to cancellation: sin(x) = x sinc(x) by its Taylor series (|x| < 1) in the two-term form.
        {
            const int lane_ = tid & 31, warp_ = tid >> 5, nwarp_ = nthr >> 5;
            for (int n = warp_; n < N; n += nwarp_) {
                if (n == (*sh_row0)) continue;
                const double lam = tlm[n], sl = tsl[n], cl = tcl[n];
                const double lsl = lam * sl;
                double* row = Imk + (long)n * ldk;
                if (batched) {
                    // four entries per lane and pass share ONE division: r = 1 / (d0 d1 d2 d3), 1 / d_i = r * (product of the others)
                    // (prefix products; an entry near a resonance or beyond the row enters the product as 1 and takes the series form)
                    for (int kp0 = lane_; kp0 < K - 1; kp0 += 128) {
                        double den[4], num[4];
                        unsigned nearm = 0;
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int kp = kp0 + 32 * i;
                            const bool valid = kp < K - 1;
                            const int kc = valid ? kp : kp0;
                            const double mv = tmk[kc];
                            const double sp = mv + lam, sd = mv - lam;
                            num[i] = fma(tpk[kc], cl, -(tqk[kc] * lsl));
                            const bool nr = fabs(sd * Lh) < 1.0;
                            den[i] = (nr || !valid) ? 1.0 : sp * sd;
                            if (nr && valid) nearm |= 1u << i;
                        }
                        const double p1 = den[0] * den[1], p2 = p1 * den[2], p3 = p2 * den[3];
                        double r = 1.0 / p3;
                        const double i3 = r * p2; r *= den[3];
                        const double i2 = r * p1; r *= den[2];
                        const double i1 = r * den[0], i0 = r * den[1];
                        double val[4] = {num[0] * i0, num[1] * i1, num[2] * i2, num[3] * i3};
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int kp = kp0 + 32 * i;
                            if (nearm & (1u << i)) {
                                const double mv = tmk[kp];
                                const double sp = mv + lam, sd = mv - lam;
                                const double x = sd * Lh, x2 = x * x;
                                double pl = -8.22063524662432972e-18;
                                pl = fma(pl, x2, 2.81145725434552076e-15);
                                pl = fma(pl, x2, -7.64716373181981648e-13);
                                pl = fma(pl, x2, 1.60590438368216146e-10);
                                pl = fma(pl, x2, -2.50521083854417188e-08);
                                pl = fma(pl, x2, 2.75573192239858907e-06);
                                pl = fma(pl, x2, -1.98412698412698413e-04);
                                pl = fma(pl, x2, 8.33333333333333333e-03);
                                pl = fma(pl, x2, -1.66666666666666667e-01);
                                const double S = fma(x * x2, pl, x);
                                const double a = tsk[kp] * cl, b = tck[kp] * sl;
                                val[i] = tnk[kp] * (fma(a + b, sd, S * sp) / (sp * sd));
                                if (fabs(mv) == lam) val[i] = sqrt(2.0 / Nk[kp + 1]) * Lh / 2.0;
                            }
                            if (kp < K - 1) row[kp] = val[i];
                        }
                    }
                    continue;
                }
                for (int kp = lane_; kp < K - 1; kp += 32) {
                    const double mv = tmk[kp];
                    const double sp = mv + lam, sd = mv - lam;
                    const double x = sd * Lh;
                    double val = fma(tpk[kp], cl, -(tqk[kp] * lsl)) / (sp * sd);
                    if (fabs(x) < 1.0) {
                        const double x2 = x * x;
                        double pl = -8.22063524662432972e-18;                       // -1/19!
                        pl = fma(pl, x2, 2.81145725434552076e-15);                  //  1/17!
                        pl = fma(pl, x2, -7.64716373181981648e-13);                 // -1/15!
                        pl = fma(pl, x2, 1.60590438368216146e-10);                  //  1/13!
                        pl = fma(pl, x2, -2.50521083854417188e-08);                 // -1/11!
                        pl = fma(pl, x2, 2.75573192239858907e-06);                  //  1/9!
                        pl = fma(pl, x2, -1.98412698412698413e-04);                 // -1/7!
                        pl = fma(pl, x2, 8.33333333333333333e-03);                  //  1/5!
                        pl = fma(pl, x2, -1.66666666666666667e-01);                 // -1/3!
                        const double S = fma(x * x2, pl, x);
                        const double a = tsk[kp] * cl, b = tck[kp] * sl;
                        val = tnk[kp] * (fma(a + b, sd, S * sp) / (sp * sd));
                        if (fabs(mv) == lam) val = sqrt(2.0 / Nk[kp + 1]) * Lh / 2.0;
                    }
                    row[kp] = val;
                }
            }
        }
    } else {
        for (int e = tid; e < N * K; e += nthr) {
            const int n = e / K, kp = e - n * K;
            const int k = (kp == K - 1) ? 0 : kp + 1;
            Imk[(long)n * ldk + kp] = imk(HARM(n), k, g, ex, mk, Nk);
        }
    }
#undef HARM
}

// Parity export of the product coupling table (meemb200_imk_table): one CTA, identity row order, columns in the
// reference's order k = 0 .. K-1.
__global__ void __launch_bounds__(256) k_imk_dump(const void* blob, void* ws, int gi, int f, double* out) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ Exterior ex;
    __shared__ int sh_row0;
    DevView v = make_view(blob, ws);
    const GeomDesc gd = v.geom[gi];
    const Geom g = geom_of(v, gd);
    const RootSet rs = v.sets[gd.root_set];
    const int F = v.hdr->n_freq;
    const double m0 = v.roots[rs.off + f];
    const double* mk = v.roots + rs.off + F + (int64_t)f * rs.K;
    const double* Nk = v.roots + rs.off + F + (int64_t)F * rs.K + (int64_t)f * rs.K;
    if (threadIdx.x == 0) make_exterior(g, m0, ex);
    __syncthreads();
    const int N = gd.N, K = gd.K, ldk = pad8(K);
    double* sm = reinterpret_cast<double*>(smem);
    double* Imk = v.slots;                    // scratch: the first slot
    cta_imk_table(g, ex, mk, Nk, N, K, ldk, nullptr, Imk, sm, 6 * K + 3 * N, &sh_row0, true);      // the product's form
    __syncthreads();
    for (int e = threadIdx.x; e < N * K; e += blockDim.x) {
        const int n = e / K, k = e - n * K;
        out[e] = Imk[(long)n * ldk + (k == 0 ? K - 1 : k - 1)];
    }
}

__global__ void __launch_bounds__(256, 2) k_freq_gemm(const void* blob, void* ws, int item0, int item1, int* counter, int flags) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ Layout L;
    __shared__ Exterior ex;
    __shared__ int sh_item, sh_row0;
    __shared__ SyrkPlan sh_syrk;
    __shared__ __align__(8) uint64_t sh_mbar[kGemmStages];
    __shared__ __align__(8) TrRing sh_ring;
    unsigned mbar_phase = 0;                               // bit s: parity the next wait on stage s expects
    unsigned ring_bits = 0;                                // the same for the chunk ring of the triangular product
    if (threadIdx.x == 0) {
        sh_syrk.nb = -1;
        for (int i = 0; i < kGemmStages; ++i) mbar_init(&sh_mbar[i], 1);
        for (int i = 0; i < kTrMaxStages; ++i) { mbar_init(&sh_ring.full[i], 1); mbar_init(&sh_ring.empty[i], (int)(blockDim.x >> 5)); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    const bool use_bulk = (flags & 4) == 0;
    const bool use_syrk = (flags & 1) == 0;
    const bool use_stream = (flags & 8) == 0;
    const bool use_stream_syrk = (flags & 16) == 0;
    DevView v = make_view(blob, ws);
    double* sm = reinterpret_cast<double*>(smem);
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int F = v.hdr->n_freq;
    double* slot = v.slots + (int64_t)blockIdx.x * v.hdr->slot_doubles_freq;
    long long tph[4] = {0, 0, 0, 0};
    long long tprev = clock64();
    for (;;) {
        const int item = item0 + next_item(counter, &sh_item);
        if (item >= item1) break;
        const int gi = v.order[item / F], f = item % F;
        const GeomDesc gd = v.geom[gi];
        const Geom g = geom_of(v, gd);
        const RootSet rs = v.sets[gd.root_set];
        const double m0 = v.roots[rs.off + f];
        const double* mk = v.roots + rs.off + F + (int64_t)f * rs.K;
        const double* Nk = v.roots + rs.off + F + (int64_t)F * rs.K + (int64_t)f * rs.K;
        if (tid == 0) { make_layout(g, L); make_exterior(g, m0, ex); }
        __syncthreads();
        const int N = gd.N, K = gd.K, R = gd.R;
        const int ldk = pad8(K), ldn = gs_ldn(N, R);
        const int lsl = syrk_lds(K);                        // leading dimension of Imk / T in the slot = that of the shared-memory stages
        double* Imk = slot;
        double* T = Imk + (int64_t)pad16(N) * lsl;             // (re-pointed at Imk below on the chunk-stream path)
        double* sys = v.sysbuf + (int64_t)(item - item0) * v.hdr->sys_stride;
        double* P = sys;                                   // packed lower triangle
        double* aug = P + tri_doubles(K);                  // [2R][ldk] right-hand sides r_i, then q_j
        double* ext = aug + (int64_t)2 * R * ldk;          // 1. exterior diagonal (real parts)
        double* misc = ext + (int64_t)(R + 2) * ldk;
        const double* gs = v.gstore + gd.gs_off;
        const double* HT = gs;
        const double* wv = HT + (int64_t)N * ldn;
        // factorised H (k_hchol): row kk of LA / of the coupling table belongs to harmonic perm[kk]
        const int hrank = (int)gs[(int64_t)N * ldn + 2LL * R * N + (int64_t)R * R + 2 * R + 1];
        const bool lmode = hrank > 0;
        const double* LA = gs + gs_base_doubles(N, R);
        const int* perm = reinterpret_cast<const int*>(LA + (int64_t)N * ldn);
        const double* LAT = gs + gs_lat_off(N, R);
#define HARM(kk) (lmode ? perm[kk] : (kk))
        // 1. coupling table (columns permuted; rows permuted in the factorised-H form): cta_imk_table
        cta_imk_table(g, ex, mk, Nk, N, K, lsl, lmode ? perm : nullptr, Imk, sm, kGemmSmemDoubles, &sh_row0, (flags & 32) == 0);
        for (int e = tid; e < (pad8(N) - N) * lsl; e += nthr) Imk[(int64_t)N * lsl + e] = 0.0;      // the chunk stream reads whole 8-row chunks
        for (int kp = tid; kp < K; kp += nthr) {
            const int k = (kp == K - 1) ? 0 : kp + 1;
            const cplx dg = ext_diag(k, g, ex, mk);
            ext[kp] = dg.re;
            if (k == 0) misc[0] = dg.im;
        }
        PHASE_END(PH_IMK);
        // 2./3. [T ; -r_i ; q_j] = [H | u'_i | w_j]^T-contraction with Imk: rows N.. of the product are the right-hand
        // sides r_i = -(bc_i Imk[0][:] + Imk^T u_i) (augmented rows of P) and q_j = Imk^T w_j
        // Factorised form: [-r_i ; q_j ; W] = LA^T-contraction with the row-permuted Imk, W = L^T Imk; output rows
        // 2R + m only need contraction rows k >= m (L is lower trapezoidal), which the GEMM skips chunk-wise.
        if (lmode && use_stream && trmm_fits(N, K, R)) {
            // chunk-stream form (meem_dmma.cuh): row groups of W, balanced warp tiles, mbarrier ring without CTA barriers.
            // W overwrites the table IN PLACE: row group a of W replaces rows 32 a .. 32 a + 31 of Imk, which only the groups
            // <= a contract over and whose bulk reads have all landed before any warp finishes group a - the slot's second
            // half stays untouched, which halves the L2 footprint of the slots (296 x 353 KB instead of 705 KB at N = K = 200)
            T = Imk;
            auto epi_w = [&](int m, int n, double v0, double v1) {
                if (m < hrank && n < K) *reinterpret_cast<double2*>(T + (long)m * lsl + n) = make_double2(v0, v1);
            };
            auto epi_a = [&](int m, int n, double v0, double v1) {
                if (m < 2 * R && n < K) *reinterpret_cast<double2*>(aug + (long)m * ldk + n) = m < R ? make_double2(-v0, -v1) : make_double2(v0, v1);
            };
            // 16-row chunks when three stages of them fit (fewer ring hand-overs per DMMA: 3.48 -> 3.38 ms at N = K = 200), else 8
            const bool kc16 = trmm_stages<16>(K) >= 3;
#define MEEM_TRMM(NJ_, KC_) cta_trmm_stream<NJ_, KC_>(LAT, Imk, hrank, N, K, sm, sh_ring, ring_bits, -1, epi_w, epi_a)
            switch (((K + 7) >> 3) >> 3) {
                case 1: MEEM_TRMM(1, 16); break;
                case 2: MEEM_TRMM(2, 16); break;
                case 3: if (kc16) MEEM_TRMM(3, 16); else MEEM_TRMM(3, 8); break;
                default: MEEM_TRMM(4, 8); break;
            }
#undef MEEM_TRMM
        } else if (lmode) {
            cta_gemm_dmma(LA, ldn, Imk, lsl, 2 * R + hrank, K, N, false, 2 * R, sm, [&](int m, int n, double v0, double v1) {
                if (n >= K) return;
                if (m < R) *reinterpret_cast<double2*>(aug + (long)m * ldk + n) = make_double2(-v0, -v1);
                else if (m < 2 * R) *reinterpret_cast<double2*>(aug + (long)m * ldk + n) = make_double2(v0, v1);
                else if (m < 2 * R + hrank) *reinterpret_cast<double2*>(T + (long)(m - 2 * R) * lsl + n) = make_double2(v0, v1);
            });
        } else {
            cta_gemm_dmma(HT, ldn, Imk, lsl, N + 2 * R, K, N, false, -1, sm, [&](int m, int n, double v0, double v1) {
                if (n >= K) return;
                if (m < N) *reinterpret_cast<double2*>(T + (long)m * lsl + n) = make_double2(v0, v1);
                else if (m < N + R) *reinterpret_cast<double2*>(aug + (long)(m - N) * ldk + n) = make_double2(-v0, -v1);
                else if (m < N + 2 * R) *reinterpret_cast<double2*>(aug + (long)(m - N) * ldk + n) = make_double2(v0, v1);
            });
        }
        PHASE_END(PH_GEMM1);
        // 4. P = Imk^T T - Re(D)  (= W^T W - Re(D) in the factorised form), lower triangle
        auto store_p = [&](int m, int n, double v0, double v1) {
            if (m < K && n <= m) {
                if (m == n) v0 -= ext[m];
                if (m == n + 1) v1 -= ext[m];
                *reinterpret_cast<double2*>(P + tri_off(m) + n) = make_double2(v0, v1);
            }
        };
        if (lmode && use_syrk && use_stream_syrk && K >= 64 && syrk_stream_fits<16>(K)) {
            for (int e = tid; e < (pad16(hrank) - hrank) * lsl; e += nthr) T[(int64_t)hrank * lsl + e] = 0.0;
            cta_syrk_stream<16>(T, K, hrank, sm, sh_syrk, sh_ring, ring_bits, -1, store_p);
        } else if (lmode && use_syrk && K >= 64 && syrk_fits(K)) {
            const bool bulk = use_bulk && (K & 7) == 0;        // whole 64-byte rows: the bulk copies need 16-byte sizes
            if (bulk) {
                // rows [hrank, pad16(hrank)) of W are read by the last 16-row chunk: zero; then make this CTA's generic-proxy
                // stores to W visible to the async proxy (TMA reads of global memory)
                for (int e = tid; e < (pad16(hrank) - hrank) * lsl; e += nthr) T[(int64_t)hrank * lsl + e] = 0.0;
                asm volatile("fence.proxy.async;" ::: "memory");
                __syncthreads();
            }
            cta_syrk_dmma(T, lsl, K, hrank, sm, sh_syrk, bulk ? sh_mbar : nullptr, mbar_phase, store_p);
        }
        else cta_gemm_dmma(lmode ? T : Imk, lsl, T, lsl, K, K, lmode ? hrank : N, true, -1, sm, store_p);
        PHASE_END(PH_GEMM2);
    }
    if (tid == 0) {
        unsigned long long* prof = reinterpret_cast<unsigned long long*>(v.counters + 16);
        for (int i = 0; i < 4; ++i) atomicAdd(&prof[i], (unsigned long long)tph[i]);
    }
}

// RESIDENT: the packed P (+ right-hand-side rows) is pulled into shared memory with one bulk copy and the whole
// factorisation / back-substitution runs there (K <= ~208 at one CTA per SM); otherwise P stays in its L2-resident buffer.
#undef HARM
template <int THREADS, int MINB, bool RESIDENT>
__global__ void __launch_bounds__(THREADS, MINB) k_freq_solve(const void* blob, void* ws, OutPtrs out, int item0, int item1, int* counter, int flags) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ Layout L;
    __shared__ Exterior ex;
    __shared__ int sh_item, sh_fail;
    __shared__ double sh_dots[64];
    __shared__ __align__(8) uint64_t sh_bar;
    DevView v = make_view(blob, ws);
    double* sm = reinterpret_cast<double*>(smem);
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    const int F = v.hdr->n_freq;
    long long tph[4] = {0, 0, 0, 0};
    long long tprev = clock64();
    unsigned bar_phase = 0;
    if (RESIDENT) {
        if (tid == 0) mbar_init(&sh_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        __syncthreads();
    }
    for (;;) {
        const int item = item0 + next_item(counter, &sh_item);
        if (item >= item1) break;
        const int gi = v.order[item / F], f = item % F;
        const GeomDesc gd = v.geom[gi];
        const Geom g = geom_of(v, gd);
        const RootSet rs = v.sets[gd.root_set];
        const double m0 = v.roots[rs.off + f];
        const int N = gd.N, K = gd.K, R = gd.R, M = gd.M, p = gd.p;
        const int ldk = pad8(K), ldn = gs_ldn(N, R);
        double* sys = v.sysbuf + (int64_t)(item - item0) * v.hdr->sys_stride;
        double* xsg = sys + tri_doubles(K) + (int64_t)(2 * R + 1) * ldk;
        const double* misc = xsg + (int64_t)(R + 1) * ldk;
        double* xpath = sm + chol_smem_bytes(K, 2 * R) / sizeof(double);      // xc, vv (X path only)
        double* res = xpath + (out.x != nullptr ? 2 * R * K + 2 * R * N : 0);  // resident copy of Ptri | aug
        TriMat P;
        P.K = K; P.lda = ldk;
        P.p = RESIDENT ? res : sys;
        P.aug = P.p + tri_doubles(K);
        double* xs = RESIDENT ? P.aug + (int64_t)2 * R * ldk : xsg;
        if (RESIDENT && tid == 0) {
            // the previous system's generic-proxy accesses to this shared memory were ordered by next_item's barriers
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            const unsigned bytes = 8u * (unsigned)(tri_doubles(K) + 2 * R * ldk);
            mbar_expect_tx(&sh_bar, bytes);
            for (unsigned o = 0; o < bytes; o += 32768u)
                bulk_g2s(reinterpret_cast<char*>(res) + o, reinterpret_cast<const char*>(sys) + o, min(32768u, bytes - o), &sh_bar);
        }
        if (tid == 0) { make_layout(g, L); make_exterior(g, m0, ex); sh_fail = 0; }
        const double* gs = v.gstore + gd.gs_off;
        const double* s0 = gs + (int64_t)N * ldn + 2LL * R * N;
        const double* phi = s0 + R * R;
        const double* bc = phi + R;
        const int gstatus = (int)bc[R];
        if (RESIDENT) { mbar_wait(&sh_bar, bar_phase); bar_phase ^= 1u; }
        __syncthreads();
        // 5. P = L L^T; rows K.. become L^{-1} r_i and L^{-1} q_j
#ifdef MEEM_CHOL_PROF
        cta_cholesky(P, 2 * R, sm, &sh_fail, tph);
        tprev = clock64();
#else
        // P in L2: left-looking (one pass over the factor per block column, accumulators in registers); P in shared
        // memory: right-looking
        if (!RESIDENT && !(flags & 2)) cta_cholesky_ll(P, 2 * R, sm, &sh_fail);
        else cta_cholesky(P, 2 * R, sm, &sh_fail);
        PHASE_END(0);
#endif
        // 6. (X output only) back-substitution for [w_1 .. w_R, z],  z = P^{-1} e_last.  The diagonal blocks of P hold the
        // inverses of the factor's diagonal blocks, so P[K-1][K-1] = 1 / L_KK is the forward-solved e_last.
        const double* fr = P.aug;                          // (L^{-1} r_i)^T
        const double* fq = P.aug + (int64_t)R * ldk;       // (L^{-1} q_j)^T
        const double linv_last = P.row(K - 1)[K - 1];
        if (out.x != nullptr) {
            for (int t = tid; t < (R + 1) * K; t += nthr) {
                const int r = t / K, kp = t - r * K;
                xs[(long)r * ldk + kp] = (r < R) ? fr[(long)r * ldk + kp] : (kp == K - 1 ? linv_last : 0.0);
            }
            cta_chol_backsolve(P, xs, ldk, R + 1);
        }
        PHASE_END_S(1);
        // 7. coefficients (eng.py:370-391).  x2_i = w_i + i beta x_e z,  x_e = w_e / (1 - i beta z_e), with
        // q_j . w_i = fq_j . fr_i,  q_j . z = fq_j[K-1] / L_KK,  w_e = fr_i[K-1] / L_KK,  z_e = 1 / L_KK^2
        const double beta = misc[0];
        const double* zv = xs + (long)R * ldk;
        for (int t = warp; t < R * R; t += nwarp) {
            const double* qa = fq + (long)(t / R) * ldk;
            const double* xb = fr + (long)(t % R) * ldk;
            double a = 0.0;
            for (int k = lane; k < K; k += 32) a = fma(qa[k], xb[k], a);
            a = warp_sum(a);
            if (lane == 0 && t < 64) sh_dots[t] = a;
        }
        __syncthreads();
        const int64_t sysi = gd.out_sys + f;
        const int flag_every = flags >> 16;                         // test hook: send every n-th system down the LU re-solve
        // a system whose geometry stage already failed (overflow: NaN by the reference's convention too) is NOT sent to the
        // pivoted-LU re-solve: its outputs are NaN whatever that would compute (C5: 8 such systems of 400 modes cost 10 ms there)
        const int fail = sh_fail | (flag_every > 0 && sysi % flag_every == 0);
        const int status = gstatus | ((fail && !gstatus) ? 4 : 0);
        if (tid == 0) { out.m0[sysi] = m0; out.status[sysi] = status; }
        const double ze = linv_last * linv_last;
        const double den = 1.0 + beta * beta * ze * ze;
        for (int t = tid; t < R * R; t += nthr) {
            const int j = t / R, i = t - j * R;
            const double we = fr[(long)i * ldk + K - 1] * linv_last;
            const double xer = we / den, xei = we * beta * ze / den;      // x_e = we (1 + i beta ze) / den
            const double qz = fq[(long)j * ldk + K - 1] * linv_last;
            double qw;
            if (R * R <= 64) qw = sh_dots[t];
            else {   // many modes: recompute serially (rare)
                qw = 0.0;
                for (int k = 0; k < K; ++k) qw = fma(fq[(long)j * ldk + k], fr[(long)i * ldk + k], qw);
            }
            const double cre = 2.0 * kPi * ((s0[j * R + i] + (qw - beta * xei * qz)) + phi[j]);
            const double cim = 2.0 * kPi * (beta * xer * qz);
            double am = cre * v.hdr->rho;
            double dpv = ex.inf_m0 ? 0.0 : cim * omega_of_m0(m0, g.h) * v.hdr->rho;
            if (status) { am = NAN; dpv = NAN; }
            const int64_t o = gd.out_rr + ((int64_t)f * R + j) * R + i;
            out.am[o] = am;
            out.dp[o] = dpv;
            if (i == j) {
                const int64_t o1 = gd.out_r + (int64_t)f * R + j;
                double ph = ex.inf_m0 ? -(kPi / 2.0) : -(kPi / 2.0) + atan2(xei, xer) - hankel0_angle(m0 * g.a[M - 1]);
                double fo = excitation_force(dpv, m0, g.h, kRhoDefault);
                if (status) { ph = NAN; fo = NAN; }
                out.force[o1] = fo;
                out.phase[o1] = ph;
            }
        }
        PHASE_END_S(2);
        // 8. optional X = [y0 + G v ; x2],  v = Imk x2 (Imk entries recomputed: the gemm kernel's slot is gone)
        if (out.x != nullptr && gd.xg_off >= 0) {
            const double* mk = v.roots + rs.off + F + (int64_t)f * rs.K;
            const double* Nk = v.roots + rs.off + F + (int64_t)F * rs.K + (int64_t)f * rs.K;
            double* xc = xpath;                                            // x2_i[k] complex, natural k order
            double* vv = xc + 2 * R * K;                                  // v_i[n] complex
            for (int e = tid; e < R * K; e += nthr) {
                const int i = e / K, kp = e - i * K;
                const int k = (kp == K - 1) ? 0 : kp + 1;
                const double we = xs[(long)i * ldk + K - 1];
                const double xer = we / den, xei = we * beta * ze / den;
                xc[2 * (i * K + k)] = xs[(long)i * ldk + kp] - beta * xei * zv[kp];
                xc[2 * (i * K + k) + 1] = beta * xer * zv[kp];
            }
            __syncthreads();
            for (int t = warp; t < N; t += nwarp) {
                double ar[4] = {0, 0, 0, 0}, ai[4] = {0, 0, 0, 0};        // up to 4 modes per pass
                for (int i0 = 0; i0 < R; i0 += 4) {
                    for (int q = 0; q < 4; ++q) ar[q] = ai[q] = 0.0;
                    for (int k = lane; k < K; k += 32) {
                        const double c = imk(t, k, g, ex, mk, Nk);
                        for (int q = 0; q < 4 && i0 + q < R; ++q) {
                            ar[q] = fma(c, xc[2 * ((i0 + q) * K + k)], ar[q]);
                            ai[q] = fma(c, xc[2 * ((i0 + q) * K + k) + 1], ai[q]);
                        }
                    }
                    for (int q = 0; q < 4 && i0 + q < R; ++q) {
                        const double sr = warp_sum(ar[q]), si = warp_sum(ai[q]);
                        if (lane == 0) { vv[2 * ((i0 + q) * N + t)] = sr; vv[2 * ((i0 + q) * N + t) + 1] = si; }
                    }
                }
            }
            __syncthreads();
            const double* xg = v.xstore + gd.xg_off;
            const int NT = gd.NT;
            for (int t = warp; t < R * p; t += nwarp) {
                const int i = t / p, r = t - i * p;
                const double* grow = xg + (int64_t)r * (N + R);
                double ar = 0.0, ai = 0.0;
                for (int n = lane; n < N; n += 32) {
                    ar = fma(grow[n], vv[2 * (i * N + n)], ar);
                    ai = fma(grow[n], vv[2 * (i * N + n) + 1], ai);
                }
                ar = warp_sum(ar); ai = warp_sum(ai);
                if (lane == 0) {
                    double* xo = out.x + 2 * (gd.out_x + ((int64_t)f * R + i) * NT + r);
                    xo[0] = status ? NAN : grow[N + i] + ar;
                    xo[1] = status ? NAN : ai;
                }
            }
            for (int e = tid; e < R * K; e += nthr) {
                const int i = e / K, k = e - i * K;
                double* xo = out.x + 2 * (gd.out_x + ((int64_t)f * R + i) * NT + p + k);
                xo[0] = status ? NAN : xc[2 * e];
                xo[1] = status ? NAN : xc[2 * e + 1];
            }
            PHASE_END_S(3);
        }
    }
    if (tid == 0) {
        unsigned long long* prof = reinterpret_cast<unsigned long long*>(v.counters + 16);
        for (int i = 0; i < 4; ++i) atomicAdd(&prof[4 + i], (unsigned long long)tph[i]);
    }
}
#undef PHASE_END

// ------------------------------------------------------------------------------------------------
// Schur path, frequency stage for SMALL systems (N <= 16, K <= 16, R <= 4, no X output): one WARP per (geometry, omega)
// system, everything in the warp's slice of shared memory - coupling table, [-r; q; W] = LA^T Imk, P = W^T W - Re D,
// the register-resident 16 x 16 Cholesky (warp_chol16), forward solves as products with the inverse factor, Sherman-
// Morrison closure and coefficients.  No CTA barriers, no HBM round trip between the stages: the design-sweep workload
// (10^5 geometries x 64 omega at 16 modes per region) and the mini_bicylinder replica batch spend ~40 ns of SM time per
// system here instead of a 256-thread CTA with 64 x 128 DMMA tiles per 16 x 16 system.
// Per-system scalars of the one-warp kernel, one THREAD per (geometry, omega): the exterior-region exponentials, the complex
// diagonal entry of the propagating mode (Bessel J/Y of m0 a_M), the Hankel phase and the constants of the force formula.
// In k_freq_small these were evaluated by one lane while 31 waited (about a fifth of its instruction stream); here they run
// at full lane occupancy.  rec[8] per system: tail, sinh(m0 L), Re D_0, Im D_0, arg H0(m0 a), omega(m0), force numerator
// ((2 c rho) g^2), force denominator (omega m0) - the same operation order as excitation_force().
constexpr int kSysRec = 8;
__global__ void k_sys_pre(const void* blob, void* ws, int n_sys) {
    DevView v = make_view(blob, ws);
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_sys) return;
    const int F = v.hdr->n_freq;
    const int gi = t / F, f = t - gi * F;
    const GeomDesc gd = v.geom[gi];
    const Geom g = geom_of(v, gd);
    const RootSet rs = v.sets[gd.root_set];
    const double m0 = v.roots[rs.off + f];
    const double* mk = v.roots + rs.off + F + (int64_t)f * rs.K;
    Exterior e;
    make_exterior(g, m0, e);
    const cplx dg = ext_diag(0, g, e, mk);
    double* rec = v.sysbuf + (int64_t)(gd.out_sys + f) * kSysRec;
    rec[0] = e.tail; rec[1] = e.sinh_m0L; rec[2] = dg.re; rec[3] = dg.im;
    rec[4] = e.inf_m0 ? 0.0 : hankel0_angle(m0 * g.a[g.M - 1]);
    const double om = omega_of_m0(m0, g.h);
    rec[5] = om;
    const double T = tanh(m0 * g.h);
    const double c = T + m0 * g.h * (1.0 - T * T);
    rec[6] = 2.0 * c * kRhoDefault * (kG * kG);
    rec[7] = om * m0;
}

constexpr int kSmallMax = 16, kSmallMaxR = 4, kSmallWarps = 4;
// e / d for the small index ranges of the one-warp kernel (e < 1024, d <= 32) by a float multiply: the quotient's fraction
// (r + 1/2) / d stays >= 1/64 away from an integer, far beyond the float rounding error, so the floor is exact and the
// ~20-instruction integer division disappears from the inner loops.
__device__ __forceinline__ int fdiv_small(int e, float rcp) { return (int)(((float)e + 0.5f) * rcp); }
constexpr int kSmallLdA = kSmallMax + 2 * kSmallMaxR + 1;   // 25: odd stride, conflict-free column reads
struct SmallWs {
    union {                                     // the geometry's operand block (steps 1-2), then the Cholesky block (3-4)
        double A[kSmallMax][kSmallLdA];
        double Dg[kCholNB * (kCholNB + 1) + kCholNB];
    };
    union {                                     // coupling table (steps 1-3), then the inverse factor (4-6)
        double Imk[kSmallMax][kSmallMax + 1];
        double Li[kCholNB * kCholLdP];
    };
    double T[kSmallMax][kSmallMax + 1];
    double aug[2 * kSmallMaxR][kSmallMax];      // rows r_i then q_j; overwritten by their forward-solved forms
    double ext[kSmallMax];
    Exterior ex;
    double beta;
    int fail;
};

__global__ void __launch_bounds__(32 * kSmallWarps, 6) k_freq_small(const void* blob, void* ws, OutPtrs out, int n_items, int* counter,
                                                                   int flag_every) {
    extern __shared__ __align__(16) unsigned char smem[];
    DevView v = make_view(blob, ws);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    SmallWs& S = reinterpret_cast<SmallWs*>(smem)[warp];
    const int F = v.hdr->n_freq;
    for (;;) {
        int item = 0;
        if (lane == 0) item = atomicAdd(counter, 1);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= n_items) break;
        const int gi = v.order[item / F], f = item % F;
        const GeomDesc gd = v.geom[gi];
        const Geom g = geom_of(v, gd);
        const RootSet rs = v.sets[gd.root_set];
        const double m0 = v.roots[rs.off + f];
        const double* mk = v.roots + rs.off + F + (int64_t)f * rs.K;
        const double* Nk = v.roots + rs.off + F + (int64_t)F * rs.K + (int64_t)f * rs.K;
        const int N = gd.N, K = gd.K, R = gd.R, M = gd.M;
        const int ldn = gs_ldn(N, R);
        const double* gs = v.gstore + gd.gs_off;
        const double* HT = gs;
        const double* s0 = gs + (int64_t)N * ldn + 2LL * R * N;
        const double* phi = s0 + R * R;
        const double* bc = phi + R;
        const int gstatus = (int)bc[R];
        const int hrank = (int)bc[R + 1];
        const bool lmode = hrank > 0;
        const double* LA = gs + gs_base_doubles(N, R);
        const int* perm = reinterpret_cast<const int*>(LA + (int64_t)N * ldn);
        __syncwarp();
        const double* rec = v.sysbuf + (int64_t)(gd.out_sys + f) * kSysRec;       // k_sys_pre
        if (lane == 0) {
            S.ex.m0 = m0; S.ex.h = g.h; S.ex.m0h = m0 * g.h;
            S.ex.inf_m0 = isinf(m0); S.ex.lowbranch = S.ex.m0h < kM0HThresh;
            S.ex.tail = rec[0]; S.ex.sinh_m0L = rec[1];
            S.fail = 0;
        }
        // the geometry's operand block into shared memory (coalesced, one latency): LA (factorised H) or HT
        const int rows_out = (lmode ? hrank : N) + 2 * R;
        const float rl_ldn = 1.0f / (float)ldn, rl_k = 1.0f / (float)K, rl_k1 = 1.0f / (float)max(K - 1, 1);
        {
            const double* Ag = lmode ? LA : HT;
            for (int e = lane; e < N * ldn; e += 32) {
                const int k = fdiv_small(e, rl_ldn), m = e - k * ldn;
                if (m < rows_out) S.A[k][m] = Ag[e];
            }
        }
        __syncwarp();
        // 1. coupling table, column order kp = k - 1, k = 0 last.  Interior entries by the angle-addition form of
        // k_freq_gemm (2K + 2N sines / cosines tabulated per system, sinc series near a resonance); the row of harmonic 0
        // and the column k = 0 through the reference-ordered imk()
        if (!S.ex.inf_m0 && K > 1) {
            double* tb = &S.T[0][0];                          // T is not live yet
            double *tmk = tb, *tsk = tb + 16, *tck = tb + 32, *tnk = tb + 48, *tlm = tb + 64, *tsl = tb + 80, *tcl = tb + 96;
            const double Lh = g.h - g.d[M - 1];
            {
                const bool isk = lane < 16;
                const int idx = isk ? lane : lane - 16;
                if (isk ? (idx < K) : (idx < N)) {
                    const double val = isk ? mk[(idx == K - 1) ? 0 : idx + 1] : lambda_ni(lmode ? perm[idx] : idx, M - 1, g);
                    double sv, cv;
                    const double pr = val * Lh, er = fma(val, Lh, -pr);     // exact product = pr + er
                    sincos(pr, &sv, &cv);
                    if (isk) { tmk[idx] = val; tsk[idx] = fma(er, cv, sv); tck[idx] = fma(-er, sv, cv);
                               tnk[idx] = sqrt(2.0 / Nk[(idx == K - 1) ? 0 : idx + 1]) * 0.5; }
                    else { tlm[idx] = val; tsl[idx] = fma(er, cv, sv); tcl[idx] = fma(-er, sv, cv); }
                }
            }
            const unsigned zmask = __ballot_sync(0xffffffffu, lane < N && (lmode ? perm[lane] : lane) == 0);
            const int row0 = __ffs(zmask) - 1;
            __syncwarp();
            for (int e = lane; e < N + K - 1; e += 32) {
                if (e < N) S.Imk[e][K - 1] = imk(lmode ? perm[e] : e, 0, g, S.ex, mk, Nk);
                else S.Imk[row0][e - N] = imk(0, e - N + 1, g, S.ex, mk, Nk);
            }
            for (int e = lane; e < N * (K - 1); e += 32) {
                const int n = fdiv_small(e, rl_k1), kp = e - n * (K - 1);
                if (n == row0) continue;
                const double lam = tlm[n], sl = tsl[n], cl = tcl[n];
                const double mv = tmk[kp];
                const double a = tsk[kp] * cl, b = tck[kp] * sl;
                const double sp = mv + lam, sd = mv - lam;
                const double x = sd * Lh, x2 = x * x;
                double Ssin = a - b;
                if (fabs(x) < 1.0) {
                    double pl = -8.22063524662432972e-18;                       // -1/19!
                    pl = fma(pl, x2, 2.81145725434552076e-15);                  //  1/17!
                    pl = fma(pl, x2, -7.64716373181981648e-13);                 // -1/15!
                    pl = fma(pl, x2, 1.60590438368216146e-10);                  //  1/13!
                    pl = fma(pl, x2, -2.50521083854417188e-08);                 // -1/11!
                    pl = fma(pl, x2, 2.75573192239858907e-06);                  //  1/9!
                    pl = fma(pl, x2, -1.98412698412698413e-04);                 // -1/7!
                    pl = fma(pl, x2, 8.33333333333333333e-03);                  //  1/5!
                    pl = fma(pl, x2, -1.66666666666666667e-01);                 // -1/3!
                    Ssin = fma(x * x2, pl, x);
                }
                double val = tnk[kp] * (fma(a + b, sd, Ssin * sp) / (sp * sd));
                if (fabs(mv) == lam) val = sqrt(2.0 / Nk[kp + 1]) * Lh / 2.0;
                S.Imk[n][kp] = val;
            }
        } else {
            for (int e = lane; e < N * K; e += 32) {
                const int kk = e / K, kp = e - kk * K;
                S.Imk[kk][kp] = imk(lmode ? perm[kk] : kk, (kp == K - 1) ? 0 : kp + 1, g, S.ex, mk, Nk);
            }
        }
        if (lane < K - 1) S.ext[lane] = ext_diag(lane + 1, g, S.ex, mk).re;       // evanescent modes: K1e / K0e ratios
        else if (lane == K - 1) { S.ext[lane] = rec[2]; S.beta = rec[3]; }        // propagating mode: complex, from k_sys_pre
        __syncwarp();
        // 2. [-r_i ; q_j ; W] (factorised H) or [T ; -r_i ; q_j] (H itself): out[m][n] = sum_k A[k][m] Imk[k][n]
        const int nrow_t = lmode ? hrank : N;                 // rows of T / W
        for (int e = lane; e < rows_out * K; e += 32) {
            const int m = fdiv_small(e, rl_k), n = e - m * K;
            double acc = 0.0;
            // factorised form: column 2R + j of LA is column j of the lower-trapezoidal factor, zero above row j
            for (int k = (lmode && m >= 2 * R) ? m - 2 * R : 0; k < N; ++k) acc = fma(S.A[k][m], S.Imk[k][n], acc);
            int a = -1, trow = -1;                            // destination: aug row a (sign) or T row
            if (lmode) { if (m < 2 * R) a = m; else trow = m - 2 * R; }
            else { if (m < N) trow = m; else a = m - N; }
            if (trow >= 0) S.T[trow][n] = acc;
            else S.aug[a][n] = (a < R) ? -acc : acc;
        }
        __syncwarp();
        // 3. P = A2^T T - Re(D), lower triangle, identity-padded to 16 x 16 in warp_chol16's layout
        // padding first (upper triangle of the K x K block: 0, rows >= K: identity), then the K (K + 1) / 2 lower entries dealt
        // to the lanes as a packed triangle, so that no lane idles on an upper-triangle slot
        for (int e = lane; e < kCholNB * kCholNB; e += 32) {
            const int m = e >> 4, n = e & 15;
            if (!(m < K && n <= m)) S.Dg[m * (kCholNB + 1) + n] = (m >= K && m == n) ? 1.0 : 0.0;
        }
        for (int e = lane; e < K * (K + 1) / 2; e += 32) {
            int m = (int)((sqrtf(8.0f * (float)e + 1.0f) - 1.0f) * 0.5f);
            if (m * (m + 1) / 2 > e) --m;                      // guard the float square root at the row boundaries
            if ((m + 1) * (m + 2) / 2 <= e) ++m;
            const int n = e - m * (m + 1) / 2;
            double acc = 0.0;
            if (lmode) for (int k = 0; k < nrow_t; ++k) acc = fma(S.T[k][m], S.T[k][n], acc);
            else for (int k = 0; k < N; ++k) acc = fma(S.Imk[k][m], S.T[k][n], acc);
            S.Dg[m * (kCholNB + 1) + n] = (m == n) ? acc - S.ext[m] : acc;
        }
        __syncwarp();
        // 4. P = L L^T and L^{-1} in one warp
        warp_chol16(S.Dg, S.Li, &S.fail);
        __syncwarp();
        // 5. forward solves fr_i = L^{-1} r_i, fq_j = L^{-1} q_j as products with the inverse factor
        double fw[(2 * kSmallMaxR * kSmallMax + 31) / 32];
#pragma unroll
        for (int u = 0; u < (2 * kSmallMaxR * kSmallMax + 31) / 32; ++u) {
            const int e = lane + 32 * u, a = e >> 4, c = e & 15;
            double acc = 0.0;
            if (a < 2 * R && c < K)
                for (int q = 0; q <= c; ++q) acc = fma(S.Li[c * kCholLdP + q], S.aug[a][q], acc);
            fw[u] = acc;
        }
        __syncwarp();
#pragma unroll
        for (int u = 0; u < (2 * kSmallMaxR * kSmallMax + 31) / 32; ++u) {
            const int e = lane + 32 * u, a = e >> 4, c = e & 15;
            if (a < 2 * R && c < K) S.aug[a][c] = fw[u];
        }
        __syncwarp();
        // 6. coefficients (same closure as k_freq_solve)
        const int64_t sysi = gd.out_sys + f;
        const bool hook = flag_every > 0 && sysi % flag_every == 0;
        const int status = gstatus | (((S.fail || hook) && !gstatus) ? 4 : 0);
        if (lane == 0) { out.m0[sysi] = m0; out.status[sysi] = status; }
        const double linv_last = S.Li[(K - 1) * kCholLdP + K - 1];
        const double beta = S.beta;
        const double ze = linv_last * linv_last;
        const double den = 1.0 + beta * beta * ze * ze;
        if (lane < R * R) {
            const int j = lane / R, i = lane - j * R;
            const double* fr = S.aug[i];
            const double* fq = S.aug[R + j];
            const double we = fr[K - 1] * linv_last;
            const double xer = we / den, xei = we * beta * ze / den;
            const double qz = fq[K - 1] * linv_last;
            double qw = 0.0;
            for (int k = 0; k < K; ++k) qw = fma(fq[k], fr[k], qw);
            const double cre = 2.0 * kPi * ((s0[j * R + i] + (qw - beta * xei * qz)) + phi[j]);
            const double cim = 2.0 * kPi * (beta * xer * qz);
            double am = cre * v.hdr->rho;
            double dpv = S.ex.inf_m0 ? 0.0 : cim * rec[5] * v.hdr->rho;
            if (status) { am = NAN; dpv = NAN; }
            const int64_t o = gd.out_rr + ((int64_t)f * R + j) * R + i;
            out.am[o] = am;
            out.dp[o] = dpv;
            if (i == j) {
                const int64_t o1 = gd.out_r + (int64_t)f * R + j;
                double ph = S.ex.inf_m0 ? -(kPi / 2.0) : -(kPi / 2.0) + atan2(xei, xer) - rec[4];
                double fo = sqrt((rec[6] * dpv) / rec[7]);            // excitation_force(), constants from k_sys_pre
                if (status) { ph = NAN; fo = NAN; }
                out.force[o1] = fo;
                out.phase[o1] = ph;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// dense path: the reference's N_T x N_T complex system, one CTA (512 threads) per (geometry, frequency)
template <int NB>
__global__ void __launch_bounds__(512, 1) k_dense(const void* blob, void* ws, OutPtrs out) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ Layout L;
    __shared__ Exterior ex;
    __shared__ int sh_item, sh_status;
    DevView v = make_view(blob, ws);
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    const int F = v.hdr->n_freq;
    const int64_t total = (int64_t)v.hdr->n_geom * F;
    cplx* W = reinterpret_cast<cplx*>(v.slots + (int64_t)blockIdx.x * v.hdr->slot_doubles_dense);
    for (;;) {
        const int item = next_item(&v.counters[1], &sh_item);
        if (item >= total) break;
        const int gi = v.order[item / F], f = item % F;
        const GeomDesc gd = v.geom[gi];
        const Geom g = geom_of(v, gd);
        const RootSet rs = v.sets[gd.root_set];
        const double m0 = v.roots[rs.off + f];
        const double* mk = v.roots + rs.off + F + (int64_t)f * rs.K;
        const double* Nk = v.roots + rs.off + F + (int64_t)F * rs.K + (int64_t)f * rs.K;
        if (tid == 0) { make_layout(g, L); make_exterior(g, m0, ex); sh_status = 0; }
        __syncthreads();
        const int K = gd.K, R = gd.R, M = gd.M, p = gd.p, NT = gd.NT;
        const long ld = NT + R;
        const double* tab = v.tables + gd.tab_off;
        const uint8_t* flags = v.flags + gd.flag_off;
        cta_zero(W, (int64_t)NT * ld);
        __syncthreads();
        fill_A_static<cplx>(g, L, tab, W, ld, tid, nthr);
        fill_A_omega(g, L, tab, ex, mk, Nk, W, ld, tid, nthr);
        for (int i = 0; i < R; ++i) {
            const uint8_t* hv = flags + (int64_t)i * (M + 1);
            fill_b_static<cplx>(g, L, hv, W + NT + i, ld, tid, nthr);
            const double cst = b_end_constant(g, hv);
            for (int k = tid; k < K; k += nthr)
                set_real(W[(long)(L.vel_row[M - 1] + k) * ld + NT + i], cst * imk(0, k, g, ex, mk, Nk));
        }
        __syncthreads();
        lu_forward<cplx, NB>(W, ld, NT, NT, NT + R, smem, &sh_status);
        lu_backward<cplx, NB>(W, ld, NT, NT, NT + R, smem, NT);
        __syncthreads();
        const int status = sh_status;
        const int64_t sys = gd.out_sys + f;
        if (tid == 0) { out.m0[sys] = m0; out.status[sys] = status; }
        const double hk_angle = ex.inf_m0 ? 0.0 : hankel0_angle(m0 * g.a[M - 1]);
        // coefficients: coef_ji = 2 pi (c_j . X_i[:p] + phi_j)
        for (int t = warp; t < R * R; t += nwarp) {
            const int j = t / R, i = t - j * R;
            const uint8_t* hvj = flags + (int64_t)j * (M + 1);
            double ar = 0.0, ai = 0.0;
            for (int reg = 0; reg < M; ++reg) {
                if (!hvj[reg]) continue;
                const int Nr = g.nmk[reg];
                const double* T8 = tab + L.tab[reg] + 8 * Nr;
                for (int n = lane; n < Nr; n += 32) {
                    cplx x = W[(long)(L.col[reg] + n) * ld + NT + i];
                    ar = fma(T8[n], x.re, ar); ai = fma(T8[n], x.im, ai);
                    if (reg > 0) {
                        x = W[(long)(L.col[reg] + Nr + n) * ld + NT + i];
                        ar = fma(T8[Nr + n], x.re, ar); ai = fma(T8[Nr + n], x.im, ai);
                    }
                }
            }
            ar = warp_sum(ar); ai = warp_sum(ai);
            if (lane == 0) {
                const double cre = 2.0 * kPi * (ar + phi_sum(g, L, tab, hvj));
                const double cim = 2.0 * kPi * ai;
                double am = cre * v.hdr->rho;
                double dpv = ex.inf_m0 ? 0.0 : cim * omega_of_m0(m0, g.h) * v.hdr->rho;
                if (status) { am = NAN; dpv = NAN; }
                const int64_t o = gd.out_rr + ((int64_t)f * R + j) * R + i;
                out.am[o] = am;
                out.dp[o] = dpv;
                if (i == j) {
                    const int64_t o1 = gd.out_r + (int64_t)f * R + j;
                    const cplx xe = W[(long)p * ld + NT + i];
                    double ph = ex.inf_m0 ? -(kPi / 2.0) : -(kPi / 2.0) + atan2(xe.im, xe.re) - hk_angle;
                    double fo = excitation_force(dpv, m0, g.h, kRhoDefault);
                    if (status) { ph = NAN; fo = NAN; }
                    out.force[o1] = fo;
                    out.phase[o1] = ph;
                }
            }
        }
        if (out.x != nullptr) {
            for (int e = tid; e < R * NT; e += nthr) {
                const int i = e / NT, r = e - i * NT;
                const cplx x = W[(long)r * ld + NT + i];
                double* xo = out.x + 2 * (gd.out_x + ((int64_t)f * R + i) * NT + r);
                xo[0] = status ? NAN : x.re;
                xo[1] = status ? NAN : x.im;
            }
        }
    }
}

// debug export: dense A, b, c of one (geometry, frequency)
__global__ void k_assemble(const void* blob, void* ws, int gi, int f, cplx* A, cplx* b, cplx* c) {
    __shared__ Layout L;
    __shared__ Exterior ex;
    DevView v = make_view(blob, ws);
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int F = v.hdr->n_freq;
    const GeomDesc gd = v.geom[gi];
    const Geom g = geom_of(v, gd);
    const RootSet rs = v.sets[gd.root_set];
    const double m0 = v.roots[rs.off + f];
    const double* mk = v.roots + rs.off + F + (int64_t)f * rs.K;
    const double* Nk = v.roots + rs.off + F + (int64_t)F * rs.K + (int64_t)f * rs.K;
    if (tid == 0) { make_layout(g, L); make_exterior(g, m0, ex); }
    __syncthreads();
    const int NT = gd.NT, R = gd.R, M = gd.M, p = gd.p, K = gd.K;
    const double* tab = v.tables + gd.tab_off;
    const uint8_t* flags = v.flags + gd.flag_off;
    if (A) {
        cta_zero(A, (int64_t)NT * NT);
        __syncthreads();
        fill_A_static<cplx>(g, L, tab, A, NT, tid, nthr);
        fill_A_omega(g, L, tab, ex, mk, Nk, A, NT, tid, nthr);
    }
    for (int i = 0; i < R; ++i) {
        const uint8_t* hv = flags + (int64_t)i * (M + 1);
        if (b) {
            cplx* bi = b + (int64_t)i * NT;
            cta_zero(bi, NT);
            __syncthreads();
            fill_b_static<cplx>(g, L, hv, bi, 1, tid, nthr);
            const double cst = b_end_constant(g, hv);
            for (int k = tid; k < K; k += nthr) set_real(bi[L.vel_row[M - 1] + k], cst * imk(0, k, g, ex, mk, Nk));
        }
        if (c) {   // c_j[0:p] followed by sum_i heaving_j[i] * int_phi_p_i (eng.py:370-376)
            fill_c<cplx>(g, L, tab, hv, c + (int64_t)i * (p + 1), tid, nthr);
            if (tid == 0) set_real(c[(int64_t)i * (p + 1) + p], phi_sum(g, L, tab, hv));
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Field evaluation: potential and velocity of the homogeneous solution and of the particular
// solution at arbitrary points, one warp per point, lanes over the harmonics.  Mirrors
// MEEMEngine.calculate_potentials / calculate_velocities (eng.py:395-467, :478-553) with the radial / vertical
// eigenfunctions of me.py:192-416 in the reference's operation order; the region is chosen by r alone and the
// reference's "inside the body" masks are applied by the caller (they are index bookkeeping, not arithmetic).
// out: 9 planes of n_points doubles: Re/Im phiH, Re/Im vrH, Re/Im vzH, phiP, vrP, vzP.
__global__ void __launch_bounds__(256) k_fields(const void* blob, void* ws, int gi, int f, int mode, const double* X,
                                                long n_points, const double* rp, const double* zp, double* out) {
    __shared__ Layout L;
    __shared__ Exterior ex;
    DevView v = make_view(blob, ws);
    const int F = v.hdr->n_freq;
    const GeomDesc gd = v.geom[gi];
    const Geom g = geom_of(v, gd);
    const RootSet rs = v.sets[gd.root_set];
    const double m0 = v.roots[rs.off + f];
    const double* mk = v.roots + rs.off + F + (int64_t)f * rs.K;
    const double* Nk = v.roots + rs.off + F + (int64_t)F * rs.K + (int64_t)f * rs.K;
    if (threadIdx.x == 0) { make_layout(g, L); make_exterior(g, m0, ex); }
    __syncthreads();
    const uint8_t* hv = v.flags + gd.flag_off + (int64_t)mode * (gd.M + 1);
    const int lane = threadIdx.x & 31;
    const long wid = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((long)gridDim.x * blockDim.x) >> 5;
    const int M = g.M;
    const double h = g.h;
    for (long pt = wid; pt < n_points; pt += nw) {
        const double r = rp[pt], z = zp[pt];
        int reg = M;
        for (int i = M - 1; i >= 0; --i) if (r <= g.a[i]) reg = i;
        cplx pH = cmake(0, 0), vr = cmake(0, 0), vz = cmake(0, 0);
        double pP = 0.0, vrP = 0.0, vzP = 0.0;
        if (reg < M) {
            const int i = reg, Ni = g.nmk[i], col = L.col[i];
            const double ai = g.a[i];
            for (int n = lane; n < Ni; n += 32) {
                const cplx c1 = cmake(X[2 * (col + n)], X[2 * (col + n) + 1]);
                const cplx c2 = (i > 0) ? cmake(X[2 * (col + Ni + n)], X[2 * (col + Ni + n) + 1]) : cmake(0, 0);
                double R1, dR1, R2 = 0.0, dR2 = 0.0, Zn, dZn;
                if (n == 0) {
                    R1 = 0.5; dR1 = 0.0; Zn = 1.0; dZn = 0.0;
                    if (i > 0) { R2 = 0.5 * log(r / ai); dR2 = (r != 0.0) ? 1.0 / (2.0 * r) : INFINITY; }
                } else {
                    const double lam = lambda_ni(n, i, g);
                    const double den1 = i0e(lam * ai), e1 = exp(lam * (r - ai));
                    R1 = (r == ai) ? 1.0 : (i0e(lam * r) / den1) * e1;
                    dR1 = ((lam * i1e(lam * r)) / den1) * e1;
                    if (i > 0) {
                        const double den2 = k0e(lam * ai), e2 = exp(lam * (ai - r));
                        R2 = (r == ai) ? 1.0 : (k0e(lam * r) / den2) * e2;
                        dR2 = ((-lam * k1e(lam * r)) / den2) * e2;
                    }
                    const double arg = lam * (z + h);
                    Zn = kSqrt2 * cos(arg);
                    dZn = -lam * kSqrt2 * sin(arg);
                }
                const cplx t = cmake(c1.re * R1 + c2.re * R2, c1.im * R1 + c2.im * R2);
                const cplx dt = cmake(c1.re * dR1 + c2.re * dR2, c1.im * dR1 + c2.im * dR2);
                pH.re += t.re * Zn; pH.im += t.im * Zn;
                vr.re += dt.re * Zn; vr.im += dt.im * Zn;
                vz.re += t.re * dZn; vz.im += t.im * dZn;
            }
            if (hv[i]) {                                  // me.py:179-186
                const double hd = h - g.d[i];
                pP = (1.0 / (2.0 * hd)) * ((z + h) * (z + h) - (r * r) / 2.0);
                vrP = -r / (2.0 * hd);
                vzP = (z + h) / hd;
            }
        } else {
            const int K = L.K, p = L.p;
            const double aM = g.a[M - 1];
            for (int k = lane; k < K; k += 32) {
                const cplx c = cmake(X[2 * (p + k)], X[2 * (p + k) + 1]);
                cplx Lam, dLam;
                double Zk, dZk;
                if (k == 0) {
                    double j0a, y0a, j1a, y1a, j0, y0, j1, y1;
                    bessel_jy01(m0 * aM, &j0a, &y0a, &j1a, &y1a);
                    bessel_jy01(m0 * r, &j0, &y0, &j1, &y1);
                    const cplx H0a = cmake(j0a, y0a);
                    Lam = (r == aM) ? cmake(1.0, 0.0) : cmake(j0, y0) / H0a;
                    dLam = cmake(-(m0 * j1), -(m0 * y1)) / H0a;
                    if (ex.lowbranch) {
                        Zk = (1.0 / sqrt(Nk[0])) * cosh(m0 * (z + h));
                        dZk = (1.0 / sqrt(Nk[0])) * m0 * sinh(m0 * (z + h));
                    } else {
                        Zk = sqrt(2.0 * m0 * h) * (exp(m0 * z) + exp(-m0 * (z + 2.0 * h)));
                        dZk = m0 * sqrt(2.0 * h * m0) * (exp(m0 * z) - exp(-m0 * (z + 2.0 * h)));
                    }
                } else {
                    const double m = mk[k];
                    const double den = k0e(m * aM), e = exp(m * (aM - r));
                    Lam = cmake((r == aM) ? 1.0 : (k0e(m * r) / den) * e, 0.0);
                    dLam = cmake(((-(m * k1e(m * r))) / den) * e, 0.0);
                    const double isn = 1.0 / sqrt(Nk[k]);
                    Zk = isn * cos(m * (z + h));
                    dZk = -isn * m * sin(m * (z + h));
                }
                const cplx cl = c * Lam, cd = c * dLam;
                pH.re += cl.re * Zk; pH.im += cl.im * Zk;
                vr.re += cd.re * Zk; vr.im += cd.im * Zk;
                vz.re += cl.re * dZk; vz.im += cl.im * dZk;
            }
        }
        pH.re = warp_sum(pH.re); pH.im = warp_sum(pH.im);
        vr.re = warp_sum(vr.re); vr.im = warp_sum(vr.im);
        vz.re = warp_sum(vz.re); vz.im = warp_sum(vz.im);
        if (lane == 0) {
            out[0 * n_points + pt] = pH.re; out[1 * n_points + pt] = pH.im;
            out[2 * n_points + pt] = vr.re; out[3 * n_points + pt] = vr.im;
            out[4 * n_points + pt] = vz.re; out[5 * n_points + pt] = vz.im;
            out[6 * n_points + pt] = pP; out[7 * n_points + pt] = vrP; out[8 * n_points + pt] = vzP;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// FP64 throughput microbenchmarks (roofline denominator)
__global__ void k_dfma_peak(double* out, int iters) {
    double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double b = 1.0000001, c = 1e-9;
    for (int i = 0; i < iters; ++i) {
        a0 = fma(a0, b, c); a1 = fma(a1, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c);
        a4 = fma(a4, b, c); a5 = fma(a5, b, c); a6 = fma(a6, b, c); a7 = fma(a7, b, c);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

__global__ void k_dmma_peak(double* out, int iters) {
    double a = threadIdx.x * 1e-3, b = 1.0 + threadIdx.x * 1e-6;
    double c0[2] = {0, 0}, c1[2] = {0, 0}, c2[2] = {0, 0}, c3[2] = {0, 0};
    for (int i = 0; i < iters; ++i) {
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0[0]), "+d"(c0[1]) : "d"(a), "d"(b));
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c1[0]), "+d"(c1[1]) : "d"(a), "d"(b));
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c2[0]), "+d"(c2[1]) : "d"(a), "d"(b));
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c3[0]), "+d"(c3[1]) : "d"(a), "d"(b));
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = c0[0] + c0[1] + c1[0] + c1[1] + c2[0] + c2[1] + c3[0] + c3[1];
}

// ================================================================================================
// host side
static thread_local std::string g_err;
static int fail(int code, const std::string& msg) { g_err = msg; return code; }
#define CUDA_TRY(expr)                                                                     \
    do {                                                                                   \
        cudaError_t e_ = (expr);                                                           \
        if (e_ != cudaSuccess) return fail(MEEMB200_ECUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); \
    } while (0)

struct meemb200_ctx {
    int device;
    int sm_count;
    size_t smem_optin;
    cudaStream_t side = nullptr;       // one side stream per context (created at the first solve that uses it), shared by its plans:
    std::mutex side_mu;                // a stream per plan made every one-problem call of the class API pay a stream creation
};
// Launches go to the context's device whatever the caller's current device is; the caller's choice is restored.
struct DeviceGuard {
    int prev = -1;
    bool switched = false;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) == cudaSuccess && prev != dev) switched = cudaSetDevice(dev) == cudaSuccess;
    }
    ~DeviceGuard() { if (switched) cudaSetDevice(prev); }
};

struct meemb200_plan {
    meemb200_ctx* ctx;
    std::vector<unsigned char> blob;
    BlobHeader* hdr() { return reinterpret_cast<BlobHeader*>(blob.data()); }
    const BlobHeader* hdr() const { return reinterpret_cast<const BlobHeader*>(blob.data()); }
    MeemPlanInfo info;
    std::vector<GeomDesc> geoms;
    int n_slots_geom, n_slots_work, geom_threads;
    bool use_small;         // frequency stage by k_freq_small (one warp per system): N, K <= 16, R <= 4, no X output
    int max_N;
    int gemm_flags;         // bit 0: k_freq_gemm uses rectangular tiles for P = W^T W instead of the warp-tile-queue SYRK;
                            // bit 1: k_freq_solve uses the right-looking Cholesky also when P lives in L2 (tuning / tests)
    int nb_geom, nb_freq, nb_dense;            // panel widths chosen from shared-memory limits
    size_t smem_geom, smem_freq, smem_dense, smem_chol, smem_coop, smem_hchol;
    int algo;
    bool use_coop;          // grid-cooperative geometry stage (few large geometries)
    bool use_coop2;         // ... by window elimination with static pivots for the C1 blocks (k_geom_coop2): every boundary case 1, no X
    int coop2_wrows, coop2_wcols;
    size_t smem_coop2;
    int band_variant;       // geometry stage by window elimination (k_geom_band): -1 off, 0 small window in shared memory (128 threads),
                            // 1 window in shared memory (256 threads), 2 window in the CTA's global slot (512 threads)
    int band_wrows, band_wcols, band_grid, band_side_alias;
    bool band_static;       // k_geom_band eliminates the C1 block of case-1 windows on its diagonal (no pivot search); MEEMB200_NO_BAND_STATIC
    size_t smem_band;
    int64_t wave_items;     // systems per frequency-stage wave
    int n_waves, solve_grid;
    int solve_variant;      // k_freq_solve: 0 = P in its L2-resident buffer (128 threads), 1/2/3 = P resident in shared memory, 128/256/512 threads
    int max_tab_threads;
    int64_t max_FK;
    cudaEvent_t ev[6];      // 0 start, 1 after roots, 2 after geometry, 3 after last k_freq_gemm, 4 after last k_freq_solve, 5 end
    bool ev_ok;
    cudaStream_t side = nullptr;       // the context's side stream: roots run there, concurrently with the geometry stage (meemb200_solve)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    bool no_side = false;              // MEEMB200_NO_ROOTS_OVERLAP
    char* host_call_base;   // device allocation of meemb200_solve_host (the one call that owns device memory), kept for the next call
    size_t host_call_bytes;
};

extern "C" int meemb200_version(void) { return 200; }
extern "C" const char* meemb200_last_error(void) { return g_err.c_str(); }

extern "C" int meemb200_create(int device, meemb200_ctx** out) {
    if (!out) return fail(MEEMB200_EINVAL, "ctx pointer is NULL");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) return fail(MEEMB200_ENODEV, "no CUDA device: meemb200 has no CPU fallback");
    if (device < 0 || device >= n) return fail(MEEMB200_EINVAL, "device index out of range");
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return fail(MEEMB200_ENODEV, std::string("device is sm_") + std::to_string(prop.major * 10 + prop.minor) + ", this library is built for sm_100a only");
    meemb200_ctx* c = new meemb200_ctx;
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    c->smem_optin = prop.sharedMemPerBlockOptin;
    *out = c;
    return MEEMB200_OK;
}
extern "C" int meemb200_destroy(meemb200_ctx* c) {
    if (c && c->side) { DeviceGuard guard(c->device); cudaStreamDestroy(c->side); }
    delete c;
    return MEEMB200_OK;
}
extern "C" int meemb200_sm_count(const meemb200_ctx* c) { return c ? c->sm_count : MEEMB200_EINVAL; }

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

template <class T>
static int pick_nb(int max_rows, size_t limit, size_t reserve) {
    if (lu_smem_bytes<T, 8>(max_rows) + reserve <= limit) return 8;
    if (lu_smem_bytes<T, 4>(max_rows) + reserve <= limit) return 4;
    if (lu_smem_bytes<T, 2>(max_rows) + reserve <= limit) return 2;
    return 0;
}
template <class T>
static size_t nb_bytes(int nb, int rows) {
    return nb == 8 ? lu_smem_bytes<T, 8>(rows) : nb == 4 ? lu_smem_bytes<T, 4>(rows) : lu_smem_bytes<T, 2>(rows);
}

extern "C" int meemb200_plan_create(meemb200_ctx* ctx, const MeemBatch* b, meemb200_plan** out) {
    if (!ctx || !b || !out) return fail(MEEMB200_EINVAL, "NULL argument");
    if (b->n_geom < 1 || b->n_freq < 1) return fail(MEEMB200_EINVAL, "n_geom and n_freq must be >= 1");
    if (!b->reg_off || !b->h || !b->a || !b->d || !b->nmk || !b->n_modes || !b->flag_off || !b->heaving || !b->freq)
        return fail(MEEMB200_EINVAL, "NULL array in MeemBatch");
    if (b->algo != MEEMB200_ALGO_SCHUR && b->algo != MEEMB200_ALGO_DENSE && b->algo != MEEMB200_ALGO_SCHUR_LU) return fail(MEEMB200_EINVAL, "unknown algo");
    const bool schur = b->algo != MEEMB200_ALGO_DENSE;
    if (b->root_mode != MEEMB200_ROOTS_REFERENCE && b->root_mode != MEEMB200_ROOTS_CONVERGED) return fail(MEEMB200_EINVAL, "unknown root_mode");
    if (b->freq_kind != MEEMB200_FREQ_OMEGA && b->freq_kind != MEEMB200_FREQ_M0) return fail(MEEMB200_EINVAL, "unknown freq_kind");
    const int G = b->n_geom, F = b->n_freq;
    if ((int64_t)G * F > 0x7fffffffLL) return fail(MEEMB200_EINVAL, "n_geom * n_freq must fit 31 bits (work-item indices are int32): split the batch");
    for (int f = 0; f < F; ++f) {
        const double w = b->freq[f];
        if (!(w > 0.0)) return fail(MEEMB200_EINVAL, "frequencies must be positive (meem_problem.py:31)");
        if (std::isinf(w) && b->freq_kind == MEEMB200_FREQ_OMEGA) return fail(MEEMB200_EINVAL, "omega must be finite; pass m0=inf with FREQ_M0");
    }
    meemb200_plan* P = new meemb200_plan;
    P->ctx = ctx;
    P->ev_ok = false;
    P->host_call_base = nullptr; P->host_call_bytes = 0;
    P->geoms.resize(G);
    std::vector<RootSet> sets;
    std::map<std::pair<uint64_t, int>, int> set_index;
    std::vector<double> cost(G);
    int64_t n_rr = 0, n_r = 0, n_x = 0, n_c = 0, tab_tot = 0, gs_tot = 0, xg_tot = 0;
    int64_t slot_geom = 0, slot_freq = 0, slot_dense = 0;
    int max_p = 1, max_K = 1, max_NT = 1, max_N = 1;
    int band_wrows = 0, band_wcols = 0, band_side = 0;
    int coop2_wrows = 0, coop2_wcols = 0;
    bool all_case1 = true;
    double flops = 0.0;
    const int64_t total_regs = b->reg_off[G];
    for (int gi = 0; gi < G; ++gi) {
        const int r0 = b->reg_off[gi], M = b->reg_off[gi + 1] - r0 - 1;
        auto bad = [&](const char* m) { delete P; return fail(MEEMB200_EINVAL, std::string("geometry ") + std::to_string(gi) + ": " + m); };
        if (M < 1 || M > kMaxRegions) return bad("number of interior regions must be in [1, 64]");
        const double h = b->h[gi];
        if (!(h > 0.0) || std::isinf(h)) return bad("h must be positive and finite");
        double prev = 0.0;
        for (int i = 0; i < M; ++i) {
            const double ai = b->a[r0 + i], di = b->d[r0 + i];
            if (!(ai > prev) || std::isinf(ai)) return bad("radii must be positive, finite and strictly increasing (basic_region_geometry.py:29-30)");
            if (!(di >= 0.0) || !(di < h)) return bad("drafts must satisfy 0 <= d < h");
            prev = ai;
        }
        if (b->d[r0 + M] != 0.0) return bad("the exterior draft d[M] must be 0 (basic_region_geometry.py:154-163)");
        for (int i = 0; i <= M; ++i)
            if (b->nmk[r0 + i] < 1) return bad("NMK entries must be >= 1");
        const int R = b->n_modes[gi];
        if (R < 1) return bad("n_modes must be >= 1");
        if (b->flag_off[gi + 1] - b->flag_off[gi] != (int64_t)R * (M + 1)) return bad("heaving flags must hold n_modes*(M+1) bytes");
        GeomDesc& gd = P->geoms[gi];
        Geom g{M, h, b->a + r0, b->d + r0, b->nmk + r0};
        Layout L;
        make_layout(g, L);
        gd.M = M; gd.R = R; gd.K = L.K; gd.N = b->nmk[r0 + M - 1];
        gd.NT = L.NT; gd.p = L.p; gd.reg_off = r0; gd.h = h;
        gd.flag_off = b->flag_off[gi];
        uint64_t hbits;
        memcpy(&hbits, &h, 8);
        auto ins = set_index.emplace(std::make_pair(hbits, L.K), (int)sets.size());       // roots depend on (h, K) only
        if (ins.second) { RootSet rs; rs.h = h; rs.K = L.K; rs.pad = 0; rs.off = 0; sets.push_back(rs); }
        gd.root_set = ins.first->second;
        gd.tab_off = tab_tot; tab_tot += table_doubles(L);
        gd.gs_off = gs_tot; gs_tot += gs_doubles(gd.N, R, gd.K);
        if (b->want_x && schur) { gd.xg_off = xg_tot; xg_tot += (int64_t)gd.p * (gd.N + R); }
        else gd.xg_off = -1;
        gd.out_rr = n_rr; n_rr += (int64_t)F * R * R;
        gd.out_r = n_r; n_r += (int64_t)F * R;
        gd.out_sys = (int64_t)gi * F;
        gd.out_x = n_x; if (b->want_x) n_x += (int64_t)F * R * gd.NT;
        gd.out_c = n_c; n_c += (int64_t)R * gd.p;
        slot_geom = std::max<int64_t>(slot_geom, (int64_t)gd.p * (gd.p + gd.N + R));
        {
            BandDims D;
            band_dims(g, L, R, D);
            band_wrows = std::max(band_wrows, D.wrows); band_wcols = std::max(band_wcols, D.wcols); band_side = std::max(band_side, D.side);
            Coop2Dims D2;
            coop2_dims(g, L, R, D2);
            coop2_wrows = std::max(coop2_wrows, D2.wrows); coop2_wcols = std::max(coop2_wcols, D2.wcols);
            for (int i = 0; i + 1 < M; ++i) all_case1 = all_case1 && (b->d[r0 + i] > b->d[r0 + i + 1]);
        }
        slot_freq = std::max<int64_t>(slot_freq, std::max(freq_slot_doubles(gd.N, gd.K, R), freq2_slot_doubles(gd.N, gd.K, R)));
        slot_dense = std::max<int64_t>(slot_dense, 2LL * gd.NT * (gd.NT + R));
        max_p = std::max(max_p, gd.p); max_K = std::max(max_K, gd.K); max_NT = std::max(max_NT, gd.NT); max_N = std::max(max_N, gd.N);
        const double p_ = gd.p, N_ = gd.N, K_ = gd.K, NT_ = gd.NT, R_ = R;
        if (schur) {
            cost[gi] = 2.0 / 3.0 * p_ * p_ * p_ + 2.0 * p_ * p_ * (N_ + R_);
            // frequency stage: T = H Imk (2N^2K) + lower triangle of Imk^T T (N K^2) + Cholesky (K^3/3) + solves, or
            // for the pivoted-LU variant the full second GEMM (2NK^2) and LU (2K^3/3)
            const double per_f = b->algo == MEEMB200_ALGO_SCHUR
                ? 2.0 * N_ * N_ * K_ + N_ * K_ * K_ + 1.0 / 3.0 * K_ * K_ * K_ + 2.0 * K_ * K_ * (1 + 2 * R_) + 4.0 * N_ * K_ * R_
                : 2.0 * N_ * N_ * K_ + 2.0 * N_ * K_ * K_ + 2.0 / 3.0 * K_ * K_ * K_ + 2.0 * K_ * K_ * (1 + 2 * R_) + 4.0 * N_ * K_ * R_;
            flops += cost[gi] + per_f * F;
            cost[gi] += per_f * F;
        } else {
            cost[gi] = 8.0 / 3.0 * NT_ * NT_ * NT_ + 8.0 * NT_ * NT_ * R_;
            flops += cost[gi] * F;
        }
    }
    (void)total_regs;
    // cost-sorted order (largest first) for the work queues
    std::vector<int32_t> order(G);
    for (int i = 0; i < G; ++i) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return cost[x] > cost[y]; });
    int64_t roots_tot = 0;
    P->max_FK = 1;
    for (auto& rs : sets) { rs.off = roots_tot; roots_tot += (int64_t)F * (1 + 2 * rs.K); P->max_FK = std::max<int64_t>(P->max_FK, (int64_t)F * rs.K); }
    // shared-memory driven panel widths
    const size_t limit = ctx->smem_optin;
    P->nb_geom = pick_nb<double>(max_p, limit, 4096);
    P->nb_freq = pick_nb<double>(max_K, limit / 2 > 100 * 1024 ? limit / 2 - 8192 : limit, 24 * 1024);
    P->nb_dense = pick_nb<cplx>(max_NT, limit, 4096);
    P->algo = b->algo;
    int max_R = 1;
    for (int gi = 0; gi < G; ++gi) max_R = std::max(max_R, P->geoms[gi].R);
    // k_freq_solve variant: P resident in shared memory when it fits (2 KB reserved for the kernel's static shared memory)
    P->smem_chol = freq2_solve_smem_bytes(max_K, max_R, max_N, b->want_x, true);
    // Resident by default only while >= 5 CTAs still fit an SM (K <~ 75): the factorisation is a latency chain (13 panels
    // at K = 200, each with a 16-step register Cholesky) that is hidden by CTAs in flight, not by threads per CTA -
    // measured on big_tricylinder: L2-resident 5 CTAs/SM 2.9 ms, shared-memory-resident 1 CTA/SM x 512 threads 3.0 ms.
    P->max_N = max_N;
    P->smem_hchol = std::min<size_t>(hchol_smem_bytes(max_N), limit - 1024);
    if (getenv("MEEMB200_NO_HCHOL")) P->smem_hchol = 0;          // tuning / test override: frequency stage uses H itself
    P->no_side = getenv("MEEMB200_NO_ROOTS_OVERLAP") != nullptr;
    P->gemm_flags = (getenv("MEEMB200_NO_SYRK") ? 1 : 0) | (getenv("MEEMB200_NO_LL") ? 2 : 0) | (getenv("MEEMB200_NO_BULK") ? 4 : 0) |
                    (getenv("MEEMB200_NO_STREAM") ? 8 : 0) | (getenv("MEEMB200_NO_STREAM_SYRK") ? 16 : 0) |
                    (getenv("MEEMB200_NO_BATCH_INV") ? 32 : 0);
    P->use_small = b->algo == MEEMB200_ALGO_SCHUR && max_K <= kSmallMax && max_N <= kSmallMax && max_R <= kSmallMaxR && !b->want_x &&
                   !getenv("MEEMB200_NO_SMALL");
    const bool fits = P->smem_chol + 2048 <= limit;
    P->solve_variant = (fits && P->smem_chol <= 44 * 1024) ? 1 : 0;
    if (const char* ev = getenv("MEEMB200_SOLVE_VARIANT")) {      // tuning / test override
        const int want = atoi(ev);
        if (want == 0 || (fits && want >= 1 && want <= 3)) P->solve_variant = want;
    }
    if (!P->solve_variant) P->smem_chol = freq2_solve_smem_bytes(max_K, max_R, max_N, b->want_x, false);
    if (schur && P->smem_chol > limit) { delete P; return fail(MEEMB200_EINVAL, "exterior mode count too large for the shared-memory Cholesky panel"); }
    if ((schur && (!P->nb_geom || !P->nb_freq)) || (b->algo == MEEMB200_ALGO_DENSE && !P->nb_dense)) {
        delete P;
        return fail(MEEMB200_EINVAL, "system too large for the shared-memory LU panels");
    }
    {   // cooperative geometry stage when it beats one CTA per geometry.  Cost model from measured timings (ms; v16,
        // tools/experiments/size_scan.py): one CTA per geometry takes ~1.2e-5 p^2 per wave by window elimination (~6e-8 p^3 by
        // the dense LU of the X path); the cooperative kernels process the geometries one after another at ~0.08 + 7.8e-4 p each
        // (static-pivot windows) or ~1.5e-3 p each (dense LU).
        const double waves = std::ceil((double)G / ctx->sm_count);
        const double pp = (double)max_p;
        P->smem_coop = coop_smem_bytes(max_p);
        P->coop2_wrows = coop2_wrows; P->coop2_wcols = coop2_wcols;
        P->smem_coop2 = coop2_smem_bytes(coop2_wrows, max_N);
        const bool coop2_ok = schur && all_case1 && !b->want_x && P->smem_coop2 + 8192 <= limit && !getenv("MEEMB200_NO_COOP2");
        const bool band_ok = schur && !b->want_x && !getenv("MEEMB200_NO_BAND");
        const double t_one = band_ok ? 1.2e-5 * pp * pp : 6e-8 * pp * pp * pp;
        const double t_coop = coop2_ok ? 0.08 + 7.8e-4 * pp : 1.5e-3 * pp;
        bool want_coop = schur && (double)G * t_coop < waves * t_one;
        if (getenv("MEEMB200_NO_COOP")) want_coop = false;
        if (getenv("MEEMB200_FORCE_COOP")) want_coop = schur;                                      // tests
        P->use_coop = want_coop && P->smem_coop + 8192 <= limit;
        if (P->use_coop) slot_geom = 0;
        if (P->use_coop)
            for (int gi = 0; gi < G; ++gi) slot_geom = std::max<int64_t>(slot_geom, coop_slot_doubles(P->geoms[gi].p, P->geoms[gi].N, P->geoms[gi].R));
        // static-pivot window variant: depths decreasing outwards everywhere (the C1 blocks then sit on diagonal pivot blocks), no X
        P->use_coop2 = want_coop && coop2_ok;
        if (P->use_coop2) { P->use_coop = true; slot_geom = coop2_slot_doubles(coop2_wrows, coop2_wcols); }
    }
    P->band_static = !getenv("MEEMB200_NO_BAND_STATIC");
    P->band_variant = -1; P->band_wrows = band_wrows; P->band_wcols = band_wcols; P->band_grid = 0; P->smem_band = 0;
    if (schur && !b->want_x && !P->use_coop && !getenv("MEEMB200_NO_BAND")) {
        // window elimination instead of the dense p x p LU (meem_geom_band.cuh); X needs all of G and keeps the dense form
        const size_t win_s = 8 * ((size_t)band_wrows * band_ld(band_wcols, true) + band_side);
        // small variant: the hand-over buffer aliases the LU scratch when it fits there (6 instead of 5 CTAs per SM on the C4 sweep)
        const size_t lu0 = align_up(lu_smem_bytes<double, 8, kBandSmallChunk>(band_wrows), 16);
        const bool alias0 = 8 * (size_t)band_side + 64 <= lu0;
        const size_t s0 = lu0 + (alias0 ? 8 * (size_t)band_wrows * band_ld(band_wcols, true) : win_s);
        P->band_side_alias = 0;
        const size_t s1 = align_up(lu_smem_bytes<double, 8, kLuChunk>(band_wrows), 16) + win_s;
        const size_t s2 = align_up(lu_smem_bytes<double, 8, kLuChunk>(band_wrows), 16);
        int want = -1;
        if (const char* ev = getenv("MEEMB200_BAND_VARIANT")) want = atoi(ev);     // tuning / test override
        if ((want < 0 || want == 0) && s0 + 2048 <= limit / 4) { P->band_variant = 0; P->smem_band = s0; P->band_side_alias = alias0 ? 1 : 0; }
        else if ((want < 0 || want == 1 || want == 0) && s1 + 2048 <= limit) { P->band_variant = 1; P->smem_band = s1; }
        else if (s2 + 2048 <= limit) { P->band_variant = 2; P->smem_band = s2; }
        if (P->band_variant >= 0) slot_geom = P->band_variant == 2 ? (int64_t)band_wrows * band_ld(band_wcols, false) + band_side : 0;
    }
    P->smem_geom = P->nb_geom ? nb_bytes<double>(P->nb_geom, max_p) : 0;
    P->smem_freq = P->nb_freq ? nb_bytes<double>(P->nb_freq, max_K) : 0;
    P->smem_dense = P->nb_dense ? nb_bytes<cplx>(P->nb_dense, max_NT) : 0;
    const int sms = ctx->sm_count;
    const int64_t n_sys = (int64_t)G * F;
    // One CTA per geometry is a latency chain (pivot search, row swaps and panel staging through L2): small geometries run
    // 256-thread CTAs, several per SM (register-limited), large ones 512 threads
    P->geom_threads = max_p <= 320 ? 256 : 512;
    if (const char* ev = getenv("MEEMB200_GEOM_THREADS")) { const int t = atoi(ev); if (t == 128 || t == 256 || t == 512) P->geom_threads = t; }
    P->n_slots_geom = P->use_coop ? 1 : (int)std::min<int64_t>(G, (int64_t)sms * (P->geom_threads == 512 ? 1 : 1024 / P->geom_threads));
    if (P->band_variant >= 0) P->n_slots_geom = (int)std::min<int64_t>(G, sms);
    if (schur) {
        const int per_sm = (std::max(P->smem_freq + 24 * 1024, kGemmSmemBytes + 4096) <= limit / 2) ? 2 : 1;
        P->n_slots_work = (int)std::min<int64_t>(n_sys, (int64_t)sms * per_sm);
    } else {
        P->n_slots_work = (int)std::min<int64_t>(n_sys, sms);
    }
    // blob
    size_t off = align_up(sizeof(BlobHeader), 16);
    BlobHeader H;
    memset(&H, 0, sizeof(H));
    H.n_geom = G; H.n_freq = F; H.n_sets = (int)sets.size(); H.freq_kind = b->freq_kind; H.root_mode = b->root_mode;
    H.algo = b->algo; H.want_x = b->want_x ? 1 : 0; H.rho = b->rho;
    const int64_t nreg = b->reg_off[G];
    const int64_t nflag = b->flag_off[G];
    H.off_geom = off; off = align_up(off + sizeof(GeomDesc) * G, 16);
    H.off_sets = off; off = align_up(off + sizeof(RootSet) * sets.size(), 16);
    H.off_a = off; off = align_up(off + 8 * nreg, 16);
    H.off_d = off; off = align_up(off + 8 * nreg, 16);
    H.off_nmk = off; off = align_up(off + 4 * nreg, 16);
    H.off_flags = off; off = align_up(off + nflag, 16);
    H.off_freq = off; off = align_up(off + 8 * (size_t)F, 16);
    H.off_order = off; off = align_up(off + 4 * (size_t)G, 16);
    const size_t blob_bytes = off;
    // workspace
    size_t w = 0;
    // frequency stage waves: systems per wave bounded by a ~1.5 GB system-buffer budget
    int64_t sys_stride = 8;
    for (int gi = 0; gi < G; ++gi) sys_stride = std::max<int64_t>(sys_stride, freq2_sys_doubles(P->geoms[gi].K, P->geoms[gi].R));
    sys_stride = (sys_stride + 7) & ~7LL;
    P->wave_items = (int64_t)std::max<int64_t>(4 * P->n_slots_work, (int64_t)(1.5e9 / (8.0 * sys_stride)));
    if (const char* ev = getenv("MEEMB200_WAVE_ITEMS")) {          // tuning / test override: systems per frequency-stage wave
        const long long n = atoll(ev);
        if (n > 0) P->wave_items = n;
    }
    P->wave_items = std::min<int64_t>(P->wave_items, n_sys);
    P->n_waves = b->algo == MEEMB200_ALGO_SCHUR ? (int)((n_sys + P->wave_items - 1) / P->wave_items) : 0;
    H.sys_stride = sys_stride;
    H.ws_counters = w; w = align_up(w + 256 + 8 * (size_t)P->n_waves, 256);
    H.ws_roots = w; w = align_up(w + 8 * (size_t)roots_tot, 256);
    H.ws_tables = w; w = align_up(w + 8 * (size_t)tab_tot, 256);
    H.ws_gstore = w; w = align_up(w + 8 * (size_t)gs_tot, 256);
    H.ws_xstore = w; w = align_up(w + 8 * (size_t)xg_tot, 256);
    H.slot_doubles_geom = (slot_geom + 31) & ~31LL;
    H.slot_doubles_freq = (slot_freq + 31) & ~31LL;
    H.slot_doubles_dense = (slot_dense + 31) & ~31LL;
    size_t slots_bytes;
    if (schur)
        slots_bytes = 8 * (size_t)std::max<int64_t>(H.slot_doubles_geom * P->n_slots_geom, H.slot_doubles_freq * P->n_slots_work);
    else
        slots_bytes = 8 * (size_t)(H.slot_doubles_dense * P->n_slots_work);
    H.ws_slots = w; w = align_up(w + slots_bytes, 256);
    size_t sysbuf_bytes = b->algo == MEEMB200_ALGO_SCHUR ? 8 * (size_t)(sys_stride * P->wave_items) : 0;
    if (P->use_small) sysbuf_bytes = std::max(sysbuf_bytes, 8 * (size_t)kSysRec * (size_t)n_sys);      // k_sys_pre records
    H.ws_sysbuf = w; w = align_up(w + sysbuf_bytes, 256);
    P->blob.assign(blob_bytes, 0);
    unsigned char* bp = P->blob.data();
    memcpy(bp, &H, sizeof(H));
    memcpy(bp + H.off_geom, P->geoms.data(), sizeof(GeomDesc) * G);
    memcpy(bp + H.off_sets, sets.data(), sizeof(RootSet) * sets.size());
    memcpy(bp + H.off_a, b->a, 8 * nreg);
    memcpy(bp + H.off_d, b->d, 8 * nreg);
    memcpy(bp + H.off_nmk, b->nmk, 4 * nreg);
    memcpy(bp + H.off_flags, b->heaving, nflag);
    memcpy(bp + H.off_freq, b->freq, 8 * (size_t)F);
    memcpy(bp + H.off_order, order.data(), 4 * (size_t)G);
    // a[M] of each geometry is documented as inf; it is never read
    P->info.input_bytes = blob_bytes;
    P->info.workspace_bytes = w;
    P->info.n_rr = n_rr; P->info.n_r = n_r; P->info.n_sys = n_sys; P->info.n_x = n_x; P->info.n_c = n_c;
    P->info.geom_kernel = !schur ? -1 : P->use_coop2 ? 5 : P->use_coop ? 1 : P->band_variant >= 0 ? 2 + P->band_variant : 0;
    P->info.freq_kernel = !schur ? 3 : b->algo == MEEMB200_ALGO_SCHUR_LU ? 2 : P->use_small ? 1 : 0;
    P->info.n_launches = b->algo == MEEMB200_ALGO_SCHUR ? (P->use_small ? 8 : 6 + 2 * P->n_waves) : b->algo == MEEMB200_ALGO_SCHUR_LU ? 5 : 4;
    P->solve_grid = 0;
    P->info.flops = flops;
    *out = P;
    return MEEMB200_OK;
}

extern "C" int meemb200_plan_destroy(meemb200_plan* p) {
    if (p && p->ev_ok) for (auto& e : p->ev) cudaEventDestroy(e);
    if (p && p->ev_fork) { cudaEventDestroy(p->ev_fork); cudaEventDestroy(p->ev_join); }
    if (p && p->host_call_base) { DeviceGuard guard(p->ctx->device); cudaFree(p->host_call_base); }
    delete p;
    return MEEMB200_OK;
}
extern "C" int meemb200_plan_info(const meemb200_plan* p, MeemPlanInfo* info) {
    if (!p || !info) return fail(MEEMB200_EINVAL, "NULL argument");
    *info = p->info;
    return MEEMB200_OK;
}
extern "C" const void* meemb200_plan_blob(const meemb200_plan* p) { return p ? p->blob.data() : nullptr; }
extern "C" int meemb200_plan_offsets(const meemb200_plan* p, int32_t g, int64_t out[8]) {
    if (!p || !out || g < 0 || g >= (int)p->geoms.size()) return fail(MEEMB200_EINVAL, "bad geometry index");
    const GeomDesc& gd = p->geoms[g];
    out[0] = gd.out_rr; out[1] = gd.out_r; out[2] = gd.out_sys; out[3] = gd.out_x; out[4] = gd.NT; out[5] = gd.p;
    out[6] = gd.out_c; out[7] = gd.R;
    return MEEMB200_OK;
}
// Test support (no reference counterpart): every n-th system of the next solves is reported as a failed Cholesky, which sends
// it down the pivoted-LU re-solve (k_freq_lu) - the only way to reach that path with physical inputs.  n = 0 switches it off.
extern "C" int meemb200_plan_set_test_hook(meemb200_plan* p, int32_t flag_every) {
    if (!p || flag_every < 0 || flag_every >= (1 << 20)) return fail(MEEMB200_EINVAL, "bad argument");
    p->gemm_flags = (p->gemm_flags & 0xffff) | (flag_every << 16);
    return MEEMB200_OK;
}
extern "C" int meemb200_plan_set_freq(meemb200_plan* p, const double* freq) {
    if (!p || !freq) return fail(MEEMB200_EINVAL, "NULL argument");
    const BlobHeader* h = p->hdr();
    for (int f = 0; f < h->n_freq; ++f) {
        if (!(freq[f] > 0.0)) return fail(MEEMB200_EINVAL, "frequencies must be positive (meem_problem.py:31)");
        if (std::isinf(freq[f]) && h->freq_kind == MEEMB200_FREQ_OMEGA) return fail(MEEMB200_EINVAL, "omega must be finite; pass m0=inf with FREQ_M0");
    }
    memcpy(p->blob.data() + h->off_freq, freq, 8 * (size_t)h->n_freq);
    return MEEMB200_OK;
}

static cudaError_t launch_geom_coop2(meemb200_plan* P, const void* in, void* ws, cudaStream_t s) {
    cudaError_t e = cudaFuncSetAttribute(k_geom_coop2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P->smem_coop2);
    if (e != cudaSuccess) return e;
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_geom_coop2, kCoopThreads, P->smem_coop2);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorLaunchOutOfResources;
    void* args[] = {(void*)&in, (void*)&ws, (void*)&P->coop2_wrows, (void*)&P->coop2_wcols};
    return cudaLaunchCooperativeKernel((void*)k_geom_coop2, dim3(P->ctx->sm_count), dim3(kCoopThreads), args, P->smem_coop2, s);
}
static cudaError_t launch_geom_coop(meemb200_plan* P, const void* in, void* ws, cudaStream_t s) {
    if (P->use_coop2) return launch_geom_coop2(P, in, ws, s);
    cudaError_t e = cudaFuncSetAttribute(k_geom_coop, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P->smem_coop);
    if (e != cudaSuccess) return e;
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_geom_coop, kCoopThreads, P->smem_coop);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorLaunchOutOfResources;
    void* args[] = {(void*)&in, (void*)&ws};
    return cudaLaunchCooperativeKernel((void*)k_geom_coop, dim3(P->ctx->sm_count), dim3(kCoopThreads), args, P->smem_coop, s);
}
static cudaError_t launch_geom_band(meemb200_plan* P, const void* in, void* ws, cudaStream_t s) {
    typedef void (*band_fn)(const void*, void*, int, int, int, int);
    static const band_fn fns[3] = {k_geom_band<8, 128, 6, true, true, kBandSmallChunk>, k_geom_band<8, 256, 2, false, true, kLuChunk>,
                                   k_geom_band<8, 512, 1, false, false, kLuChunk>};
    static const int thr[3] = {128, 256, 512};
    const band_fn kfn = fns[P->band_variant];
    const int nthr = thr[P->band_variant];
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P->smem_band);
    if (e != cudaSuccess) return e;
    if (!P->band_grid) {
        int per_sm = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kfn, nthr, P->smem_band);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) return cudaErrorLaunchOutOfResources;
        P->band_grid = P->band_variant == 2 ? P->n_slots_geom : (int)std::min<int64_t>(P->hdr()->n_geom, (int64_t)per_sm * P->ctx->sm_count);
    }
    kfn<<<P->band_grid, nthr, P->smem_band, s>>>(in, ws, P->band_wrows, P->band_wcols, P->band_side_alias, P->band_static ? 1 : 0);
    return cudaGetLastError();
}
template <int NB> static cudaError_t launch_geom(meemb200_plan* P, const void* in, void* ws, cudaStream_t s) {
    if (P->use_coop) return launch_geom_coop(P, in, ws, s);
    if (P->band_variant >= 0) return launch_geom_band(P, in, ws, s);
    auto kfn = P->geom_threads == 512 ? k_geom<NB, 512, 1> : P->geom_threads == 256 ? k_geom<NB, 256, 4> : k_geom<NB, 128, 8>;
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P->smem_geom);
    if (e != cudaSuccess) return e;
    kfn<<<P->n_slots_geom, P->geom_threads, P->smem_geom, s>>>(in, ws, 0);
    return cudaGetLastError();
}
template <int NB> static cudaError_t launch_freq(meemb200_plan* P, const void* in, void* ws, OutPtrs o, cudaStream_t s) {
    cudaError_t e = cudaFuncSetAttribute(k_freq_lu<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P->smem_freq);
    if (e != cudaSuccess) return e;
    if (P->algo == MEEMB200_ALGO_SCHUR && P->use_small) {
        const size_t smem_small = sizeof(SmallWs) * kSmallWarps;
        e = cudaFuncSetAttribute(k_freq_small, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_small);
        if (e != cudaSuccess) return e;
        if (!P->solve_grid) {
            int per_sm = 0;
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_freq_small, 32 * kSmallWarps, smem_small);
            if (e != cudaSuccess) return e;
            if (per_sm < 1) return cudaErrorInvalidConfiguration;
            P->solve_grid = per_sm * P->ctx->sm_count;
        }
        const BlobHeader* h = P->hdr();
        int* wave_counters = reinterpret_cast<int*>(static_cast<char*>(ws) + h->ws_counters + 256);
        const int64_t n_sys = P->info.n_sys;
        const int grid = (int)std::min<int64_t>((n_sys + kSmallWarps - 1) / kSmallWarps, P->solve_grid);
        k_sys_pre<<<(unsigned)((n_sys + 127) / 128), 128, 0, s>>>(in, ws, (int)n_sys);
        k_freq_small<<<grid, 32 * kSmallWarps, smem_small, s>>>(in, ws, o, (int)n_sys, wave_counters, P->gemm_flags >> 16);
        cudaEventRecord(P->ev[3], s);
        cudaEventRecord(P->ev[4], s);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    } else if (P->algo == MEEMB200_ALGO_SCHUR) {
        e = cudaFuncSetAttribute(k_freq_gemm, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGemmSmemBytes);
        if (e != cudaSuccess) return e;
        typedef void (*solve_fn)(const void*, void*, OutPtrs, int, int, int*, int);
        static const solve_fn solve_k[4] = {k_freq_solve<128, 4, false>, k_freq_solve<128, 5, true>, k_freq_solve<256, 2, true>,
                                            k_freq_solve<512, 1, true>};
        static const int solve_thr[4] = {128, 128, 256, 512};
        const solve_fn ksolve = solve_k[P->solve_variant];
        const int nthr_solve = solve_thr[P->solve_variant];
        e = cudaFuncSetAttribute(ksolve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P->smem_chol);
        if (e != cudaSuccess) return e;
        if (!P->solve_grid) {
            int per_sm = 0;
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ksolve, nthr_solve, P->smem_chol);
            if (e != cudaSuccess) return e;
            if (per_sm < 1) return cudaErrorInvalidConfiguration;
            P->solve_grid = per_sm * P->ctx->sm_count;
        }
        const BlobHeader* h = P->hdr();
        int* wave_counters = reinterpret_cast<int*>(static_cast<char*>(ws) + h->ws_counters + 256);
        const int64_t n_sys = P->info.n_sys;
        for (int w = 0; w < P->n_waves; ++w) {
            const int item0 = (int)(w * P->wave_items), item1 = (int)std::min<int64_t>(n_sys, (w + 1) * P->wave_items);
            const int n = item1 - item0;
            k_freq_gemm<<<std::min(n, P->n_slots_work), 256, kGemmSmemBytes, s>>>(in, ws, item0, item1, wave_counters + 2 * w, P->gemm_flags);
            if (w == P->n_waves - 1) cudaEventRecord(P->ev[3], s);
            ksolve<<<std::min(n, P->solve_grid), nthr_solve, P->smem_chol, s>>>(in, ws, o, item0, item1, wave_counters + 2 * w + 1, P->gemm_flags);
            if (w == P->n_waves - 1) cudaEventRecord(P->ev[4], s);
        }
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    k_freq_lu<NB><<<P->n_slots_work, 256, P->smem_freq, s>>>(in, ws, o, P->algo == MEEMB200_ALGO_SCHUR ? 1 : 0);
    return cudaGetLastError();
}
template <int NB> static cudaError_t launch_dense(meemb200_plan* P, const void* in, void* ws, OutPtrs o, cudaStream_t s) {
    cudaError_t e = cudaFuncSetAttribute(k_dense<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P->smem_dense);
    if (e != cudaSuccess) return e;
    k_dense<NB><<<P->n_slots_work, 512, P->smem_dense, s>>>(in, ws, o);
    return cudaGetLastError();
}

static int enqueue_prologue(meemb200_plan* P, const void* d_in, void* d_ws, cudaStream_t s, cudaStream_t sr = nullptr) {
    const BlobHeader* h = P->hdr();
    if (!sr) sr = s;                                       // sr: the stream of the wavenumber kernels
    CUDA_TRY(cudaMemsetAsync(static_cast<char*>(d_ws) + h->ws_counters, 0, 256 + 8 * (size_t)P->n_waves, s));
    const int nt = h->n_sets * h->n_freq;
    k_m0<<<(nt + 127) / 128, 128, 0, sr>>>(d_in, d_ws);
    CUDA_TRY(cudaGetLastError());
    dim3 grid((unsigned)((P->max_FK + 127) / 128), (unsigned)std::min(h->n_sets, 65535));
    k_mk<<<grid, 128, 0, sr>>>(d_in, d_ws);
    CUDA_TRY(cudaGetLastError());
    return MEEMB200_OK;
}

// side_ok: the wavenumber kernels (k_m0, k_mk) may run on the plan's side stream, forked from and joined back into `stream`
// with events (graph-capture safe), concurrently with the radial tables and the geometry stage, which do not read the roots -
// big_tricylinder: 0.07 ms of roots hidden behind k_geom_coop2 / k_hchol.  meemb200_solve_static_ws passes false: its
// contract is to create nothing but the six timing events.
static int solve_impl(meemb200_plan* P, const void* d_in, void* d_ws, size_t ws_bytes, const MeemOutputs* out, void* stream, bool side_ok) {
    if (!P || !d_in || !d_ws || !out) return fail(MEEMB200_EINVAL, "NULL argument");
    if (ws_bytes < P->info.workspace_bytes) return fail(MEEMB200_ENOMEM, "workspace too small");
    if (!out->added_mass || !out->damping || !out->exc_force || !out->exc_phase || !out->m0 || !out->status)
        return fail(MEEMB200_EINVAL, "NULL output pointer");
    const BlobHeader* h = P->hdr();
    if (h->want_x && !out->x) return fail(MEEMB200_EINVAL, "plan was built with want_x but outputs.x is NULL");
    DeviceGuard guard(P->ctx->device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (!P->ev_ok) {
        for (auto& e : P->ev) CUDA_TRY(cudaEventCreate(&e));
        P->ev_ok = true;
    }
    OutPtrs o{out->added_mass, out->damping, out->exc_force, out->exc_phase, out->m0, out->status, h->want_x ? out->x : nullptr};
    CUDA_TRY(cudaEventRecord(P->ev[0], s));
    const bool side = side_ok && !P->no_side;
    if (side && !P->side) {
        {
            std::lock_guard<std::mutex> lk(P->ctx->side_mu);
            if (!P->ctx->side) CUDA_TRY(cudaStreamCreateWithFlags(&P->ctx->side, cudaStreamNonBlocking));
        }
        P->side = P->ctx->side;
        CUDA_TRY(cudaEventCreateWithFlags(&P->ev_fork, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&P->ev_join, cudaEventDisableTiming));
    }
    if (side) {     // the fork event follows every earlier reader of the roots on `stream` (the previous solve's frequency stage)
        CUDA_TRY(cudaEventRecord(P->ev_fork, s));
        CUDA_TRY(cudaStreamWaitEvent(P->side, P->ev_fork, 0));
    }
    int rc = enqueue_prologue(P, d_in, d_ws, s, side ? P->side : s);
    if (rc) return rc;
    if (side) CUDA_TRY(cudaEventRecord(P->ev_join, P->side));
    CUDA_TRY(cudaEventRecord(P->ev[1], s));
    k_tables<<<h->n_geom, 128, 0, s>>>(d_in, d_ws);
    CUDA_TRY(cudaGetLastError());
    if (out->c_vector) {
        k_cvec<<<h->n_geom, 128, 0, s>>>(d_in, d_ws, out->c_vector);
        CUDA_TRY(cudaGetLastError());
    }
    if (h->algo != MEEMB200_ALGO_DENSE) {
        CUDA_TRY(P->nb_geom == 8 ? launch_geom<8>(P, d_in, d_ws, s) : P->nb_geom == 4 ? launch_geom<4>(P, d_in, d_ws, s) : launch_geom<2>(P, d_in, d_ws, s));
        if (h->algo == MEEMB200_ALGO_SCHUR) {
            CUDA_TRY(cudaFuncSetAttribute(k_hchol, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(P->smem_hchol, 1024)));
            const int hthreads = P->max_N > 64 ? kHcholThreads : 128;
            k_hchol<<<std::min(h->n_geom, 8 * P->ctx->sm_count), hthreads, std::max<size_t>(P->smem_hchol, 1024), s>>>(d_in, d_ws, (unsigned)P->smem_hchol);
            CUDA_TRY(cudaGetLastError());
        }
        if (side) CUDA_TRY(cudaStreamWaitEvent(s, P->ev_join, 0));
        CUDA_TRY(cudaEventRecord(P->ev[2], s));
        CUDA_TRY(P->nb_freq == 8 ? launch_freq<8>(P, d_in, d_ws, o, s) : P->nb_freq == 4 ? launch_freq<4>(P, d_in, d_ws, o, s) : launch_freq<2>(P, d_in, d_ws, o, s));
    } else {
        if (side) CUDA_TRY(cudaStreamWaitEvent(s, P->ev_join, 0));
        CUDA_TRY(cudaEventRecord(P->ev[2], s));
        CUDA_TRY(P->nb_dense == 8 ? launch_dense<8>(P, d_in, d_ws, o, s) : P->nb_dense == 4 ? launch_dense<4>(P, d_in, d_ws, o, s) : launch_dense<2>(P, d_in, d_ws, o, s));
    }
    if (h->algo != MEEMB200_ALGO_SCHUR) { CUDA_TRY(cudaEventRecord(P->ev[3], s)); CUDA_TRY(cudaEventRecord(P->ev[4], s)); }
    CUDA_TRY(cudaEventRecord(P->ev[5], s));
    return MEEMB200_OK;
}
extern "C" int meemb200_solve(meemb200_plan* P, const void* d_in, void* d_ws, size_t ws_bytes, const MeemOutputs* out, void* stream) {
    return solve_impl(P, d_in, d_ws, ws_bytes, out, stream, true);
}

extern "C" int meemb200_last_timing(meemb200_plan* P, float ms[5]) {
    if (!P || !ms || !P->ev_ok) return fail(MEEMB200_EINVAL, "no solve recorded");
    CUDA_TRY(cudaEventElapsedTime(&ms[0], P->ev[0], P->ev[1]));
    CUDA_TRY(cudaEventElapsedTime(&ms[1], P->ev[1], P->ev[2]));
    CUDA_TRY(cudaEventElapsedTime(&ms[2], P->ev[2], P->ev[5]));
    if (P->algo == MEEMB200_ALGO_SCHUR && P->n_waves == 1) {     // single wave: the two frequency kernels separately
        CUDA_TRY(cudaEventElapsedTime(&ms[3], P->ev[2], P->ev[3]));
        CUDA_TRY(cudaEventElapsedTime(&ms[4], P->ev[3], P->ev[4]));
    } else {
        ms[3] = ms[2]; ms[4] = 0.0f;
    }
    return MEEMB200_OK;
}

// Host-side view of two device layouts, for the CPU tests (pure integer functions, no device needed):
// kind 0: out[r] = tri_off(r) for r = 0 .. n, out[n + 1] = tri_doubles(n);  kind 1: SYRK warp-tile list for Nc = n columns:
// out[0] = number of tiles, then (a, b, first row block, row blocks, first column block, column blocks) per tile.
extern "C" int meemb200_debug_layout(int kind, int n, int* out, int out_len) {
    if (!out || n < 1) return fail(MEEMB200_EINVAL, "bad argument");
    if (kind == 0) {
        if (out_len < n + 2) return fail(MEEMB200_EINVAL, "output too short");
        for (int r = 0; r <= n; ++r) out[r] = tri_off(r);
        out[n + 1] = tri_doubles(n);
        return MEEMB200_OK;
    }
    if (kind == 1) {
        if (!syrk_fits(n)) return fail(MEEMB200_EINVAL, "SYRK tile list does not apply to this width");
        SyrkPlan plan;
        const int nb = (n + 7) >> 3, ng = (nb + 3) >> 2;
        syrk_build_plan(plan, nb);
        if (out_len < 1 + 6 * plan.nt) return fail(MEEMB200_EINVAL, "output too short");
        out[0] = plan.nt;
        for (int t = 0; t < plan.nt; ++t) {
            const int a = plan.ta[t], b = plan.tb[t];
            int* o = out + 1 + 6 * t;
            o[0] = a; o[1] = b; o[2] = a * nb / ng; o[3] = (a + 1) * nb / ng - o[2]; o[4] = b * nb / ng; o[5] = (b + 1) * nb / ng - o[4];
        }
        return MEEMB200_OK;
    }
    if (kind == 2) {      // chunk stream of the triangular product for N = K = n, Mw = n - 1 (one zero row of H)
        if (out_len < 8 + 8 * 10) return fail(MEEMB200_EINVAL, "output too short");
        const int nb = (n + 7) >> 3, Mw = n - 1, mbw = (Mw + 7) >> 3, ngr = (mbw + 3) >> 2;
        const bool fits = trmm_fits(n, n, 1);
        const int kc = (nb >> 3) >= 4 ? 8 : (nb >> 3) == 3 ? (trmm_stages<16>(n) >= 3 ? 16 : 8) : 16;
        out[0] = fits; out[1] = kc; out[2] = kc == 16 ? trmm_stages<16>(n) : trmm_stages<8>(n); out[3] = ngr;
        out[4] = fits ? trmm_total_chunks(n, Mw, kc) : 0; out[5] = nb; out[6] = (int)(lat_doubles(n, n, 1)); out[7] = syrk_lds(n);
        for (int a = 0; a < 8; ++a) {     // per group: NI, first row, blocks of the 8 warps
            int* o = out + 8 + 10 * a;
            const int NI = a < ngr ? std::min(4, mbw - 4 * a) : 0;
            o[0] = NI; o[1] = 32 * a;
            for (int w = 0; w < 8; ++w) o[2 + w] = NI ? trmm_warp_blocks(nb, NI, w) : 0;
        }
        return MEEMB200_OK;
    }
    if (kind == 3) {      // LAT (the factor re-tiled for the stream): tile offsets in doubles for N = n, tiles -1 .. (n + 31) / 32
        const int nt = (n + 31) / 32;
        if (out_len < nt + 2) return fail(MEEMB200_EINVAL, "output too short");
        for (int a = -1; a <= nt; ++a) out[a + 1] = (int)lat_tile_off(n, a);
        return MEEMB200_OK;
    }
    return fail(MEEMB200_EINVAL, "unknown kind");
}

extern "C" int meemb200_phase_profile(meemb200_plan* P, const void* d_ws, double cycles[16], void* stream) {
    if (!P || !d_ws || !cycles) return fail(MEEMB200_EINVAL, "NULL argument");
    unsigned long long raw[16];
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    CUDA_TRY(cudaMemcpyAsync(raw, static_cast<const char*>(d_ws) + P->hdr()->ws_counters + 64, sizeof(raw), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    for (int i = 0; i < 16; ++i) cycles[i] = (double)raw[i];
    return MEEMB200_OK;
}

extern "C" int meemb200_solve_host(meemb200_plan* P, const MeemOutputs* ho, void* stream) {
    if (!P || !ho) return fail(MEEMB200_EINVAL, "NULL argument");
    DeviceGuard guard(P->ctx->device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const MeemPlanInfo& I = P->info;
    const size_t b_rr = 8 * (size_t)I.n_rr, b_r = 8 * (size_t)I.n_r, b_m0 = 8 * (size_t)I.n_sys, b_st = 4 * (size_t)I.n_sys, b_x = 16 * (size_t)I.n_x;
    const size_t b_c = ho->c_vector ? 8 * (size_t)I.n_c : 0;
    if (P->hdr()->want_x && !ho->x) return fail(MEEMB200_EINVAL, "plan was built with want_x but outputs.x is NULL");
    size_t off = 0;
    auto take = [&](size_t n) { size_t o = off; off = align_up(off + n, 256); return o; };
    const size_t o_in = take(I.input_bytes), o_ws = take(I.workspace_bytes), o_am = take(b_rr), o_dp = take(b_rr), o_f = take(b_r),
                 o_p = take(b_r), o_m0 = take(b_m0), o_st = take(b_st), o_x = take(b_x), o_c = take(b_c);
    if (P->host_call_bytes < off) {          // first call (or a larger request): the allocation stays with the plan
        if (P->host_call_base) { CUDA_TRY(cudaFree(P->host_call_base)); P->host_call_base = nullptr; P->host_call_bytes = 0; }
        CUDA_TRY(cudaMalloc(reinterpret_cast<void**>(&P->host_call_base), off));
        P->host_call_bytes = off;
    }
    char* base = P->host_call_base;
    MeemOutputs d;
    d.added_mass = reinterpret_cast<double*>(base + o_am); d.damping = reinterpret_cast<double*>(base + o_dp);
    d.exc_force = reinterpret_cast<double*>(base + o_f); d.exc_phase = reinterpret_cast<double*>(base + o_p);
    d.m0 = reinterpret_cast<double*>(base + o_m0); d.status = reinterpret_cast<int32_t*>(base + o_st);
    d.x = I.n_x ? reinterpret_cast<double*>(base + o_x) : nullptr;
    d.c_vector = b_c ? reinterpret_cast<double*>(base + o_c) : nullptr;
    cudaError_t e = cudaMemcpyAsync(base + o_in, P->blob.data(), I.input_bytes, cudaMemcpyHostToDevice, s);
    int rc = MEEMB200_OK;
    if (e != cudaSuccess) rc = fail(MEEMB200_ECUDA, cudaGetErrorString(e));
    if (!rc) rc = meemb200_solve(P, base + o_in, base + o_ws, I.workspace_bytes, &d, s);
    if (!rc) {
        auto back = [&](void* dst, const void* src, size_t n) {
            if (!dst || !n || rc) return;
            const cudaError_t ce = cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, s);
            if (ce != cudaSuccess) rc = fail(MEEMB200_ECUDA, std::string("solve_host download: ") + cudaGetErrorString(ce));
        };
        back(ho->added_mass, d.added_mass, b_rr);
        back(ho->damping, d.damping, b_rr);
        back(ho->exc_force, d.exc_force, b_r);
        back(ho->exc_phase, d.exc_phase, b_r);
        back(ho->m0, d.m0, b_m0);
        back(ho->status, d.status, b_st);
        back(ho->x, d.x, I.n_x ? b_x : 0);
        back(ho->c_vector, d.c_vector, b_c);
    }
    e = cudaStreamSynchronize(s);
    if (!rc && e != cudaSuccess) rc = fail(MEEMB200_ECUDA, std::string("solve_host: ") + cudaGetErrorString(e));
    return rc;
}

// ---- allocation-free solve: the workspace (scratch of every kernel of the pipeline) lives in the library image.
// g_static_ws is a DEVICE symbol: in host code its name is only a handle for cudaGetSymbolAddress, never a device address.
// Each GPU that loads the library holds its own MEEMB200_STATIC_WS_BYTES copy until the process exits.
static __device__ __align__(256) unsigned char g_static_ws[MEEMB200_STATIC_WS_BYTES];
static const int kStaticWsMaxDevices = 64;
static void* g_static_ws_ptr[kStaticWsMaxDevices];       // device address of g_static_ws per device ordinal, resolved once
static std::mutex g_static_ws_mu;

// `device` must be the current device of the calling thread (the symbol resolves in the current context).
static int static_ws_address(int device, void** out) {
    if (device < 0 || device >= kStaticWsMaxDevices) return fail(MEEMB200_EINVAL, "device ordinal beyond the static-workspace table");
    std::lock_guard<std::mutex> lock(g_static_ws_mu);
    if (!g_static_ws_ptr[device]) {
        int cur = -1;
        CUDA_TRY(cudaGetDevice(&cur));
        if (cur != device) return fail(MEEMB200_ECUDA, "static workspace: the plan's device could not be made current");
        void* p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(&p, g_static_ws));
        if (!p || (reinterpret_cast<uintptr_t>(p) & 255) != 0) return fail(MEEMB200_ECUDA, "static workspace: bad symbol address");
        g_static_ws_ptr[device] = p;
    }
    *out = g_static_ws_ptr[device];
    return MEEMB200_OK;
}

extern "C" int meemb200_solve_static_ws(meemb200_plan* P, const void* d_in, const MeemOutputs* out, void* stream) {
    if (!P || !d_in || !out) return fail(MEEMB200_EINVAL, "NULL argument");
    if (P->info.workspace_bytes > (size_t)MEEMB200_STATIC_WS_BYTES)
        return fail(MEEMB200_ENOMEM, "plan needs " + std::to_string(P->info.workspace_bytes) + " bytes of workspace, the library's static workspace holds " +
                                         std::to_string((size_t)MEEMB200_STATIC_WS_BYTES) + " (MEEMB200_STATIC_WS_BYTES): use meemb200_solve with a caller-owned workspace");
    DeviceGuard guard(P->ctx->device);
    void* ws = nullptr;
    int rc = static_ws_address(P->ctx->device, &ws);
    if (rc) return rc;
    // Same kernels, same order as meemb200_solve; its prologue zeroes the work-queue counters, the only part of the
    // workspace that has to start at zero - everything else is written before it is read, whatever the last call left.
    return solve_impl(P, d_in, ws, (size_t)MEEMB200_STATIC_WS_BYTES, out, stream, false);
}

extern "C" int meemb200_assemble(meemb200_plan* P, const void* d_in, void* d_ws, size_t ws_bytes, int32_t g, int32_t f,
                                 double* d_A, double* d_b, double* d_c, void* stream) {
    if (!P || !d_in || !d_ws) return fail(MEEMB200_EINVAL, "NULL argument");
    if (ws_bytes < P->info.workspace_bytes) return fail(MEEMB200_ENOMEM, "workspace too small");
    const BlobHeader* h = P->hdr();
    if (g < 0 || g >= h->n_geom || f < 0 || f >= h->n_freq) return fail(MEEMB200_EINVAL, "geometry/frequency index out of range");
    DeviceGuard guard(P->ctx->device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int rc = enqueue_prologue(P, d_in, d_ws, s);
    if (rc) return rc;
    k_tables<<<h->n_geom, 128, 0, s>>>(d_in, d_ws);
    CUDA_TRY(cudaGetLastError());
    k_assemble<<<1, 512, 0, s>>>(d_in, d_ws, g, f, reinterpret_cast<cplx*>(d_A), reinterpret_cast<cplx*>(d_b), reinterpret_cast<cplx*>(d_c));
    CUDA_TRY(cudaGetLastError());
    return MEEMB200_OK;
}

extern "C" int meemb200_imk_table(meemb200_plan* P, const void* d_in, void* d_ws, size_t ws_bytes, int32_t g, int32_t f,
                                  double* d_out, void* stream) {
    if (!P || !d_in || !d_ws || !d_out) return fail(MEEMB200_EINVAL, "NULL argument");
    if (ws_bytes < P->info.workspace_bytes) return fail(MEEMB200_ENOMEM, "workspace too small");
    const BlobHeader* h = P->hdr();
    if (g < 0 || g >= h->n_geom || f < 0 || f >= h->n_freq) return fail(MEEMB200_EINVAL, "geometry/frequency index out of range");
    if (h->algo != MEEMB200_ALGO_SCHUR) return fail(MEEMB200_EINVAL, "the coupling-table export needs a plan of the default algorithm");
    const GeomDesc& gd = P->geoms[g];
    const size_t smem = sizeof(double) * (size_t)(6 * gd.K + 3 * gd.N);
    if (smem + 1024 > P->ctx->smem_optin || (int64_t)pad16(gd.N) * pad8(gd.K) > h->slot_doubles_freq)
        return fail(MEEMB200_EINVAL, "system too large for the coupling-table export");
    DeviceGuard guard(P->ctx->device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int rc = enqueue_prologue(P, d_in, d_ws, s);
    if (rc) return rc;
    CUDA_TRY(cudaFuncSetAttribute(k_imk_dump, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_imk_dump<<<1, 256, smem, s>>>(d_in, d_ws, g, f, d_out);
    CUDA_TRY(cudaGetLastError());
    return MEEMB200_OK;
}

extern "C" int meemb200_coefficients(meemb200_plan* P, const void* d_in, void* d_ws, size_t ws_bytes, int32_t g, int32_t f,
                                     const double* d_X, double* d_out, void* stream) {
    if (!P || !d_in || !d_ws || !d_X || !d_out) return fail(MEEMB200_EINVAL, "NULL argument");
    if (ws_bytes < P->info.workspace_bytes) return fail(MEEMB200_ENOMEM, "workspace too small");
    const BlobHeader* h = P->hdr();
    if (g < 0 || g >= h->n_geom || f < 0 || f >= h->n_freq) return fail(MEEMB200_EINVAL, "geometry/frequency index out of range");
    DeviceGuard guard(P->ctx->device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int rc = enqueue_prologue(P, d_in, d_ws, s);
    if (rc) return rc;
    k_tables<<<h->n_geom, 128, 0, s>>>(d_in, d_ws);
    CUDA_TRY(cudaGetLastError());
    k_reduce<<<1, 128, 0, s>>>(d_in, d_ws, g, f, reinterpret_cast<const cplx*>(d_X), d_out);
    CUDA_TRY(cudaGetLastError());
    return MEEMB200_OK;
}

extern "C" int meemb200_fields(meemb200_plan* P, const void* d_in, void* d_ws, size_t ws_bytes, int32_t g, int32_t f, int32_t mode,
                               const double* d_X, int64_t n_points, const double* d_r, const double* d_z, double* d_out, void* stream) {
    if (!P || !d_in || !d_ws || !d_X || !d_r || !d_z || !d_out) return fail(MEEMB200_EINVAL, "NULL argument");
    if (ws_bytes < P->info.workspace_bytes) return fail(MEEMB200_ENOMEM, "workspace too small");
    const BlobHeader* h = P->hdr();
    if (g < 0 || g >= h->n_geom || f < 0 || f >= h->n_freq) return fail(MEEMB200_EINVAL, "geometry/frequency index out of range");
    if (mode < 0 || mode >= P->geoms[g].R) return fail(MEEMB200_EINVAL, "mode index out of range");
    if (n_points < 1) return fail(MEEMB200_EINVAL, "n_points must be >= 1");
    const double fr = reinterpret_cast<const double*>(P->blob.data() + h->off_freq)[f];
    if (std::isinf(fr)) return fail(MEEMB200_EINVAL, "field evaluation needs a finite wavenumber");
    DeviceGuard guard(P->ctx->device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int rc = enqueue_prologue(P, d_in, d_ws, s);
    if (rc) return rc;
    const long warps = (long)n_points;
    const int blocks = (int)std::min<long>((warps + 7) / 8, (long)P->ctx->sm_count * 8);
    k_fields<<<blocks, 256, 0, s>>>(d_in, d_ws, g, f, mode, d_X, (long)n_points, d_r, d_z, d_out);
    CUDA_TRY(cudaGetLastError());
    return MEEMB200_OK;
}

extern "C" int meemb200_roots(meemb200_ctx* ctx, int32_t n, int32_t K, const double* d_freq, const double* d_h, int32_t freq_kind,
                              int32_t root_mode, double* d_m0, double* d_mk, double* d_Nk, void* stream) {
    if (!ctx || n < 1 || K < 1 || !d_freq || !d_h || !d_m0 || !d_mk || !d_Nk) return fail(MEEMB200_EINVAL, "bad argument");
    DeviceGuard guard(ctx->device);
    const int64_t t = (int64_t)n * K;
    k_roots_flat<<<(unsigned)((t + 127) / 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(n, K, d_freq, d_h, freq_kind, root_mode, d_m0, d_mk, d_Nk);
    CUDA_TRY(cudaGetLastError());
    return MEEMB200_OK;
}

extern "C" int meemb200_bessel(meemb200_ctx* ctx, int32_t kind, int64_t n, const double* d_x, double* d_out, void* stream) {
    if (!ctx || kind < 0 || kind > 7 || n < 1 || !d_x || !d_out) return fail(MEEMB200_EINVAL, "bad argument");
    DeviceGuard guard(ctx->device);
    k_bessel<<<(unsigned)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(kind, n, d_x, d_out);
    CUDA_TRY(cudaGetLastError());
    return MEEMB200_OK;
}

extern "C" int meemb200_fp64_peak(meemb200_ctx* ctx, int32_t kind, double* flops, void* stream) {
    if (!ctx || !flops) return fail(MEEMB200_EINVAL, "bad argument");
    DeviceGuard guard(ctx->device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int blocks = ctx->sm_count * 8, threads = 256, iters = 1 << 15;
    double* buf = nullptr;
    CUDA_TRY(cudaMalloc(&buf, sizeof(double) * blocks * threads));
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0, s);
        if (kind == 0) k_dfma_peak<<<blocks, threads, 0, s>>>(buf, iters);
        else k_dmma_peak<<<blocks, threads, 0, s>>>(buf, iters);
        cudaEventRecord(e1, s);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0 && ms < best) best = ms;
    }
    cudaError_t e = cudaGetLastError();
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(buf);
    if (e != cudaSuccess) return fail(MEEMB200_ECUDA, cudaGetErrorString(e));
    const double per_thread = kind == 0 ? 8.0 * 2.0 * iters : 4.0 * (8 * 8 * 4 * 2.0 / 32.0) * iters;
    *flops = per_thread * blocks * threads / (best * 1e-3);
    return MEEMB200_OK;
}
