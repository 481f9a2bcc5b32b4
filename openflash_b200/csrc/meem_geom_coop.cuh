// Grid-cooperative geometry stage: partial-pivot LU of A11 (p x p real) with the right-hand sides
// [E | b1_i] riding along, spread over every SM of the B200 (cooperative launch, grid.sync()).
// Used when a batch has few, large geometries (big_tricylinder: one geometry, p = 1000), where the
// one-CTA-per-geometry kernel k_geom would leave 147 SMs idle.
//
// Per panel of kCoopNB columns:
//   owner CTA   factorises the panel in shared memory (pivot = max |a|, first wins, like LAPACK
//               dgetf2), publishes L21, L11, U11 and the NET row permutation of the panel as a list
//               of (dst, src) row moves
//   grid.sync
//   all CTAs    own 8-column chunks of everything right of the panel: gather the moved rows, solve
//               U12 = L11^{-1} A12 (one thread per column), A22 -= L21 U12 on the DMMA pipe with
//               the A fragments read straight from L2 (each L21 element is used once per chunk)
//   grid.sync
// Back-substitution needs no grid synchronisation: every right-hand-side column is independent, so
// each CTA owns whole columns, keeps them in shared memory and streams U from L2.
// Included by meem_kernels.cu after DevView / GeomDesc are defined.
#pragma once
#include <cooperative_groups.h>
#include "meem_dmma.cuh"

namespace cg = cooperative_groups;

namespace meem {

constexpr int kCoopNB = 16;
constexpr int kCoopThreads = 512;
constexpr int kCoopMaxMoves = 2 * kCoopNB;

struct CoopShared {
    double wbest[kCoopThreads / 32];
    int widx[kCoopThreads / 32];
    double L11[kCoopNB][kCoopNB + 1];
    double top[kCoopNB][8];
    double Us[kCoopNB][12];
    int mv_dst[kCoopMaxMoves], mv_src[kCoopMaxMoves];
    int n_mv;
    int status;
};

// scratch published by the panel owner (global memory, after the matrix in the slot)
struct CoopScratch {
    double* Lbuf;     // [rows][16] multipliers of the rows below the panel's diagonal block (panel-local row index)
    double* L11g;     // [16][16]
    int* moves;       // [0] = n_mv, then dst[32], src[32]
};

// Panel factorisation by one CTA.  panel: shared memory, column-major [nb][prow], prow >= rows.
__device__ void coop_panel_factor(double* W, long ld, int k0, int nb, int rows, double* panel, int prow,
                                  CoopShared& S, const CoopScratch& sc) {
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    __shared__ int piv[kCoopNB];
    for (int i = tid; i < rows; i += nthr) {
        const double* src = W + (long)(k0 + i) * ld + k0;
        for (int q = 0; q < nb; ++q) panel[q * prow + i] = src[q];
    }
    __syncthreads();
    for (int j = 0; j < nb; ++j) {
        double best = -1.0;
        int bi = 0x7fffffff;
        for (int i = j + tid; i < rows; i += nthr) {
            const double v = fabs(panel[j * prow + i]);
            if (v > best) { best = v; bi = i; }
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
            const double ov = __shfl_down_sync(0xffffffffu, best, off);
            const int oi = __shfl_down_sync(0xffffffffu, bi, off);
            if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        }
        if (lane == 0) { S.wbest[warp] = best; S.widx[warp] = bi; }
        __syncthreads();
        best = S.wbest[0]; bi = S.widx[0];
        for (int w = 1; w < nwarp; ++w) {
            const double ov = S.wbest[w];
            const int oi = S.widx[w];
            if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        }
        if (bi >= rows) bi = j;                                  // column of NaNs
        if (tid == 0) {
            piv[j] = bi;
            if (!(best < INFINITY) || best < 0.0) atomicOr(&S.status, 2);
            else if (best == 0.0) atomicOr(&S.status, 1);
        }
        if (bi != j && tid < nb) {
            const double t = panel[tid * prow + j];
            panel[tid * prow + j] = panel[tid * prow + bi];
            panel[tid * prow + bi] = t;
        }
        __syncthreads();
        const double pv = panel[j * prow + j];
        for (int i = j + 1 + tid; i < rows; i += nthr) {
            const double l = panel[j * prow + i] / pv;
            panel[j * prow + i] = l;
            for (int q = j + 1; q < nb; ++q) panel[q * prow + i] = fma(-l, panel[q * prow + j], panel[q * prow + i]);
        }
    }
    __syncthreads();
    // publish: U11 -> W, L11 -> scratch, L21 -> Lbuf (zero-padded to 16 columns), net permutation
    for (int idx = tid; idx < kCoopNB * kCoopNB; idx += nthr) {
        const int i = idx >> 4, q = idx & 15;
        if (i < nb && q < nb) {
            if (q >= i) W[(long)(k0 + i) * ld + k0 + q] = panel[q * prow + i];
            sc.L11g[idx] = (q < i) ? panel[q * prow + i] : 0.0;
        } else {
            sc.L11g[idx] = 0.0;
        }
    }
    for (int idx = tid; idx < (rows - nb) * kCoopNB; idx += nthr) {
        const int i = nb + (idx >> 4), q = idx & 15;
        sc.Lbuf[(long)i * kCoopNB + q] = (q < nb) ? panel[q * prow + i] : 0.0;
    }
    if (tid == 0) {
        // simulate the nb swaps on the (at most 2 nb) rows they touch
        int idx[kCoopMaxMoves], org[kCoopMaxMoves], n = 0;
        for (int q = 0; q < nb; ++q) { idx[n] = q; org[n] = q; ++n; }
        for (int j = 0; j < nb; ++j) {
            const int pr = piv[j];
            int b = -1;
            for (int s = 0; s < n; ++s) if (idx[s] == pr) { b = s; break; }
            if (b < 0) { idx[n] = pr; org[n] = pr; b = n++; }
            const int t = org[j]; org[j] = org[b]; org[b] = t;
        }
        // the first nb entries are always the top rows (possibly identity moves): the update phase needs them
        int m = 0;
        for (int s = 0; s < n; ++s)
            if (s < nb || org[s] != idx[s]) { sc.moves[1 + m] = k0 + idx[s]; sc.moves[1 + kCoopMaxMoves + m] = k0 + org[s]; ++m; }
        sc.moves[0] = m;
    }
}

// Update phase of one panel, run by every CTA on its own 8-column chunks of the columns >= kb.
__device__ void coop_panel_update(double* W, long ld, int p, int ncols, int k0, int nb, CoopShared& S, const CoopScratch& sc) {
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    const int lr = lane >> 2, lc = lane & 3;
    const int kb = k0 + nb;
    const int nchunks = (ncols - kb + 7) >> 3;
    if ((int)blockIdx.x >= nchunks) return;
    for (int idx = tid; idx < kCoopNB * kCoopNB; idx += nthr) S.L11[idx >> 4][idx & 15] = sc.L11g[idx];
    if (tid < kCoopMaxMoves) { S.mv_dst[tid] = sc.moves[1 + tid]; S.mv_src[tid] = sc.moves[1 + kCoopMaxMoves + tid]; }
    if (tid == 0) S.n_mv = sc.moves[0];
    __syncthreads();
    const int n_mv = S.n_mv;
    const int rows_below = p - kb;
    for (int c = blockIdx.x; c < nchunks; c += gridDim.x) {
        const int col0 = kb + 8 * c;
        const int cw = min(8, ncols - col0);
        // 1. gather the permuted rows (all reads before all writes)
        double val = 0.0;
        const int s = tid >> 3, cc = tid & 7;
        const bool mine = (s < n_mv) && (cc < cw);
        if (mine) val = W[(long)S.mv_src[s] * ld + col0 + cc];
        __syncthreads();
        if (mine) {
            if (S.mv_dst[s] != S.mv_src[s]) W[(long)S.mv_dst[s] * ld + col0 + cc] = val;
            if (s < nb) S.top[s][cc] = val;
        }
        __syncthreads();
        // 2. U12 = L11^{-1} A12, one thread per column
        if (tid < 8) {
            double u[kCoopNB];
#pragma unroll
            for (int q = 0; q < kCoopNB; ++q) {
                double a = (q < nb && tid < cw) ? S.top[q][tid] : 0.0;
#pragma unroll
                for (int q2 = 0; q2 < kCoopNB; ++q2)
                    if (q2 < q) a = fma(-S.L11[q][q2], u[q2], a);
                u[q] = a;
                S.Us[q][tid] = a;
                if (q < nb && tid < cw) W[(long)(k0 + q) * ld + col0 + tid] = a;
            }
        }
        __syncthreads();
        // 3. A22 -= L21 U12 (8-row blocks per warp, A fragments straight from L2)
        const int nrb = (rows_below + 7) >> 3;
        double bf[kCoopNB / 4];
#pragma unroll
        for (int ks = 0; ks < kCoopNB / 4; ++ks) bf[ks] = S.Us[ks * 4 + lc][lr];
        for (int rb = warp; rb < nrb; rb += nwarp) {
            const int i = nb + rb * 8 + lr;                  // panel-local row of this lane's A fragment
            const bool rok = (rb * 8 + lr) < rows_below;
            double c0 = 0.0, c1 = 0.0;
            double af[kCoopNB / 4];
#pragma unroll
            for (int ks = 0; ks < kCoopNB / 4; ++ks) af[ks] = rok ? sc.Lbuf[(long)i * kCoopNB + ks * 4 + lc] : 0.0;
#pragma unroll
            for (int ks = 0; ks < kCoopNB / 4; ++ks) dmma(c0, c1, af[ks], bf[ks]);
            const int col = col0 + 2 * lc;
            if (rok && col < ncols) {
                double* pw = W + (long)(kb + rb * 8 + lr) * ld + col;
                if ((col & 1) == 0) {          // ld is even: 16-byte aligned pair (the pad column absorbs col + 1 == ncols)
                    double2* ptr = reinterpret_cast<double2*>(pw);
                    double2 v = *ptr;
                    v.x -= c0; v.y -= c1;
                    *ptr = v;
                } else {
                    pw[0] -= c0;
                    if (col + 1 < ncols) pw[1] -= c1;
                }
            }
        }
        __syncthreads();
    }
}

// Back-substitution U x = y for the right-hand-side columns owned by this CTA (no grid sync).
// ycol: shared memory, p doubles.
__device__ void coop_backsolve(double* W, long ld, int p, int nrhs, double* ycol, CoopShared& S) {
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5;
    const int nblk = (p + kCoopNB - 1) / kCoopNB;
    for (int rc = blockIdx.x; rc < nrhs; rc += gridDim.x) {
        const int col = p + rc;
        __syncthreads();
        for (int i = tid; i < p; i += nthr) ycol[i] = W[(long)i * ld + col];
        for (int blk = nblk - 1; blk >= 0; --blk) {
            const int k0 = blk * kCoopNB;
            const int nb = min(kCoopNB, p - k0);
            __syncthreads();
            for (int idx = tid; idx < kCoopNB * kCoopNB; idx += nthr) {
                const int q = idx >> 4, c = idx & 15;
                S.L11[q][c] = (q < nb && c < nb && c >= q) ? W[(long)(k0 + q) * ld + k0 + c] : (q == c ? 1.0 : 0.0);
            }
            __syncthreads();
            if (warp == 0) {
                double c = (lane < nb) ? ycol[k0 + lane] : 0.0;
                for (int q = nb - 1; q >= 0; --q) {
                    double x = 0.0;
                    if (lane == q) x = c / S.L11[q][q];
                    x = __shfl_sync(0xffffffffu, x, q);
                    if (lane == q) c = x;
                    else if (lane < q) c = fma(-S.L11[lane][q], x, c);
                }
                if (lane < nb) ycol[k0 + lane] = c;
            }
            __syncthreads();
            for (int i = tid; i < k0; i += nthr) {
                const double* urow = W + (long)i * ld + k0;
                double acc = ycol[i];
                if (nb == kCoopNB) {
#pragma unroll
                    for (int q = 0; q < kCoopNB; q += 2) {
                        const double2 u2 = *reinterpret_cast<const double2*>(urow + q);
                        acc = fma(-u2.x, ycol[k0 + q], acc);
                        acc = fma(-u2.y, ycol[k0 + q + 1], acc);
                    }
                } else {
                    for (int q = 0; q < nb; ++q) acc = fma(-urow[q], ycol[k0 + q], acc);
                }
                ycol[i] = acc;
            }
        }
        __syncthreads();
        for (int i = tid; i < p; i += nthr) W[(long)i * ld + col] = ycol[i];
    }
}

}  // namespace meem
