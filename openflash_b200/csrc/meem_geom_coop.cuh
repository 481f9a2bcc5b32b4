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
    unsigned long long wkey[kCoopThreads / 32];
    int widx[kCoopThreads / 32];
    double L11[kCoopNB][kCoopNB + 1];
    double top[kCoopNB][8];
    double Us[kCoopNB][12];
    double xrow[2][kCoopNB];          // pivot-row / row-j exchange of the register-resident panel factorisation
    int mv_dst[kCoopMaxMoves], mv_src[kCoopMaxMoves];
    int piv[kCoopNB];
    int n_mv;
    int status;
};

// scratch published by the panel owner (global memory, after the matrix in the slot)
struct CoopScratch {
    double* Lbuf;     // [rows][16] multipliers of the rows below the panel's diagonal block (panel-local row index)
    double* L11g;     // [16][16] inverse of the unit-lower diagonal block L11
    int* moves;       // [0] = n_mv, then dst[32], src[32]
};

__device__ void coop_publish_block(int k0, int nb, CoopShared& S, const CoopScratch& sc);

// Warp-wide arg-max of 64-bit keys (bit patterns of non-negative doubles), smallest index among equal keys.
// On return every lane holds the winning key and index.
__device__ __forceinline__ void coop_warp_argmax(unsigned long long& key, int& idx) {
    const unsigned khi = (unsigned)(key >> 32), klo = (unsigned)key;
    const unsigned mhi = __reduce_max_sync(0xffffffffu, khi);
    const unsigned mlo = __reduce_max_sync(0xffffffffu, khi == mhi ? klo : 0u);
    const bool win = (khi == mhi) && (klo == mlo);
    idx = (int)__reduce_min_sync(0xffffffffu, win ? (unsigned)idx : 0x7fffffffu);
    key = ((unsigned long long)mhi << 32) | mlo;
}

// Panel factorisation by one CTA (fallback for panels taller than the register-resident variant handles).  panel: shared memory, column-major [nb][prow], prow >= rows.
// srows: only the first srows rows of the panel are pivot candidates (rows beyond it - the c rows of the static-pivot
// window kernel - are eliminated but never chosen).
__device__ void coop_panel_factor(double* W, long ld, int k0, int nb, int rows, int srows, double* panel, int prow,
                                  CoopShared& S, const CoopScratch& sc) {
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    for (int i = tid; i < rows; i += nthr) {
        const double* src = W + (long)(k0 + i) * ld + k0;
        for (int q = 0; q < nb; ++q) panel[q * prow + i] = src[q];
    }
    __syncthreads();
    for (int j = 0; j < nb; ++j) {
        double best = -1.0;
        int bi = 0x7fffffff;
        for (int i = j + tid; i < srows; i += nthr) {
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
        if (bi >= srows) bi = j;                                 // column of NaNs
        if (tid == 0) {
            S.piv[j] = bi;
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
    // publish: U11 -> W, strict-lower L11 -> S.L11, L21 -> Lbuf (zero-padded to 16 columns)
    for (int idx = tid; idx < kCoopNB * kCoopNB; idx += nthr) {
        const int i = idx >> 4, q = idx & 15;
        if (i < nb && q < nb && q >= i) W[(long)(k0 + i) * ld + k0 + q] = panel[q * prow + i];
        S.L11[i][q] = (i < nb && q < i) ? panel[q * prow + i] : 0.0;
    }
    for (int idx = tid; idx < (rows - nb) * kCoopNB; idx += nthr) {
        const int i = nb + (idx >> 4), q = idx & 15;
        sc.Lbuf[(long)i * kCoopNB + q] = (q < nb) ? panel[q * prow + i] : 0.0;
    }
    coop_publish_block(k0, nb, S, sc);
}

// After a panel factorisation: S.L11 holds the unit-lower L11 (strictly-lower part), S.piv the pivots.  Publishes
// L11^{-1} and the net row permutation.  Called by all threads of the owner CTA.
__device__ void coop_publish_block(int k0, int nb, CoopShared& S, const CoopScratch& sc) {
    const int tid = threadIdx.x;
    __syncthreads();
    if (tid < kCoopNB) {      // column `tid` of L11^{-1}: x = e_c - strict_lower(L11) x, forward
        double x[kCoopNB];
#pragma unroll
        for (int i = 0; i < kCoopNB; ++i) {
            double acc = (i == tid) ? 1.0 : 0.0;
#pragma unroll
            for (int q = 0; q < kCoopNB; ++q)
                if (q < i) acc = fma(-S.L11[i][q], x[q], acc);
            x[i] = acc;
        }
#pragma unroll
        for (int i = 0; i < kCoopNB; ++i) sc.L11g[i * kCoopNB + tid] = x[i];
    }
    if (tid >= 32 && tid < 64) {      // warp 1: net permutation of the nb swaps on the (<= 2 nb) rows they touch
        const int lane = tid & 31;
        int idx = (lane < nb) ? lane : -1, org = idx, n = nb;
        for (int j = 0; j < nb; ++j) {
            const int pr = S.piv[j];
            const unsigned hit = __ballot_sync(0xffffffffu, idx == pr);
            int b;
            if (hit) b = __ffs(hit) - 1;
            else { b = n++; if (lane == b) { idx = pr; org = pr; } }
            const int oj = __shfl_sync(0xffffffffu, org, j), ob = __shfl_sync(0xffffffffu, org, b);
            if (lane == j) org = ob;
            if (lane == b) org = oj;
        }
        // the first nb entries are always the top rows (possibly identity moves): the update phase needs them
        const bool keep = (lane < n) && (lane < nb || org != idx);
        const unsigned km = __ballot_sync(0xffffffffu, keep);
        if (keep) {
            const int m = __popc(km & ((1u << lane) - 1u));
            sc.moves[1 + m] = k0 + idx;
            sc.moves[1 + kCoopMaxMoves + m] = k0 + org;
        }
        if (lane == 0) sc.moves[0] = __popc(km);
    }
}

// Register-resident panel factorisation: thread t owns rows t, t + nthr, ... (RPT of them) with their 16 panel
// entries in registers; per column one warp-shuffle arg-max, two block barriers and a 2 x 16-double row exchange
// through shared memory.  rows <= RPT * blockDim.x.
template <int RPT>
__device__ void coop_panel_factor_reg(double* W, long ld, int k0, int nb, int rows, int srows, CoopShared& S, const CoopScratch& sc) {
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    double r[RPT][kCoopNB];
#pragma unroll
    for (int t = 0; t < RPT; ++t) {
        const int i = tid + t * nthr;
        const double* src = W + (long)(k0 + i) * ld + k0;
#pragma unroll
        for (int q = 0; q < kCoopNB; ++q) r[t][q] = (i < rows && q < nb) ? src[q] : 0.0;
    }
#pragma unroll
    for (int j = 0; j < kCoopNB; ++j) {
        if (j >= nb) break;
        // arg-max of |a| with the smallest row index on ties.  |a| >= 0, so the IEEE bit pattern orders like an
        // unsigned integer and the warp reduction is three redux.sync instead of five double shuffles + compares.
        unsigned long long key = 0ull;
        int bi = 0x7fffffff;
#pragma unroll
        for (int t = 0; t < RPT; ++t) {
            const int i = tid + t * nthr;
            const unsigned long long kv = (unsigned long long)__double_as_longlong(fabs(r[t][j]));
            if (i >= j && i < srows && (bi == 0x7fffffff || kv > key)) { key = kv; bi = i; }
        }
        coop_warp_argmax(key, bi);
        if (lane == 0) { S.wkey[warp] = key; S.widx[warp] = bi; }
        __syncthreads();
        key = (lane < nwarp) ? S.wkey[lane] : 0ull;
        bi = (lane < nwarp) ? S.widx[lane] : 0x7fffffff;
        coop_warp_argmax(key, bi);
        const double best = __longlong_as_double((long long)key);
        if (bi >= srows) bi = j;                                 // no candidate row
        if (tid == 0) {
            S.piv[j] = bi;
            if (!(best < INFINITY)) atomicOr(&S.status, 2);      // inf or NaN pivot column
            else if (best == 0.0) atomicOr(&S.status, 1);
        }
#pragma unroll
        for (int t = 0; t < RPT; ++t) {
            const int i = tid + t * nthr;
            if (i == bi) {
#pragma unroll
                for (int q = 0; q < kCoopNB; ++q) S.xrow[0][q] = r[t][q];
            }
            if (i == j) {
#pragma unroll
                for (int q = 0; q < kCoopNB; ++q) S.xrow[1][q] = r[t][q];
            }
        }
        __syncthreads();
        if (bi != j) {
#pragma unroll
            for (int t = 0; t < RPT; ++t) {
                const int i = tid + t * nthr;
                if (i == j) {
#pragma unroll
                    for (int q = 0; q < kCoopNB; ++q) r[t][q] = S.xrow[0][q];
                } else if (i == bi) {
#pragma unroll
                    for (int q = 0; q < kCoopNB; ++q) r[t][q] = S.xrow[1][q];
                }
            }
        }
        const double rpv = 1.0 / S.xrow[0][j];
#pragma unroll
        for (int t = 0; t < RPT; ++t) {
            const int i = tid + t * nthr;
            if (i > j && i < rows) {
                const double l = r[t][j] * rpv;                  // LAPACK dgetf2 scales by the reciprocal pivot too
                r[t][j] = l;
#pragma unroll
                for (int q = 0; q < kCoopNB; ++q)
                    if (q > j) r[t][q] = fma(-l, S.xrow[0][q], r[t][q]);
            }
        }
    }
    // publish: U11 -> W, strict-lower L11 -> S.L11, L21 -> Lbuf (zero beyond nb)
#pragma unroll
    for (int t = 0; t < RPT; ++t) {
        const int i = tid + t * nthr;
        if (i < kCoopNB) {
#pragma unroll
            for (int q = 0; q < kCoopNB; ++q) S.L11[i][q] = (i < nb && q < i) ? r[t][q] : 0.0;
        }
        if (i < nb) {
#pragma unroll
            for (int q = 0; q < kCoopNB; ++q)
                if (q < nb && q >= i) W[(long)(k0 + i) * ld + k0 + q] = r[t][q];
        } else if (i < rows) {
            double2* dst = reinterpret_cast<double2*>(sc.Lbuf + (long)i * kCoopNB);
#pragma unroll
            for (int q = 0; q < kCoopNB; q += 2) dst[q >> 1] = make_double2(r[t][q], r[t][q + 1]);
        }
    }
    coop_publish_block(k0, nb, S, sc);
}

// Update phase of one panel, run by every CTA on its own 8-column chunks of the columns >= kb.
__device__ void coop_panel_update(double* W, long ld, int p, int ncols, int k0, int nb, CoopShared& S, const CoopScratch& sc,
                                  unsigned long long* tmr) {
    const long long tu0 = clock64();
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
    const long long tu1 = clock64();
    if (tmr && blockIdx.x == 0 && threadIdx.x == 32) tmr[0] += (unsigned long long)(tu1 - tu0);
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
        if (tmr && blockIdx.x == 0 && threadIdx.x == 32 && c == 0) tmr[1] += (unsigned long long)(clock64() - tu1);
        // 2. U12 = L11^{-1} A12: thread (q, column) takes one dot product with row q of the inverse
        if (tid < kCoopNB * 8) {
            const int q = tid >> 3, cx = tid & 7;
            double a = 0.0;
            if (q < nb && cx < cw) {
#pragma unroll
                for (int q2 = 0; q2 < kCoopNB; ++q2)
                    if (q2 <= q) a = fma(S.L11[q][q2], S.top[q2][cx], a);
                W[(long)(k0 + q) * ld + col0 + cx] = a;
            }
            S.Us[q][cx] = a;
        }
        __syncthreads();
        // 3. A22 -= L21 U12 (8-row blocks per warp, A fragments straight from L2, four blocks in flight per warp)
        const int nrb = (rows_below + 7) >> 3;
        double bf[kCoopNB / 4];
#pragma unroll
        for (int ks = 0; ks < kCoopNB / 4; ++ks) bf[ks] = S.Us[ks * 4 + lc][lr];
        const int col = col0 + 2 * lc;
        const bool pair_ok = ((reinterpret_cast<size_t>(W + col) & 15) == 0);   // ld is even: 16-byte aligned pair (the pad column absorbs col + 1 == ncols)
        for (int rb0 = warp; rb0 < nrb; rb0 += 4 * nwarp) {
            double af[4][kCoopNB / 4], cv[4][2];
            bool rok[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int rb = rb0 + u * nwarp;
                rok[u] = (rb < nrb) && (rb * 8 + lr) < rows_below;
                const long i = nb + rb * 8 + lr;             // panel-local row of this lane's A fragment
#pragma unroll
                for (int ks = 0; ks < kCoopNB / 4; ++ks) af[u][ks] = rok[u] ? -sc.Lbuf[i * kCoopNB + ks * 4 + lc] : 0.0;
                cv[u][0] = cv[u][1] = 0.0;
                if (rok[u] && col < ncols) {
                    const double* pw = W + (long)(kb + rb * 8 + lr) * ld + col;
                    if (pair_ok) { const double2 v = *reinterpret_cast<const double2*>(pw); cv[u][0] = v.x; cv[u][1] = v.y; }
                    else { cv[u][0] = pw[0]; if (col + 1 < ncols) cv[u][1] = pw[1]; }
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int ks = 0; ks < kCoopNB / 4; ++ks) dmma(cv[u][0], cv[u][1], af[u][ks], bf[ks]);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int rb = rb0 + u * nwarp;
                if (rok[u] && col < ncols) {
                    double* pw = W + (long)(kb + rb * 8 + lr) * ld + col;
                    if (pair_ok) *reinterpret_cast<double2*>(pw) = make_double2(cv[u][0], cv[u][1]);
                    else { pw[0] = cv[u][0]; if (col + 1 < ncols) pw[1] = cv[u][1]; }
                }
            }
        }
        __syncthreads();
    }
}

// Back-substitution U x = y for the right-hand-side columns owned by this CTA (no grid sync): the CTA's columns are
// taken two at a time (shared U traffic), each kept whole in shared memory.  Per 16-row block, bottom-up: the rows
// of U that the update needs are prefetched into registers while one warp per column solves the 16 x 16 triangle.
// ycol: shared memory, 2 * p doubles.  p <= 2 * blockDim.x rows per prefetch pass (else the pass repeats).
__device__ void coop_backsolve(double* W, long ld, int p, int nrhs, double* ycol, CoopShared& S) {
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5;
    const int nblk = (p + kCoopNB - 1) / kCoopNB;
    const int pp = (p + kCoopNB - 1) & ~(kCoopNB - 1);          // columns padded with zeros to a whole block
    double* y0 = ycol;
    double* y1 = ycol + pp;
    __shared__ double rdiag[kCoopNB];
    for (int rc = 2 * blockIdx.x; rc < nrhs; rc += 2 * gridDim.x) {
        const int colA = p + rc;
        const bool two = rc + 1 < nrhs;
        __syncthreads();
        for (int i = tid; i < pp; i += nthr) {
            y0[i] = (i < p) ? W[(long)i * ld + colA] : 0.0;
            y1[i] = (two && i < p) ? W[(long)i * ld + colA + 1] : 0.0;
        }
        for (int blk = nblk - 1; blk >= 0; --blk) {
            const int k0 = blk * kCoopNB;
            const int nb = min(kCoopNB, p - k0);
            // prefetch this thread's first two update rows of U[:, k0 : k0 + 16) (the common case p <= 2 * nthr)
            double u[2][kCoopNB];
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                const int i = tid + t * nthr;
                if (i < k0) {
                    const double* urow = W + (long)i * ld + k0;
#pragma unroll
                    for (int q = 0; q < kCoopNB; q += 2) {
                        const double2 u2 = *reinterpret_cast<const double2*>(urow + q);
                        u[t][q] = (q < nb) ? u2.x : 0.0; u[t][q + 1] = (q + 1 < nb) ? u2.y : 0.0;
                    }
                }
            }
            for (int idx = tid; idx < kCoopNB * kCoopNB; idx += nthr) {
                const int q = idx >> 4, c = idx & 15;
                const double v = (q < nb && c < nb && c >= q) ? W[(long)(k0 + q) * ld + k0 + c] : (q == c ? 1.0 : 0.0);
                S.L11[q][c] = v;
                if (q == c) rdiag[q] = 1.0 / v;
            }
            __syncthreads();
            if (warp < 2) {
                double* y = warp ? y1 : y0;
                double c = (lane < nb) ? y[k0 + lane] : 0.0;
                for (int q = nb - 1; q >= 0; --q) {
                    const double x = __shfl_sync(0xffffffffu, c * rdiag[q], q);
                    if (lane == q) c = x;
                    else if (lane < q) c = fma(-S.L11[lane][q], x, c);
                }
                if (lane < nb) y[k0 + lane] = c;
            }
            __syncthreads();
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                const int i = tid + t * nthr;
                if (i < k0) {
                    double a0 = y0[i], a1 = y1[i];
#pragma unroll
                    for (int q = 0; q < kCoopNB; ++q) {
                        a0 = fma(-u[t][q], y0[k0 + q], a0);
                        a1 = fma(-u[t][q], y1[k0 + q], a1);
                    }
                    y0[i] = a0; y1[i] = a1;
                }
            }
            for (int i = tid + 2 * nthr; i < k0; i += nthr) {     // tall systems: remaining rows straight from L2
                const double* urow = W + (long)i * ld + k0;
                double a0 = y0[i], a1 = y1[i];
                for (int q = 0; q < nb; ++q) { a0 = fma(-urow[q], y0[k0 + q], a0); a1 = fma(-urow[q], y1[k0 + q], a1); }
                y0[i] = a0; y1[i] = a1;
            }
            __syncthreads();
        }
        for (int i = tid; i < p; i += nthr) {
            W[(long)i * ld + colA] = y0[i];
            if (two) W[(long)i * ld + colA + 1] = y1[i];
        }
    }
}

// Static-pivot elimination of a block of Nb unknowns whose pivot rows form a DIAGONAL block (the C1 columns of a region
// against the potential rows of its outer boundary, k_geom_coop2): for every row r < nrows and column c < ncols
//   W[r][c] -= sum_k (W[r][c1L + k] * dinv[k]) * W[potr0 + k][c]
// as a grid-wide DMMA product, 8 x 32 tiles per warp, operand fragments straight from L2 (one use each).
__device__ void coop2_static_update(double* W, long ld, int nrows, int ncols, int c1L, int potr0, int Nb, const double* dinv) {
    const int lane = threadIdx.x & 31;
    const long gwarp = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, gnwarp = ((long)gridDim.x * blockDim.x) >> 5;
    const int lr = lane >> 2, lc = lane & 3;
    const int ntr = (nrows + 7) >> 3, ntc = (ncols + 31) >> 5;
    for (long t = gwarp; t < (long)ntr * ntc; t += gnwarp) {
        const int r = (int)(t / ntc) * 8 + lr, cb = (int)(t % ntc) * 32;
        const bool rok = r < nrows;
        double acc[4][2];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = cb + j * 8 + 2 * lc;
            acc[j][0] = (rok && c < ncols) ? W[(long)r * ld + c] : 0.0;
            acc[j][1] = (rok && c + 1 < ncols) ? W[(long)r * ld + c + 1] : 0.0;
        }
        const double* arow = W + (long)(rok ? r : 0) * ld + c1L;
#pragma unroll 4
        for (int k0 = 0; k0 < Nb; k0 += 4) {
            const int k = k0 + lc;
            const bool kok = k < Nb;
            const double a = (rok && kok) ? -(arow[k] * dinv[k]) : 0.0;
            const double* brow = W + (long)(potr0 + (kok ? k : 0)) * ld + cb + lr;
            double b[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = (kok && cb + j * 8 + lr < ncols) ? brow[j * 8] : 0.0;
#pragma unroll
            for (int j = 0; j < 4; ++j) dmma(acc[j][0], acc[j][1], a, b[j]);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = cb + j * 8 + 2 * lc;
            if (rok && c < ncols) W[(long)r * ld + c] = acc[j][0];
            if (rok && c + 1 < ncols) W[(long)r * ld + c + 1] = acc[j][1];
        }
    }
}

}  // namespace meem
