// Geometry stage by block-bidiagonal ("window") elimination of A11 = A[:p,:p]  (SURVEY §8f rank 4).
//
// The reference's A (meem_engine.py:164-276) is, after grouping the potential and velocity rows of each boundary, block
// bi-diagonal in the unknown blocks u_0 = C1^0 (N_0), u_i = (C1^i, C2^i) (2 N_i): boundary bd couples only u_bd and
// u_bd+1 (paper: pubs/JFM/FLMguide.tex:1470-1475, "block bi-diagonal ... Thomas algorithm"; many-region generator
// dev/python/slants/helpers.py:80-125).  Window bd holds
//
//      rows   [ left over from window bd-1 (N_bd) | potential rows of bd | velocity rows of bd | c rows (R) ]
//      cols   [ u_bd | u_bd+1 (last window: E, the N columns that carry -I_mk x2) | b_i (R) ]
//
// and eliminates the columns of u_bd with partial pivoting among the window's rows.  Every row of A11 with a non-zero in
// a pivot column is in the window, so the pivots are exactly those of LAPACK's dense getrf on A11 (the reference's
// scipy.linalg.lu_factor, meem_engine.py:609) - only the zero blocks are skipped: O(M N^3) instead of O(M^3 N^3) flops,
// and the working set is one (3N+R) x (4N+R) window (25 KB at N = 16: shared memory) instead of the p x (p+N+R)
// augmented matrix.  The c vectors of the radiating modes (meem_engine.py:65-98) ride along as R extra rows that are
// eliminated but never chosen as pivots; when the last unknown is gone their right-hand-side part holds
// -c_j^T A11^-1 [E | b_i] = -[w_j | s0_ji], so neither G = A11^-1 E nor the factors are ever stored: the frequency stage
// needs H, u (rows of the LAST region of [G | y0]: one back-substitution inside the last window), w and s0 only.
// Numerical probe against a dense LAPACK solve: tools/experiments/exp_band.py (<= 7e-15 on every case tried).
#pragma once
#include "meem_lu.cuh"

namespace meem {

struct BandDims {
    int wrows, wcols;      // largest window (rows incl. the R c rows, columns incl. the R right-hand sides)
    int side;              // doubles of the hand-over buffer between two windows
};

MEEM_HDX int band_unknowns(const Geom& g, int i) { return (i == 0 ? 1 : 2) * g.nmk[i]; }

MEEM_HDX void band_dims(const Geom& g, const Layout& L, int R, BandDims& D) {
    const int M = g.M;
    D.wrows = D.wcols = D.side = 0;
    int nc = 0;
    for (int bd = 0; bd < M; ++bd) {
        const bool last = bd == M - 1;
        const int n0 = band_unknowns(g, bd), n1 = last ? g.nmk[M - 1] : band_unknowns(g, bd + 1);
        const int nsearch = nc + L.pot_h[bd] + (last ? 0 : L.vel_h[bd]);
        if (nsearch + R > D.wrows) D.wrows = nsearch + R;
        if (n0 + n1 + R > D.wcols) D.wcols = n0 + n1 + R;
        if (!last) {
            const int s = (nsearch - n0 + R) * (n1 + R);
            if (s > D.side) D.side = s;
        }
        nc = nsearch - n0;
    }
}

#ifdef __CUDACC__

// All threads of the CTA must call.  W: window buffer [wrows][ld] (shared or global), side: hand-over buffer,
// lu_smem: lu_smem_bytes<double, NB, CHUNK>(wrows).  On return, with n = unknowns of the last region and N = nmk[M-1]:
// rows [0, n) x cols [n, n+N+R) = rows of the last region of [G | y0_i]; rows [n, n+R) x the same columns = -[w_j | s0_ji].
// Static-pivot step of a case-1 window (potential rows FIRST): the C1 block of the potential rows is the diagonal matrix
// (h - d_bd) diag R_1n(a_bd), so its Nb columns are eliminated without a pivot search and without a panel factorisation -
// per NB columns: multipliers W[r][k] / D1_k of the rows below and the pivot rows' other columns are staged, then one
// rank-NB update by all warps.  Probe: tools/experiments/exp_band.py --static (<= 9e-15 incl. mixed case-1 / case-2 geometries).
template <int NB, bool SMALL, int CHUNK>
__device__ void band_static_step(double* W, long ld, int nrows, int ncols, int Nb, unsigned char* smem_raw) {
    double* panel = reinterpret_cast<double*>(smem_raw);     // [NB][nrows]: multipliers of rows Nb .. nrows-1
    double* ust = panel + (size_t)NB * nrows;                // [NB][CHUNK]
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int below = nrows - Nb;
    for (int k0 = 0; k0 < Nb; k0 += NB) {
        const int nbj = min(NB, Nb - k0);
        for (int idx = tid; idx < below * nbj; idx += nthr) {
            const int i = idx / nbj, q = idx - i * nbj;
            panel[q * nrows + i] = W[(long)(Nb + i) * ld + k0 + q] / W[(long)(k0 + q) * ld + k0 + q];
        }
        for (int cc0 = Nb; cc0 < ncols; cc0 += CHUNK) {
            const int cc1 = min(ncols, cc0 + CHUNK);
            for (int idx = tid; idx < nbj * (cc1 - cc0); idx += nthr) {
                const int q = idx / (cc1 - cc0), c = idx - q * (cc1 - cc0);
                ust[q * CHUNK + c] = W[(long)(k0 + q) * ld + cc0 + c];
            }
            __syncthreads();
            lu_rank_update<double, NB, SMALL, CHUNK>(W, ld, Nb, nrows, Nb, panel, nrows, cc0, cc1, ust, nbj);
            __syncthreads();
        }
    }
}

template <int NB, bool SMALL, int CHUNK>
__device__ void band_eliminate(const Geom& g, const Layout& L, const double* tab, const unsigned char* flags, int R,
                               double* W, long ld, double* side, unsigned char* lu_smem, int* status, bool use_static) {
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int M = g.M;
    int nc = 0;
    for (int bd = 0; bd < M; ++bd) {
        const bool last = bd == M - 1;
        const int n0 = band_unknowns(g, bd), n1 = last ? g.nmk[M - 1] : band_unknowns(g, bd + 1);
        const int hp = L.pot_h[bd], hv = last ? 0 : L.vel_h[bd];
        const int nsearch = nc + hp + hv, nrows = nsearch + R, ncols = n0 + n1 + R;
        // case-1 boundary (and the last one): potential rows first, their diagonal C1 block is eliminated without a search
        const bool stat = use_static && (last || g.d[bd] > g.d[bd + 1]);
        const int r_car = stat ? hp : 0, r_pot = stat ? 0 : nc, r_vel = nc + hp;
        for (int e = tid; e < nrows * ncols; e += nthr) {
            const int r = e / ncols, c = e - r * ncols;
            W[(long)r * ld + c] = 0.0;
        }
        __syncthreads();
        if (bd > 0) {          // rows left over from the previous window and the c rows: columns [u_bd | b]
            const int sw = n0 + R;
            for (int e = tid; e < (nc + R) * sw; e += nthr) {
                const int r = e / sw, c = e - r * sw;
                const int wr = r < nc ? r_car + r : nsearch + (r - nc);
                const int wc = c < n0 ? c : n0 + n1 + (c - n0);
                W[(long)wr * ld + wc] = side[e];
            }
        } else {
            for (int e = tid; e < R * n0; e += nthr) {
                const int j = e / n0, n = e - j * n0;
                W[(long)(nsearch + j) * ld + n] = c_entry(g, L, tab, flags + (long)j * (M + 1), 0, 0, n);
            }
        }
        if (!last) {           // c entries of region bd+1 enter with its columns
            const int Nr = g.nmk[bd + 1];
            for (int e = tid; e < R * n1; e += nthr) {
                const int j = e / n1, c = e - j * n1;
                W[(long)(nsearch + j) * ld + n0 + c] =
                    c_entry(g, L, tab, flags + (long)j * (M + 1), bd + 1, c < Nr ? 0 : 1, c < Nr ? c : c - Nr);
            }
        } else {
            for (int n = tid; n < n1; n += nthr) W[(long)(r_pot + n) * ld + n0 + n] = 1.0;     // E
        }
        fill_A_boundary<double>(g, L, tab, bd, W, ld, r_pot, r_vel, L.col[bd], tid, nthr);
        for (int e = tid; e < R * (hp + hv); e += nthr) {
            const int i = e / (hp + hv), n = e - i * (hp + hv);
            const bool vel = n >= hp;
            W[(long)(vel ? r_vel + (n - hp) : r_pot + n) * ld + n0 + n1 + i] =
                b_static_entry(g, L, flags + (long)i * (M + 1), bd, vel, vel ? n - hp : n);
        }
        __syncthreads();
        if (stat) {
            band_static_step<NB, SMALL, CHUNK>(W, ld, nrows, ncols, hp, lu_smem);
            lu_forward<double, NB, SMALL, CHUNK>(W + (long)hp * ld + hp, ld, nrows - hp, n0 - hp, ncols - hp, lu_smem, status, nsearch - hp);
        } else {
            lu_forward<double, NB, SMALL, CHUNK>(W, ld, nrows, n0, ncols, lu_smem, status, nsearch);
        }
        if (last) {
            lu_backward<double, NB, SMALL, CHUNK>(W, ld, n0, n0, ncols, lu_smem, n0);
            __syncthreads();
            break;
        }
        const int sw = n1 + R;
        for (int e = tid; e < (nrows - n0) * sw; e += nthr) {
            const int r = e / sw, c = e - r * sw;
            side[e] = W[(long)(n0 + r) * ld + n0 + c];
        }
        __syncthreads();
        nc = nsearch - n0;
    }
}

#endif  // __CUDACC__
}  // namespace meem
