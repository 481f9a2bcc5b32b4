// CTA-cooperative blocked Gaussian elimination with partial pivoting (FP64 real or complex).
//
// Replaces scipy.linalg.lu_factor / lu_solve / solve (LAPACK zgetrf / zgetrs / zgesv) on the
// reference's hot path (meem_engine.py:312, :609, :633).  Same algorithm as LAPACK's right-looking
// getrf: unscaled partial pivoting with the |re|+|im| pivot measure (izamax), first maximum wins.
// No row equilibration — it destroys these systems.
//
// The matrix W is row-major [nrows x ld] in generic memory (shared or L2-resident global) and holds
// the right-hand sides as extra columns, so the forward substitution is fused into the elimination
// and L is never stored.  Panels of NB columns are staged column-major in shared memory and
// factorised by ONE warp (pivot search = warp-shuffle arg-max, no block barrier per column);
// the trailing matrix is updated by all warps with register-tiled rank-NB updates whose L operand
// is broadcast from shared memory and whose U operand is staged per column chunk.
#pragma once
#include "meem_forms.cuh"

namespace meem {

constexpr int kLuChunk = 512;   // columns of U12 staged per pass

template <class T, bool SMALL = false> struct LuTile;            // register tile of the rank-NB update
template <> struct LuTile<double, false> { static constexpr int TR = 8, TC = 4; };
template <> struct LuTile<cplx, false>   { static constexpr int TR = 4, TC = 2; };
// SMALL: half the accumulators, for the many-small-geometries kernel that trades tile efficiency for resident warps
template <> struct LuTile<double, true>  { static constexpr int TR = 4, TC = 2; };
template <> struct LuTile<cplx, true>    { static constexpr int TR = 4, TC = 2; };

MEEM_HD double zero_of(double) { return 0.0; }
MEEM_HD cplx zero_of(cplx) { return cmake(0.0, 0.0); }
MEEM_HD void fnma(double& acc, double a, double b) { acc = fma(-a, b, acc); }
MEEM_HD void fnma(cplx& acc, cplx a, cplx b) {
    acc.re = fma(-a.re, b.re, acc.re); acc.re = fma(a.im, b.im, acc.re);
    acc.im = fma(-a.re, b.im, acc.im); acc.im = fma(-a.im, b.re, acc.im);
}
MEEM_HD bool finite_of(double a) { return isfinite(a); }
MEEM_HD bool finite_of(cplx a) { return isfinite(a.re) && isfinite(a.im); }

template <class T, int NB, int CHUNK = kLuChunk>
__host__ __device__ constexpr size_t lu_smem_bytes(int max_rows) {
    return sizeof(T) * (size_t)NB * ((size_t)max_rows + CHUNK) + 16 * sizeof(int);
}

#ifdef __CUDACC__

// W[i][c] -= sum_q Lp[q*lstride + (i - row0)] * Us[q*CHUNK + (c - c0)]  for i in [r_begin, r_end), c in [c0, c1)
template <class T, int NB, bool SMALL = false, int CHUNK = kLuChunk>
__device__ void lu_rank_update(T* W, long ld, int r_begin, int r_end, int row0, const T* Lp, int lstride,
                               int c0, int c1, const T* Us, int nbj) {
    constexpr int TR = LuTile<T, SMALL>::TR, TC = LuTile<T, SMALL>::TC;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    const int nrt = (r_end - r_begin + TR - 1) / TR;
    const int nct = (c1 - c0 + 32 * TC - 1) / (32 * TC);
    for (int t = warp; t < nrt * nct; t += nwarp) {
        const int rt = t / nct, ct = t - rt * nct;
        const int i0 = r_begin + rt * TR;
        const int cb = c0 + ct * 32 * TC + lane;
        // the accumulators start from the old entries of W: their (L2-latency) loads are in flight during the products
        T acc[TR][TC];
#pragma unroll
        for (int r = 0; r < TR; ++r)
#pragma unroll
            for (int c = 0; c < TC; ++c) {
                const int cc = cb + 32 * c;
                acc[r][c] = (i0 + r < r_end && cc < c1) ? W[(long)(i0 + r) * ld + cc] : zero_of(T());
            }
        // rows / columns past the end re-read the last valid one (their accumulators are never stored), so the
        // operand loads of the product loop are unconditional and address a running pointer
        int ro[TR], co[TC];
#pragma unroll
        for (int r = 0; r < TR; ++r) ro[r] = min(r, r_end - 1 - i0);
#pragma unroll
        for (int c = 0; c < TC; ++c) co[c] = min(cb + 32 * c, c1 - 1) - cb;
        const T* lp = Lp + (i0 - row0);
        const T* up = Us + (cb - c0);
        auto step = [&]() {
            T l[TR], u[TC];
#pragma unroll
            for (int r = 0; r < TR; ++r) l[r] = lp[ro[r]];
#pragma unroll
            for (int c = 0; c < TC; ++c) u[c] = up[co[c]];
#pragma unroll
            for (int r = 0; r < TR; ++r)
#pragma unroll
                for (int c = 0; c < TC; ++c) fnma(acc[r][c], l[r], u[c]);
            lp += lstride;
            up += CHUNK;
        };
        if (nbj == NB) {
#pragma unroll
            for (int q = 0; q < NB; ++q) step();
        } else {
            for (int q = 0; q < nbj; ++q) step();
        }
#pragma unroll
        for (int r = 0; r < TR; ++r) {
            if (i0 + r >= r_end) break;
#pragma unroll
            for (int c = 0; c < TC; ++c) {
                const int cc = cb + 32 * c;
                if (cc < c1) W[(long)(i0 + r) * ld + cc] = acc[r][c];
            }
        }
    }
}

// Forward elimination over pivot columns [0, npiv) of the nrows x ncols row-major matrix W
// (nrows >= npiv; columns >= npiv ride along as right-hand sides).  On return rows [0, npiv) hold
// U and the transformed right-hand sides; rows >= npiv hold the Schur remainder in columns >= npiv.
// status: bit 0 set on an exactly-zero pivot, bit 1 on a non-finite pivot column.
// smem: lu_smem_bytes<T,NB,CHUNK>(nrows).  All threads of the CTA must call.
// nsearch (default nrows): only rows < nsearch are pivot candidates; rows in [nsearch, nrows) are eliminated like every
// other row but never become a pivot row (the c rows of the window elimination, k_geom_band).
template <class T, int NB, bool SMALL = false, int CHUNK = kLuChunk>
__device__ void lu_forward(T* W, long ld, int nrows, int npiv, int ncols, unsigned char* smem_raw, int* status, int nsearch = -1) {
    T* panel = reinterpret_cast<T*>(smem_raw);               // [NB][nrows] column-major
    T* ust = panel + (size_t)NB * nrows;                     // [NB][CHUNK]
    int* piv = reinterpret_cast<int*>(ust + (size_t)NB * CHUNK);
    if (nsearch < 0) nsearch = nrows;
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5;
    for (int k0 = 0; k0 < npiv; k0 += NB) {
        const int nbj = min(NB, npiv - k0);
        const int rows = nrows - k0;
        const int srows = nsearch - k0;
        // stage the panel
        for (int idx = tid; idx < rows * nbj; idx += nthr) {
            const int i = idx / nbj, q = idx - i * nbj;
            panel[q * nrows + i] = W[(long)(k0 + i) * ld + k0 + q];
        }
        __syncthreads();
        if (warp == 0) {
            for (int j = 0; j < nbj; ++j) {
                double best = -1.0;
                int bi = 0x7fffffff;
                for (int i = j + lane; i < srows; i += 32) {
                    const double v = cabs1(panel[j * nrows + i]);
                    if (v > best) { best = v; bi = i; }
                }
#pragma unroll
                for (int off = 16; off >= 1; off >>= 1) {
                    const double ov = __shfl_down_sync(0xffffffffu, best, off);
                    const int oi = __shfl_down_sync(0xffffffffu, bi, off);
                    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
                }
                best = __shfl_sync(0xffffffffu, best, 0);
                bi = __shfl_sync(0xffffffffu, bi, 0);
                if (bi >= srows) bi = j;                      // column of NaNs
                if (lane == 0) {
                    piv[j] = k0 + bi;
                    if (!(best < INFINITY) || best < 0.0) atomicOr(status, 2);
                    else if (best == 0.0) atomicOr(status, 1);
                }
                if (bi != j && lane < nbj) {
                    const T t = panel[lane * nrows + j];
                    panel[lane * nrows + j] = panel[lane * nrows + bi];
                    panel[lane * nrows + bi] = t;
                }
                __syncwarp();
                const T pv = panel[j * nrows + j];
                for (int i = j + 1 + lane; i < rows; i += 32) {
                    const T l = panel[j * nrows + i] / pv;
                    panel[j * nrows + i] = l;
                    for (int q = j + 1; q < nbj; ++q) fnma(panel[q * nrows + i], l, panel[q * nrows + j]);
                }
                __syncwarp();
            }
        }
        __syncthreads();
        // U11 back to W (back-substitution reads it there)
        for (int idx = tid; idx < nbj * nbj; idx += nthr) {
            const int i = idx / nbj, q = idx - i * nbj;
            W[(long)(k0 + i) * ld + k0 + q] = panel[q * nrows + i];
        }
        const int cr0 = k0 + nbj;
        for (int cc0 = cr0; cc0 < ncols; cc0 += CHUNK) {
            const int cc1 = min(ncols, cc0 + CHUNK);
            // row swaps + U12 = L11^{-1} A12, one column per thread
            for (int c = cc0 + tid; c < cc1; c += nthr) {
                T u[NB];
#pragma unroll
                for (int j = 0; j < NB; ++j) {
                    if (j < nbj) {
                        const int r2 = piv[j];
                        if (r2 != k0 + j) {
                            T* p1 = &W[(long)(k0 + j) * ld + c];
                            T* p2 = &W[(long)r2 * ld + c];
                            const T t = *p1; *p1 = *p2; *p2 = t;
                        }
                    }
                }
#pragma unroll
                for (int q = 0; q < NB; ++q) {
                    if (q < nbj) {
                        T v = W[(long)(k0 + q) * ld + c];
#pragma unroll
                        for (int q2 = 0; q2 < NB; ++q2)
                            if (q2 < q) fnma(v, panel[q2 * nrows + q], u[q2]);
                        u[q] = v;
                        W[(long)(k0 + q) * ld + c] = v;
                        ust[q * CHUNK + (c - cc0)] = v;
                    }
                }
            }
            __syncthreads();
            lu_rank_update<T, NB, SMALL, CHUNK>(W, ld, k0 + nbj, nrows, k0, panel, nrows, cc0, cc1, ust, nbj);
            __syncthreads();
        }
    }
}

// Back-substitution U X = Y for right-hand-side columns [c0, c1) after lu_forward (U = rows/cols
// [0, npiv)); X overwrites Y.  Blocked bottom-up, reusing the rank-NB update for the rows above.
template <class T, int NB, bool SMALL = false, int CHUNK = kLuChunk>
__device__ void lu_backward(T* W, long ld, int npiv, int c0, int c1, unsigned char* smem_raw, int max_rows) {
    T* panel = reinterpret_cast<T*>(smem_raw);               // [NB][max_rows]: U[0:k0+nbj, k0:k0+nbj]
    T* ust = panel + (size_t)NB * max_rows;
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int nblk = (npiv + NB - 1) / NB;
    for (int blk = nblk - 1; blk >= 0; --blk) {
        const int k0 = blk * NB;
        const int nbj = min(NB, npiv - k0);
        const int rows = k0 + nbj;
        for (int idx = tid; idx < rows * nbj; idx += nthr) {
            const int i = idx / nbj, q = idx - i * nbj;
            panel[q * max_rows + i] = W[(long)i * ld + k0 + q];
        }
        __syncthreads();
        for (int cc0 = c0; cc0 < c1; cc0 += CHUNK) {
            const int cc1 = min(c1, cc0 + CHUNK);
            for (int c = cc0 + tid; c < cc1; c += nthr) {
                T x[NB];
#pragma unroll
                for (int q = NB - 1; q >= 0; --q) {
                    if (q < nbj) {
                        T v = W[(long)(k0 + q) * ld + c];
#pragma unroll
                        for (int q2 = NB - 1; q2 >= 0; --q2)
                            if (q2 > q && q2 < nbj) fnma(v, panel[q2 * max_rows + k0 + q], x[q2]);
                        v = v / panel[q * max_rows + k0 + q];
                        x[q] = v;
                        W[(long)(k0 + q) * ld + c] = v;
                        ust[q * CHUNK + (c - cc0)] = v;
                    }
                }
            }
            __syncthreads();
            lu_rank_update<T, NB, SMALL, CHUNK>(W, ld, 0, k0, 0, panel, max_rows, cc0, cc1, ust, nbj);
            __syncthreads();
        }
    }
}

#endif  // __CUDACC__
}  // namespace meem
