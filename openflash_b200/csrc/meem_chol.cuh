// CTA-cooperative blocked Cholesky of the per-frequency Schur matrix, FP64, 256 threads.
//
// P = -Re(S) = Imk^T H Imk - Re(D) is symmetric positive definite for every geometry probed
// (tools/experiments/exp_spd.py; physically it is the quadratic form of the exterior Dirichlet-to-
// Neumann map), so the K x K frequency system needs no pivoting: P = L L^T, K^3/3 flops instead of
// the 2K^3/3 of a pivoted LU and no pivot search on the critical path.  A non-positive or non-finite
// pivot raises a flag and the system is re-solved by the pivoted-LU kernel (k_freq_lu).
//
// The matrix is the PACKED lower triangle (row r holds columns 0..r, rows padded to an even length, tri_off(r)
// doubles before it: 162 KB for K = 200), either in the system's L2-resident buffer or - when it fits - in shared
// memory, where the whole factorisation then runs (k_freq_solve<.., true>).  `naug` extra full rows (leading
// dimension lda) carry right-hand sides, so the forward substitution L^{-1} r falls out of the panel solves.
// Right-looking with panels of kCholNB = 16 columns:
//   diagonal block  one warp, in shared memory
//   panel solve     one thread per row against the 16 x 16 factor (broadcast from shared memory)
//   trailing update P22 -= L21 L21^T on the DMMA pipe, 32 x 32 warp tiles of the lower triangle,
//                   both operands read from the shared-memory copy of the panel (row stride 20)
#pragma once
#include "meem_dmma.cuh"

namespace meem {

constexpr int kCholNB = 16;
constexpr int kCholLdP = kCholNB + 4;   // 20 = 4 (mod 16): conflict-free fragment loads

// packed lower triangle: offset of row r (rows padded to even length so that double2 accesses stay aligned)
__host__ __device__ inline int tri_off(int r) { const int h = r >> 1; return 2 * h * (h + 1) + ((r & 1) ? 2 * (h + 1) : 0); }
__host__ __device__ inline int tri_doubles(int K) { return (tri_off(K) + 7) & ~7; }
struct TriMat {
    double* p;      // packed triangle
    double* aug;    // augmented rows K.., row-major with leading dimension lda
    int K, lda;
    __device__ __forceinline__ double* row(int r) const { return r < K ? p + tri_off(r) : aug + (long)(r - K) * lda; }
};

__host__ __device__ inline size_t chol_smem_bytes(int K, int naug) {
    const int rows = ((K + naug + 31) / 32) * 32;
    return sizeof(double) * ((size_t)kCholNB * (kCholNB + 1) + kCholNB + (size_t)kCholNB * kCholLdP + (size_t)rows * kCholLdP);
}

// Cholesky of one 16 x 16 diagonal block by one warp, rows in registers (lane i owns row i), then its inverse.
// Dg (shared, [16][17]) holds the block on entry (identity-padded beyond nb); on exit Dg = L11 with the RECIPROCAL
// diagonal, Li (shared, [16][20]) = L11^{-1} (row-major, zero above the diagonal).
__device__ __forceinline__ void warp_chol16(double* Dg, double* Li, int* fail) {
    const int lane = threadIdx.x & 31;
    const int row = lane & 15;
    double r[kCholNB];
#pragma unroll
    for (int c = 0; c < kCholNB; ++c) r[c] = Dg[row * (kCholNB + 1) + c];
    int bad = 0;
    // right-looking, every update unconditional (entries above the diagonal hold garbage that is never read): keeps
    // r[] in registers - a predicated update made the compiler spill the row to local memory
#pragma unroll
    for (int j = 0; j < kCholNB; ++j) {
        const double d = __shfl_sync(0xffffffffu, r[j], j);
        if (!(d > 0.0) || d == INFINITY) bad |= (d == d && fabs(d) != INFINITY) ? 1 : 2;
        const double rinv = rsqrt(d);
        const double lij = r[j] * rinv;               // lane j: sqrt(d)
        r[j] = (row == j) ? rinv : lij;               // reciprocal diagonal
#pragma unroll
        for (int c = j + 1; c < kCholNB; ++c) {
            const double lcj = __shfl_sync(0xffffffffu, lij, c);
            r[c] = fma(-lij, lcj, r[c]);
        }
    }
    if (bad && lane == 0) atomicOr(fail, bad);
    if (lane < kCholNB) {
#pragma unroll
        for (int c = 0; c < kCholNB; ++c) Dg[row * (kCholNB + 1) + c] = (c <= row) ? r[c] : 0.0;
    }
    __syncwarp();
    // inverse, column `row` per lane, right-looking so that the dependent chain is one multiply + one FMA per row:
    // x_q = acc_q / L_qq, then acc_i -= L_iq x_q for every i > q
    {
        double acc[kCholNB];
#pragma unroll
        for (int i = 0; i < kCholNB; ++i) acc[i] = (i == row) ? 1.0 : 0.0;
#pragma unroll
        for (int q = 0; q < kCholNB; ++q) {
            const double xq = acc[q] * Dg[q * (kCholNB + 1) + q];
            if (lane < kCholNB) Li[q * kCholLdP + row] = xq;                     // Linv[q][c]
#pragma unroll
            for (int i = q + 1; i < kCholNB; ++i) acc[i] = fma(-Dg[i * (kCholNB + 1) + q], xq, acc[i]);
        }
    }
}

// returns through *fail (shared int, OR-ed): bit 0 non-positive pivot, bit 1 non-finite pivot.
// On exit the strictly-below-block-diagonal part of P holds L, each 16 x 16 diagonal block holds the INVERSE of its
// factor (lower triangular; its diagonal is 1 / L_ii) - the back-substitution only ever needs the inverse - and rows
// K.. hold (L^{-1} r_i)^T.
#ifdef MEEM_CHOL_PROF
#define CHOL_T(i) do { const long long tn_ = clock64(); if (tsub) tsub[i] += tn_ - tprev_; tprev_ = tn_; } while (0)
#else
#define CHOL_T(i) do { } while (0)
#endif
// KEEP_FACTOR: the diagonal blocks keep the factor itself (with the reciprocal diagonal) instead of its inverse (k_hchol).
template <bool KEEP_FACTOR = false>
__device__ void cta_cholesky(const TriMat P, int naug, double* sm, int* fail, long long* tsub = nullptr) {
    const int K = P.K;
#ifdef MEEM_CHOL_PROF
    long long tprev_ = clock64();
#endif
    double* Dg = sm;                                   // [16][17]
    double* Li = Dg + kCholNB * (kCholNB + 1) + kCholNB;   // [16][20] inverse of the current diagonal block
    double* Pn = Li + kCholNB * kCholLdP;              // [rows][20] current panel L21
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    const int lr = lane >> 2, lc = lane & 3;
    const int KR = K + naug;
    for (int k0 = 0; k0 < K; k0 += kCholNB) {
        const int nb = min(kCholNB, K - k0);
        const int kb = k0 + nb;
        __syncthreads();
        for (int idx = tid; idx < kCholNB * kCholNB; idx += nthr) {
            const int i = idx >> 4, c = idx & 15;
            Dg[i * (kCholNB + 1) + c] = (i < nb && c <= i) ? P.row(k0 + i)[k0 + c] : (i == c ? 1.0 : 0.0);
        }
        __syncthreads();
        CHOL_T(0);
        if (warp == 0) warp_chol16(Dg, Li, fail);
        __syncthreads();
        CHOL_T(1);
        // the inverse of the factor block replaces the block
        for (int idx = tid; idx < kCholNB * kCholNB; idx += nthr) {
            const int i = idx >> 4, c = idx & 15;
            if (i < nb && c <= i) P.row(k0 + i)[k0 + c] = KEEP_FACTOR ? Dg[i * (kCholNB + 1) + c] : Li[i * kCholLdP + c];
        }
        const int nrow = KR - kb;
        if (nrow <= 0) break;
        const int ncol = K - kb;
        const int prow = ((nrow + 31) / 32) * 32;
        // panel solve on the DMMA pipe: L21 = A21 Linv^T, A fragments straight from L2
        {
            double bf[kCholNB / 4][2];
#pragma unroll
            for (int ks = 0; ks < kCholNB / 4; ++ks)
#pragma unroll
                for (int j = 0; j < 2; ++j) bf[ks][j] = Li[(8 * j + lr) * kCholLdP + ks * 4 + lc];   // b[k=q][n=c] = Linv[c][q]
            for (int rb = warp; rb < prow / 8; rb += nwarp) {
                const int ri = rb * 8 + lr;
                double* src = P.row(kb + min(ri, nrow - 1)) + k0;
                double af[kCholNB / 4];
#pragma unroll
                for (int ks = 0; ks < kCholNB / 4; ++ks) af[ks] = (ri < nrow && ks * 4 + lc < nb) ? src[ks * 4 + lc] : 0.0;
                double c[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
#pragma unroll
                for (int ks = 0; ks < kCholNB / 4; ++ks)
#pragma unroll
                    for (int j = 0; j < 2; ++j) dmma(c[j][0], c[j][1], af[ks], bf[ks][j]);
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const int col = 8 * j + 2 * lc;
                    *reinterpret_cast<double2*>(Pn + ri * kCholLdP + col) = make_double2(c[j][0], c[j][1]);
                    if (ri < nrow && col < nb) {
                        double* dst = src + col;
                        if (col + 1 < nb) *reinterpret_cast<double2*>(dst) = make_double2(c[j][0], c[j][1]);
                        else dst[0] = c[j][0];
                    }
                }
            }
        }
        __syncthreads();
        CHOL_T(2);
        if (ncol <= 0) continue;
        // trailing update on the DMMA pipe: 16 x 32 warp tiles of the lower triangle, C tile is the accumulator
        const int ntr = prow / 16, ntc = (ncol + 31) / 32;
        for (int t = warp; t < ntr * ntc; t += nwarp) {
            const int ti = t / ntc, tj = t - ti * ntc;
            const int ri0 = ti * 16, cj0 = tj * 32;
            if (cj0 > ri0 + 15 || ri0 >= nrow) continue;
            double acc[2][4][2];
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int r = ri0 + 8 * i + lr, c = cj0 + 8 * j + 2 * lc;
                    double2 v = make_double2(0.0, 0.0);
                    if (r < nrow && c <= r && c < ncol)
                        v = *reinterpret_cast<const double2*>(P.row(kb + r) + kb + c);
                    acc[i][j][0] = v.x; acc[i][j][1] = v.y;
                }
#pragma unroll
            for (int ks = 0; ks < kCholNB / 4; ++ks) {
                double af[2], bf[4];
#pragma unroll
                for (int i = 0; i < 2; ++i) af[i] = -Pn[(ri0 + 8 * i + lr) * kCholLdP + ks * 4 + lc];
#pragma unroll
                for (int j = 0; j < 4; ++j) bf[j] = Pn[(cj0 + 8 * j + lr) * kCholLdP + ks * 4 + lc];
#pragma unroll
                for (int i = 0; i < 2; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) dmma(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
            }
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int r = ri0 + 8 * i + lr, c = cj0 + 8 * j + 2 * lc;
                    if (r < nrow && c <= r && c < ncol)
                        *reinterpret_cast<double2*>(P.row(kb + r) + kb + c) = make_double2(acc[i][j][0], acc[i][j][1]);
                }
        }
#ifdef MEEM_CHOL_PROF
        __syncthreads();
        CHOL_T(3);
#endif
    }
    __syncthreads();
}

// Back-substitution L^T X = C for nrhs right-hand sides held as rows of X ([nrhs][ldx]).  Blocked bottom-up with
// the inverted diagonal blocks: x_b = Linv_bb^T c_b is a 16 x 16 mat-vec (no serial chain), then the rows above
// are updated.  nrhs <= 8 * blockDim.x / 16.
__device__ void cta_chol_backsolve(const TriMat P, double* X, int ldx, int nrhs) {
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int K = P.K;
    const int nblk = (K + kCholNB - 1) / kCholNB;
    for (int blk = nblk - 1; blk >= 0; --blk) {
        const int k0 = blk * kCholNB;
        const int nb = min(kCholNB, K - k0);
        __syncthreads();
        double xr[8];                                   // up to 8 passes of nthr / 16 right-hand sides
        const int c = tid & 15, rstep = nthr >> 4;
#pragma unroll
        for (int ps = 0; ps < 8; ++ps) {
            const int r = (tid >> 4) + ps * rstep;
            double x = 0.0;
            if (r < nrhs && c < nb)
                for (int q = c; q < nb; ++q) x = fma(P.row(k0 + q)[k0 + c], X[(long)r * ldx + k0 + q], x);
            xr[ps] = x;
        }
        __syncthreads();
#pragma unroll
        for (int ps = 0; ps < 8; ++ps) {
            const int r = (tid >> 4) + ps * rstep;
            if (r < nrhs && c < nb) X[(long)r * ldx + k0 + c] = xr[ps];
        }
        __syncthreads();
        for (int idx = tid; idx < nrhs * k0; idx += nthr) {
            const int rr = idx / k0, j = idx - rr * k0;
            double acc = X[(long)rr * ldx + j];
            for (int q = 0; q < nb; ++q) acc = fma(-P.row(k0 + q)[j], X[(long)rr * ldx + k0 + q], acc);
            X[(long)rr * ldx + j] = acc;
        }
    }
    __syncthreads();
}

}  // namespace meem
