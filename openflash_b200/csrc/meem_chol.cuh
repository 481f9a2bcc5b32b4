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
        // only tiles that touch the lower triangle, dealt round-robin to the warps (a strided walk over the full ntr x ntc
        // grid with skips left the warps up to 20 % out of balance)
        int tcount = 0;
        for (int ti = 0; ti < ntr; ++ti) {
          const int ri0 = ti * 16;
          if (ri0 >= nrow) break;
          const int ntj = min(ntc, (ri0 + 15) / 32 + 1);
          for (int tj = 0; tj < ntj; ++tj) {
            if ((tcount++ % nwarp) != warp) continue;
            const int cj0 = tj * 32;
            double acc[2][4][2];
            // the lane's two row pointers once per tile (tri_off is ~10 integer instructions: 16 evaluations per tile before)
            double* rp[2];
            int rlim[2];                                  // columns c <= rlim are stored for that row (-1: row out of range)
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int r = ri0 + 8 * i + lr;
                rp[i] = P.row(kb + min(r, nrow - 1)) + kb;
                rlim[i] = (r < nrow) ? min(r, ncol - 1) : -1;
            }
            const int cbase = cj0 + 2 * lc;
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int c = cbase + 8 * j;
                    double2 v = make_double2(0.0, 0.0);
                    if (c <= rlim[i]) v = *reinterpret_cast<const double2*>(rp[i] + c);
                    acc[i][j][0] = v.x; acc[i][j][1] = v.y;
                }
            const double* pa = Pn + (ri0 + lr) * kCholLdP + lc;
            const double* pb = Pn + (cj0 + lr) * kCholLdP + lc;
#pragma unroll
            for (int ks = 0; ks < kCholNB / 4; ++ks) {
                double af[2], bf[4];
#pragma unroll
                for (int i = 0; i < 2; ++i) af[i] = -pa[8 * i * kCholLdP + ks * 4];
#pragma unroll
                for (int j = 0; j < 4; ++j) bf[j] = pb[8 * j * kCholLdP + ks * 4];
#pragma unroll
                for (int i = 0; i < 2; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) dmma(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
            }
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int c = cbase + 8 * j;
                    if (c <= rlim[i]) *reinterpret_cast<double2*>(rp[i] + c) = make_double2(acc[i][j][0], acc[i][j][1]);
                }
          }
        }
#ifdef MEEM_CHOL_PROF
        __syncthreads();
        CHOL_T(3);
#endif
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------------
// LEFT-LOOKING variant for a matrix that lives in L2 / HBM: block column j is formed as
//   C_j = P[:, j] - L[:, 0:j] L[j, 0:j]^T
// with the accumulators in registers over the whole contraction (the 16 pivot rows L[j, 0:j] staged in shared memory, the
// other operand streamed from L2 with independent, prefetchable loads), then factorised / solved and written ONCE.  The
// right-looking form above reads and writes the whole trailing triangle at every one of the K / 16 steps, each tile a
// dependent L2 round trip; this one reads K^3 / 96 doubles per system and writes the factor once.
// Shared memory: Dg | Li as above, then Lrow[16][ldr], ldr = chol_ll_ldr(K).  Same result layout as cta_cholesky.
__host__ __device__ inline int chol_ll_ldr(int K) { const int n8 = (K + 7) & ~7; return n8 + ((n8 & 15) ? 12 : 4); }
__host__ __device__ inline size_t chol_ll_smem_bytes(int K) {
    return sizeof(double) * ((size_t)kCholNB * (kCholNB + 1) + kCholNB + (size_t)kCholNB * kCholLdP + (size_t)kCholNB * chol_ll_ldr(K));
}

// C blocks (8 rows x 16 columns each) of NBLK row blocks row0 + u * rstride against the staged pivot rows:
// acc[u][j][e] = C_u[lr][8 j + 2 lc + e].  The NBLK operand streams are independent, so 4 NBLK L2 loads are in flight per
// lane (unroll 4); from_smem: the rows ARE the staged pivot rows (diagonal block), read them from shared memory.
template <int NBLK>
__device__ __forceinline__ void chol_ll_blocks(const TriMat& P, int row0, int rstride, int KR, int k0, int nb, bool diag,
                                               const double* Lrow, int ldr, int lr, int lc, double (&acc)[NBLK][2][2]) {
    const double* pa[NBLK];
    bool rok[NBLK];
#pragma unroll
    for (int u = 0; u < NBLK; ++u) {
        const int row = row0 + u * rstride + lr;
        rok[u] = row < KR;
        const double* rp = P.row(rok[u] ? row : KR - 1);
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int c = 8 * j + 2 * lc + e;
                const bool ok = rok[u] && c < nb && (!diag || k0 + c <= row);
                acc[u][j][e] = ok ? rp[k0 + c] : 0.0;
            }
        pa[u] = diag ? Lrow + (row - k0) * ldr + lc : rp + lc;
        if (diag && !rok[u]) pa[u] = Lrow + lc;
    }
    const double* pb = Lrow + lr * ldr + lc;
#pragma unroll 4
    for (int ks = 0; ks < (k0 >> 2); ++ks) {
        double a[NBLK];
#pragma unroll
        for (int u = 0; u < NBLK; ++u) a[u] = rok[u] ? -pa[u][4 * ks] : 0.0;
        const double b0 = pb[4 * ks], b1 = pb[8 * ldr + 4 * ks];
#pragma unroll
        for (int u = 0; u < NBLK; ++u) {
            dmma(acc[u][0][0], acc[u][0][1], a[u], b0);
            dmma(acc[u][1][0], acc[u][1][1], a[u], b1);
        }
    }
}

// L21 block = C block Linv^T (the C fragments re-laid out by shuffles), stored once
__device__ __forceinline__ void chol_ll_solve_store(const TriMat& P, int row0, int KR, int k0, int nb, const double (&bf)[kCholNB / 4][2],
                                                    int lane, const double (&acc)[2][2]) {
    const int lr = lane >> 2, lc = lane & 3;
    double out[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
#pragma unroll
    for (int ks = 0; ks < kCholNB / 4; ++ks) {
        // A fragment a = C[lr][4 ks + lc] sits in lane (lr, 2 (ks & 1) + (lc >> 1)) as acc[ks >> 1][lc & 1]
        const int src = (lane & ~3) | (2 * (ks & 1) + (lc >> 1));
        const double v0 = __shfl_sync(0xffffffffu, acc[ks >> 1][0], src);
        const double v1 = __shfl_sync(0xffffffffu, acc[ks >> 1][1], src);
        const double a = (lc & 1) ? v1 : v0;
#pragma unroll
        for (int j = 0; j < 2; ++j) dmma(out[j][0], out[j][1], a, bf[ks][j]);
    }
    const int row = row0 + lr;
    if (row < KR) {
        double* rp = P.row(row) + k0;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int c = 8 * j + 2 * lc;
            if (c + 1 < nb) *reinterpret_cast<double2*>(rp + c) = make_double2(out[j][0], out[j][1]);
            else if (c < nb) rp[c] = out[j][0];
        }
    }
}

__device__ void cta_cholesky_ll(const TriMat P, int naug, double* sm, int* fail) {
    const int K = P.K;
    double* Dg = sm;
    double* Li = Dg + kCholNB * (kCholNB + 1) + kCholNB;
    double* Lrow = Li + kCholNB * kCholLdP;
    const int ldr = chol_ll_ldr(K);
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    const int lr = lane >> 2, lc = lane & 3;
    const int KR = K + naug;
    for (int k0 = 0; k0 < K; k0 += kCholNB) {
        const int nb = min(kCholNB, K - k0);
        const int kb = k0 + nb;
        __syncthreads();                                   // previous column's readers of Lrow / Li are done
        // pivot rows L[k0 .. k0+15][0 .. k0) -> shared memory (rows beyond nb: zeros)
        for (int idx = tid; idx < kCholNB * (k0 >> 1); idx += nthr) {
            const int i = idx / (k0 >> 1), g = idx - i * (k0 >> 1);
            double2 v = make_double2(0.0, 0.0);
            if (i < nb) v = *reinterpret_cast<const double2*>(P.row(k0 + i) + 2 * g);
            *reinterpret_cast<double2*>(Lrow + i * ldr + 2 * g) = v;
        }
        __syncthreads();
        // diagonal block: row blocks 0, 1 of the block column (lower triangle), identity-padded into Dg
        for (int w = warp; w < 2; w += nwarp) {
            double acc[1][2][2];
            chol_ll_blocks<1>(P, k0 + 8 * w, 0, k0 + nb, k0, nb, true, Lrow, ldr, lr, lc, acc);
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int i = 8 * w + lr, c = 8 * j + 2 * lc + e;
                    Dg[i * (kCholNB + 1) + c] = (i < nb && c < nb) ? acc[0][j][e] : (i == c ? 1.0 : 0.0);
                }
        }
        __syncthreads();
        // warp 0 factorises the diagonal block; meanwhile the other warps already form their first two C blocks each (these
        // need the staged pivot rows only, not the factor) and hold them in registers across the barrier
        const int nrow = KR - kb;
        const int nrb = nrow > 0 ? (nrow + 7) >> 3 : 0;
        const int nhelp = nwarp - 1;
        double hacc[2][2][2];
        if (warp == 0) warp_chol16(Dg, Li, fail);
        else if (warp - 1 < nrb) chol_ll_blocks<2>(P, kb + 8 * (warp - 1), 8 * nhelp, KR, k0, nb, false, Lrow, ldr, lr, lc, hacc);
        __syncthreads();
        for (int idx = tid; idx < kCholNB * kCholNB; idx += nthr) {      // the inverse of the factor block replaces the block
            const int i = idx >> 4, c = idx & 15;
            if (i < nb && c <= i) P.row(k0 + i)[k0 + c] = Li[i * kCholLdP + c];
        }
        if (nrow <= 0) break;
        // rows below: C block, then L21 = C Linv^T with the C fragments re-laid out by shuffles, stored once
        double bf[kCholNB / 4][2];
#pragma unroll
        for (int ks = 0; ks < kCholNB / 4; ++ks)
#pragma unroll
            for (int j = 0; j < 2; ++j) bf[ks][j] = Li[(8 * j + lr) * kCholLdP + ks * 4 + lc];
        if (warp > 0 && warp - 1 < nrb) {
            chol_ll_solve_store(P, kb + 8 * (warp - 1), KR, k0, nb, bf, lane, hacc[0]);
            if (warp - 1 + nhelp < nrb) chol_ll_solve_store(P, kb + 8 * (warp - 1 + nhelp), KR, k0, nb, bf, lane, hacc[1]);
        }
        const int rb_first = min(nrb, 2 * nhelp);            // row blocks [0, rb_first) are done
        // two row blocks in flight per warp while whole pairs remain (8 independent L2 loads per lane), then singles
        int rb = rb_first + warp;
        for (; rb + nwarp < nrb; rb += 2 * nwarp) {
            double acc[2][2][2];
            chol_ll_blocks<2>(P, kb + 8 * rb, 8 * nwarp, KR, k0, nb, false, Lrow, ldr, lr, lc, acc);
            chol_ll_solve_store(P, kb + 8 * rb, KR, k0, nb, bf, lane, acc[0]);
            chol_ll_solve_store(P, kb + 8 * (rb + nwarp), KR, k0, nb, bf, lane, acc[1]);
        }
        if (rb < nrb) {
            double acc[1][2][2];
            chol_ll_blocks<1>(P, kb + 8 * rb, 0, KR, k0, nb, false, Lrow, ldr, lr, lc, acc);
            chol_ll_solve_store(P, kb + 8 * rb, KR, k0, nb, bf, lane, acc[0]);
        }
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
