// CTA-cooperative blocked Cholesky of the per-frequency Schur matrix, FP64, 256 threads.
//
// P = -Re(S) = Imk^T H Imk - Re(D) is symmetric positive definite for every geometry probed
// (tools/experiments/exp_spd.py; physically it is the quadratic form of the exterior Dirichlet-to-
// Neumann map), so the K x K frequency system needs no pivoting: P = L L^T, K^3/3 flops instead of
// the 2K^3/3 of a pivoted LU and no pivot search on the critical path.  A non-positive or non-finite
// pivot raises a flag and the system is re-solved by the pivoted-LU kernel (k_freq_lu).
//
// The matrix lives row-major in the CTA's L2-resident slot (lower triangle, leading dimension ld, a
// multiple of 8); `naug` extra rows below row K carry right-hand sides, so the forward substitution
// L^{-1} r falls out of the panel solves.  Right-looking with panels of kCholNB = 16 columns:
//   diagonal block  one warp, in shared memory
//   panel solve     one thread per row against the 16 x 16 factor (broadcast from shared memory)
//   trailing update P22 -= L21 L21^T on the DMMA pipe, 32 x 32 warp tiles of the lower triangle,
//                   both operands read from the shared-memory copy of the panel (row stride 20)
#pragma once
#include "meem_dmma.cuh"

namespace meem {

constexpr int kCholNB = 16;
constexpr int kCholLdP = kCholNB + 4;   // 20 = 4 (mod 16): conflict-free fragment loads

__host__ __device__ inline size_t chol_smem_bytes(int K, int naug) {
    const int rows = ((K + naug + 31) / 32) * 32;
    return sizeof(double) * ((size_t)kCholNB * (kCholNB + 1) + kCholNB + (size_t)rows * kCholLdP);
}

// returns through *fail (shared int, OR-ed): bit 0 non-positive pivot, bit 1 non-finite pivot
__device__ void cta_cholesky(double* P, int ld, int K, int naug, double* sm, int* fail) {
    double* Dg = sm;                                   // [16][17]
    double* rd = Dg + kCholNB * (kCholNB + 1);         // reciprocal diagonal of the current block
    double* Pn = rd + kCholNB;                         // [rows][20] current panel L21
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
            Dg[i * (kCholNB + 1) + c] = (i < nb && c <= i) ? P[(long)(k0 + i) * ld + k0 + c] : (i == c ? 1.0 : 0.0);
        }
        __syncthreads();
        if (warp == 0) {
            for (int j = 0; j < nb; ++j) {
                const double d = Dg[j * (kCholNB + 1) + j];
                if (lane == 0) {
                    if (!(d == d) || fabs(d) == INFINITY) atomicOr(fail, 2);
                    else if (!(d > 0.0)) atomicOr(fail, 1);
                }
                const double s = sqrt(d);
                const double rinv = 1.0 / s;
                double lij = 0.0;
                if (lane > j && lane < nb) {
                    lij = Dg[lane * (kCholNB + 1) + j] * rinv;
                    Dg[lane * (kCholNB + 1) + j] = lij;
                }
                if (lane == j) { Dg[j * (kCholNB + 1) + j] = s; rd[j] = rinv; }
                __syncwarp();
                if (lane > j && lane < nb) {
                    for (int c = j + 1; c <= lane; ++c)
                        Dg[lane * (kCholNB + 1) + c] = fma(-lij, Dg[c * (kCholNB + 1) + j], Dg[lane * (kCholNB + 1) + c]);
                }
                __syncwarp();
            }
            if (lane >= nb && lane < kCholNB) rd[lane] = 1.0;
        }
        __syncthreads();
        // factor block back to the slot (back-substitution reads it there)
        for (int idx = tid; idx < nb * nb; idx += nthr) {
            const int i = idx / nb, c = idx - i * nb;
            if (c <= i) P[(long)(k0 + i) * ld + k0 + c] = Dg[i * (kCholNB + 1) + c];
        }
        const int nrow = KR - kb;
        if (nrow <= 0) break;
        const int ncol = K - kb;
        const int prow = ((nrow + 31) / 32) * 32;
        // panel solve: L21 = A21 L11^{-T}
        for (int ri = tid; ri < prow; ri += nthr) {
            double l[kCholNB];
            if (ri < nrow) {
                double* src = P + (long)(kb + ri) * ld + k0;
#pragma unroll
                for (int c = 0; c < kCholNB; c += 2) {
                    const double2 v = *reinterpret_cast<const double2*>(src + c);
                    l[c] = v.x; l[c + 1] = v.y;
                }
#pragma unroll
                for (int c = 0; c < kCholNB; ++c) {
                    double a = (c < nb) ? l[c] : 0.0;
#pragma unroll
                    for (int q = 0; q < kCholNB; ++q)
                        if (q < c) a = fma(-l[q], Dg[c * (kCholNB + 1) + q], a);
                    l[c] = a * rd[c];
                }
                for (int c = 0; c < nb; ++c) src[c] = l[c];
            } else {
#pragma unroll
                for (int c = 0; c < kCholNB; ++c) l[c] = 0.0;
            }
#pragma unroll
            for (int c = 0; c < kCholNB; c += 2)
                *reinterpret_cast<double2*>(Pn + ri * kCholLdP + c) = make_double2(l[c], l[c + 1]);
        }
        __syncthreads();
        if (ncol <= 0) continue;
        // trailing update on the DMMA pipe
        const int ntr = prow / 32, ntc = (ncol + 31) / 32;
        for (int t = warp; t < ntr * ntc; t += nwarp) {
            const int ti = t / ntc, tj = t - ti * ntc;
            const int ri0 = ti * 32, cj0 = tj * 32;
            if (cj0 > ri0 + 31) continue;
            // unmasked 4 x 4 blocks (straight-line DMMAs); panel rows beyond nrow are zero, stores are bounds-checked
            double acc[4][4][2];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
#pragma unroll
            for (int ks = 0; ks < kCholNB / 4; ++ks) {
                double af[4], bf[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) af[i] = Pn[(ri0 + 8 * i + lr) * kCholLdP + ks * 4 + lc];
#pragma unroll
                for (int j = 0; j < 4; ++j) bf[j] = Pn[(cj0 + 8 * j + lr) * kCholLdP + ks * 4 + lc];
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) dmma(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int r = ri0 + 8 * i + lr, c = cj0 + 8 * j + 2 * lc;
                    if (r < nrow && c < ncol && (ri0 + 8 * i + 7 >= cj0 + 8 * j)) {
                        double2* p = reinterpret_cast<double2*>(P + (long)(kb + r) * ld + kb + c);
                        double2 v = *p;
                        v.x -= acc[i][j][0]; v.y -= acc[i][j][1];
                        *p = v;
                    }
                }
        }
    }
    __syncthreads();
}

// Back-substitution L^T X = C for nrhs right-hand sides held as rows of X ([nrhs][ldx], global or
// shared).  Blocked bottom-up; one warp per right-hand side solves the 16 x 16 diagonal block.
__device__ void cta_chol_backsolve(const double* P, int ld, int K, double* X, int ldx, int nrhs, double* sm) {
    double* Dg = sm;                                   // [16][17]
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    const int nblk = (K + kCholNB - 1) / kCholNB;
    for (int blk = nblk - 1; blk >= 0; --blk) {
        const int k0 = blk * kCholNB;
        const int nb = min(kCholNB, K - k0);
        __syncthreads();
        for (int idx = tid; idx < kCholNB * kCholNB; idx += nthr) {
            const int i = idx >> 4, c = idx & 15;
            Dg[i * (kCholNB + 1) + c] = (i < nb && c <= i) ? P[(long)(k0 + i) * ld + k0 + c] : (i == c ? 1.0 : 0.0);
        }
        __syncthreads();
        for (int r = warp; r < nrhs; r += nwarp) {
            double c = (lane < nb) ? X[(long)r * ldx + k0 + lane] : 0.0;
            for (int q = nb - 1; q >= 0; --q) {
                double x = 0.0;
                if (lane == q) x = c / Dg[q * (kCholNB + 1) + q];
                x = __shfl_sync(0xffffffffu, x, q);
                if (lane == q) c = x;
                else if (lane < q) c = fma(-Dg[q * (kCholNB + 1) + lane], x, c);
            }
            if (lane < nb) X[(long)r * ldx + k0 + lane] = c;
        }
        __syncthreads();
        for (int idx = tid; idx < nrhs * k0; idx += nthr) {
            const int r = idx / k0, j = idx - r * k0;
            double acc = X[(long)r * ldx + j];
            for (int q = 0; q < nb; ++q) acc = fma(-P[(long)(k0 + q) * ld + j], X[(long)r * ldx + k0 + q], acc);
            X[(long)r * ldx + j] = acc;
        }
    }
    __syncthreads();
}

}  // namespace meem
