// CTA-level FP64 building blocks on the DMMA pipe (mma.sync.aligned.m8n8k4.f64) for sm_100a.
//
// Blackwell's tcgen05 / TMEM path has no FP64 kind, so the dense FP64 contractions of the MEEM
// frequency stage run on the legacy tensor pipe: one m8n8k4 DMMA = 256 FMAs per warp instruction,
// measured on B200 at the same peak as a pure DFMA stream but with 8x fewer issue slots and 1/8 of
// the operand traffic from shared memory.  Operands are staged with cp.async (LDGSTS) into padded
// shared-memory tiles whose row stride is 4 (mod 16) doubles, which makes every fragment load
// (8 rows x 4 columns of doubles per half-warp) bank-conflict free.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace meem {

// D(8x8) += A(8x4, row) * B(4x8, col).  Lane l holds a = A[l/4][l%4], b = B[l%4][l/4],
// c0/c1 = C[l/4][2*(l%4) + {0,1}].
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// 16-byte async copy global -> shared; src_bytes = 0 zero-fills the destination.
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, int src_bytes) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gsrc), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }

// ------------------------------------------------------------------------------------------------
// C[m][n] = sum_k A[k][m] * B[k][n]   ("TN": both operands stored contraction-major, row-major with
// leading dimensions lda / ldb, which must be even and the base pointers 16-byte aligned).
// 256 threads = 8 warps as a 2 x 4 grid of 32 x 32 warp tiles -> CTA tile 64 x 128; contraction is
// consumed in chunks of kGemmKC rows through a kGemmStages-deep cp.async ring.  Columns of A / B beyond
// Mm / Nn (up to the even-padded leading dimension) may hold anything finite or not: they only feed
// output rows / columns that are never stored.  Rows beyond Kd are zero-filled.
// lower_only: skip 8x8 blocks that lie strictly above the diagonal (m < n for every element).
// epi(m, n, v0, v1): called once per lane per stored 8x8 block with the two adjacent columns
// n, n+1 of row m (n is even); bounds (m < Mm, n < Nn, n+1 < Nn) are the callee's to check.
constexpr int kGemmKC = 16;
constexpr int kGemmStages = 3;          // cp.async ring: one barrier per chunk (the stage refilled at chunk c is the one read at c - 1)
constexpr int kGemmTM = 64, kGemmTN = 128;
constexpr int kGemmLdA = kGemmTM + 4, kGemmLdB = kGemmTN + 4;
constexpr int kGemmStageDoubles = kGemmKC * (kGemmLdA + kGemmLdB);
constexpr int kGemmSmemDoubles = 12800;                                  // dynamic shared memory of k_freq_gemm (2 CTAs / SM)
constexpr size_t kGemmSmemBytes = kGemmSmemDoubles * sizeof(double);
static_assert(kGemmStages * kGemmStageDoubles <= kGemmSmemDoubles, "cp.async ring does not fit");

// Thread t copies 16-byte group (t & 31) of A rows (t >> 5) + 8 it and group (t & 63) of B rows (t >> 6) + 4 it, so the
// per-thread source pointers / column predicates are chunk-invariant (GemmStageMap, built once per CTA tile) and a chunk
// only adds k0 rows: ~4 instructions per cp.async instead of ~30 of index arithmetic.
struct GemmStageMap {
    const double* pa;     // A + (t >> 5) * lda + m0 + 2 (t & 31)
    const double* pb;     // B + (t >> 6) * ldb + n0 + 2 (t & 63)
    unsigned sa, sb;      // shared-memory byte offsets of the two destinations inside a stage
    bool oka, okb;        // column inside the matrix
};
__device__ __forceinline__ GemmStageMap gemm_stage_map(const double* A, int lda, const double* B, int ldb,
                                                       int m0, int n0, int Mlim, int Nlim) {
    const int tid = threadIdx.x;
    GemmStageMap mp;
    const int ra = tid >> 5, ga = tid & 31, rb = tid >> 6, gb = tid & 63;
    mp.oka = (m0 + 2 * ga) < Mlim;
    mp.okb = (n0 + 2 * gb) < Nlim;
    mp.pa = mp.oka ? A + (long)ra * lda + m0 + 2 * ga : A;
    mp.pb = mp.okb ? B + (long)rb * ldb + n0 + 2 * gb : B;
    mp.sa = (unsigned)((ra * kGemmLdA + 2 * ga) * sizeof(double));
    mp.sb = (unsigned)((kGemmKC * kGemmLdA + rb * kGemmLdB + 2 * gb) * sizeof(double));
    return mp;
}
__device__ __forceinline__ void cp_async16s(unsigned smem_dst, const void* gsrc, int src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_dst), "l"(gsrc), "r"(src_bytes));
}
__device__ __forceinline__ void gemm_stage_load(double* st, const GemmStageMap& mp, int lda, int ldb, int k0, int Kd) {
    const unsigned s0 = (unsigned)__cvta_generic_to_shared(st);
    const int ra = threadIdx.x >> 5, rb = threadIdx.x >> 6;
    const double* pa = mp.pa + (long)k0 * lda;
    const double* pb = mp.pb + (long)k0 * ldb;
#pragma unroll
    for (int it = 0; it < kGemmKC / 8; ++it) {        // rows (t >> 5) + 8 it
        const bool ok = mp.oka && (k0 + ra + 8 * it < Kd);
        cp_async16s(s0 + mp.sa + it * 8 * kGemmLdA * (unsigned)sizeof(double), ok ? pa + (long)it * 8 * lda : pa, ok ? 16 : 0);
    }
#pragma unroll
    for (int it = 0; it < kGemmKC / 4; ++it) {        // rows (t >> 6) + 4 it
        const bool ok = mp.okb && (k0 + rb + 4 * it < Kd);
        cp_async16s(s0 + mp.sb + it * 4 * kGemmLdB * (unsigned)sizeof(double), ok ? pb + (long)it * 4 * ldb : pb, ok ? 16 : 0);
    }
}

// One contraction chunk of a warp tile of NI x NJ 8x8 blocks.  NI / NJ are compile-time so that the DMMAs sit in
// straight-line code: a runtime block mask around mma.sync costs a WARPSYNC + NOP + branch per DMMA (measured:
// 1.7x the DMMA pipe time of the unmasked loop).
template <int NI, int NJ>
__device__ __forceinline__ void gemm_chunk(const double* As, const double* Bs, int lc, double (&acc)[4][4][2]) {
#pragma unroll
    for (int ks = 0; ks < kGemmKC / 4; ++ks) {
        double af[NI], bf[NJ];
#pragma unroll
        for (int i = 0; i < NI; ++i) af[i] = As[(ks * 4 + lc) * kGemmLdA + i * 8];
#pragma unroll
        for (int j = 0; j < NJ; ++j) bf[j] = Bs[(ks * 4 + lc) * kGemmLdB + j * 8];
#pragma unroll
        for (int i = 0; i < NI; ++i)
#pragma unroll
            for (int j = 0; j < NJ; ++j) dmma(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
    }
}

template <class Epi>
// tri >= 0: A is "upper trapezoidal from column tri on" - A[k][m] = 0 for k < m - tri - so a tile whose first output row
// is m0 starts its contraction at the chunk containing m0 - tri.
__device__ void cta_gemm_dmma(const double* __restrict__ A, int lda, const double* __restrict__ B, int ldb,
                              int Mm, int Nn, int Kd, bool lower_only, int tri, double* sm, Epi epi) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int wr = warp >> 2, wc = warp & 3;            // 2 x 4 warps
    const int lr = lane >> 2, lc = lane & 3;
    const int Mlim = (Mm + 1) & ~1, Nlim = (Nn + 1) & ~1;
    const int nchunk = (Kd + kGemmKC - 1) / kGemmKC;
    // even split of the 8-blocks over CTA tiles
    const int mb = (Mm + 7) >> 3, nb = (Nn + 7) >> 3;
    const int tm = (mb + 7) / 8, tn = (nb + 15) / 16;     // number of CTA tiles per dimension
    for (int ti = 0; ti < tm; ++ti) {
        // full 8-block row tiles and one remainder tile: with a triangular A the contraction of a tile starts at its FIRST row,
        // so whole tiles (rows 0, 64, 128, ...) skip more chunks than an even split and their warp tiles are full 4 x 4
        const int bm0 = tri >= 0 ? ti * 8 : (int)((long)ti * mb / tm), bm1 = tri >= 0 ? min(mb, bm0 + 8) : (int)((long)(ti + 1) * mb / tm);
        for (int tj = 0; tj < tn; ++tj) {
            const int bn0 = (int)((long)tj * nb / tn), bn1 = (int)((long)(tj + 1) * nb / tn);   // <= 16 blocks
            if (lower_only && (bm1 * 8 <= bn0 * 8)) continue;                // whole tile above the diagonal
            const int m0 = bm0 * 8, n0 = bn0 * 8;
            // warp's blocks inside the tile: rows split in 2, cols in 4 (even split)
            const int tmb = bm1 - bm0, tnb = bn1 - bn0;
            const int rb0 = wr * tmb / 2, rb1 = (wr + 1) * tmb / 2;          // <= 4 blocks
            // column group rotated by warp row, tile and CTA so that the one-block-larger group (uneven splits) does
            // not always land on the same SM sub-partition (warp w issues on sub-partition w & 3 = wc)
            const int cg_ = (wc + 2 * wr + ti + tj + (int)blockIdx.x) & 3;
            const int cb0 = cg_ * tnb / 4, cb1 = (cg_ + 1) * tnb / 4;        // <= 4 blocks
            int nri = rb1 - rb0, nci = cb1 - cb0;
            // lower_only: a warp tile entirely above the diagonal is skipped (warp-uniform)
            if (lower_only && (m0 + rb1 * 8 <= n0 + cb0 * 8)) nri = 0;
            const int shape = (nri > 0 && nci > 0) ? (nri - 1) * 4 + (nci - 1) : -1;
            double acc[4][4][2];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
            __syncthreads();      // previous users of the stage buffers are done
            const int c0 = (tri >= 0 && m0 > tri) ? (m0 - tri) / kGemmKC : 0;
            const GemmStageMap mp = gemm_stage_map(A, lda, B, ldb, m0, n0, Mlim, Nlim);
            // prologue: chunks c0, c0 + 1 in flight (a group is committed every step so that wait_group counts stay uniform)
            gemm_stage_load(sm + (c0 % kGemmStages) * kGemmStageDoubles, mp, lda, ldb, c0 * kGemmKC, Kd);
            cp_async_commit();
            if (c0 + 1 < nchunk) gemm_stage_load(sm + ((c0 + 1) % kGemmStages) * kGemmStageDoubles, mp, lda, ldb, (c0 + 1) * kGemmKC, Kd);
            cp_async_commit();
            for (int c = c0; c < nchunk; ++c) {
                cp_async_wait<1>();                  // chunk c has landed (this thread's copies)
                __syncthreads();                     // ... everyone's; and every warp is done with chunk c - 1
                if (c + 2 < nchunk) gemm_stage_load(sm + ((c + 2) % kGemmStages) * kGemmStageDoubles, mp, lda, ldb, (c + 2) * kGemmKC, Kd);
                cp_async_commit();
                const double* cur = sm + (c % kGemmStages) * kGemmStageDoubles;
                const double* As = cur + rb0 * 8 + lr;
                const double* Bs = cur + kGemmKC * kGemmLdA + cb0 * 8 + lr;
                // triangular operand: contraction rows k of this chunk only reach output rows m <= k + tri, i.e. the first
                // `na` 8-row blocks of the warp tile (8-row granularity at no cost: the block count is a template argument)
                int shape_c = shape;
                if (tri >= 0 && shape >= 0) {
                    const int na = min(nri, ((c + 1) * kGemmKC + tri - (m0 + rb0 * 8) + 7) >> 3);
                    shape_c = na > 0 ? (na - 1) * 4 + (nci - 1) : -1;
                }
                switch (shape_c) {
#define MEEM_GEMM_CASE(NI, NJ) case (NI - 1) * 4 + (NJ - 1): gemm_chunk<NI, NJ>(As, Bs, lc, acc); break;
                    MEEM_GEMM_CASE(1, 1) MEEM_GEMM_CASE(1, 2) MEEM_GEMM_CASE(1, 3) MEEM_GEMM_CASE(1, 4)
                    MEEM_GEMM_CASE(2, 1) MEEM_GEMM_CASE(2, 2) MEEM_GEMM_CASE(2, 3) MEEM_GEMM_CASE(2, 4)
                    MEEM_GEMM_CASE(3, 1) MEEM_GEMM_CASE(3, 2) MEEM_GEMM_CASE(3, 3) MEEM_GEMM_CASE(3, 4)
                    MEEM_GEMM_CASE(4, 1) MEEM_GEMM_CASE(4, 2) MEEM_GEMM_CASE(4, 3) MEEM_GEMM_CASE(4, 4)
#undef MEEM_GEMM_CASE
                    default: break;
                }
            }
            cp_async_wait<0>();
            if (shape >= 0) {
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (i < nri && j < nci)
                            epi(m0 + (rb0 + i) * 8 + lr, n0 + (cb0 + j) * 8 + 2 * lc, acc[i][j][0], acc[i][j][1]);
            }
        }
    }
    __syncthreads();
}


// mbarrier + 1-D bulk copy (TMA unit, SASS UBLKCP): whole rows / whole matrices global -> shared with one instruction
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src), "r"(bytes), "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
    asm volatile("{\n.reg .pred p;\nWAIT_%=:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                 "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(parity) : "memory");
}

// ------------------------------------------------------------------------------------------------
// Lower triangle of C = W^T W  (W: Kd x Nc row-major, leading dimension ldw), the symmetric rank-Kd update of the
// factorised-H form of the frequency stage.  Unlike cta_gemm_dmma, whole-width chunks of W (kSyrkKC x Nc) are staged,
// so a warp can own ANY 8x8-block rectangle of C: the lower triangle is cut into warp tiles of <= 4 x 4 blocks
// (block groups a >= b; diagonal tiles a == b only compute their lower blocks and load one fragment set), the tiles
// are sorted by size and dealt to the 8 warps pass by pass - 325 useful block products in 49 block-times per warp for
// Nc = 200, against 64 for the rectangular 64 x 128 CTA tiles with warp-level skipping.
constexpr int kSyrkKC = 16;
constexpr int kSyrkMaxTiles = 96;
struct SyrkPlan {                       // lives in static shared memory of the calling kernel
    int nb, nt;
    unsigned char ta[kSyrkMaxTiles], tb[kSyrkMaxTiles];
};
__host__ __device__ inline int syrk_lds(int Nc) { const int n8 = (Nc + 7) & ~7; return n8 + ((n8 & 15) ? 12 : 4); }   // = 4 (mod 16)
__host__ __device__ inline bool syrk_fits(int Nc) {
    const int nb = (Nc + 7) >> 3, ng = (nb + 3) >> 2;
    return kGemmStages * kSyrkKC * syrk_lds(Nc) <= kGemmSmemDoubles && ng * (ng + 1) / 2 <= kSyrkMaxTiles;
}

// Tile list of the lower triangle of an nb x nb block matrix: block groups of <= 4 (even split), tiles (a >= b), sorted by
// the number of 8x8 blocks they compute (diagonal tiles: triangular), largest first.  Sets plan.nt, not plan.nb.
__host__ __device__ inline void syrk_build_plan(SyrkPlan& plan, int nb) {
    const int ng = (nb + 3) >> 2;
    int nt = 0;
    for (int a = 0; a < ng; ++a)
        for (int b = 0; b <= a; ++b) {
            const int ga = (a + 1) * nb / ng - a * nb / ng, gb = (b + 1) * nb / ng - b * nb / ng;
            const int sz = (a == b) ? ga * (ga + 1) / 2 : ga * gb;
            int pos = nt++;
            while (pos > 0) {            // insertion sort, largest first
                const int pa = plan.ta[pos - 1], pb = plan.tb[pos - 1];
                const int qa = (pa + 1) * nb / ng - pa * nb / ng, qb = (pb + 1) * nb / ng - pb * nb / ng;
                const int psz = (pa == pb) ? qa * (qa + 1) / 2 : qa * qb;
                if (psz >= sz) break;
                plan.ta[pos] = (unsigned char)pa; plan.tb[pos] = (unsigned char)pb; --pos;
            }
            plan.ta[pos] = (unsigned char)a; plan.tb[pos] = (unsigned char)b;
        }
    plan.nt = nt;
}

template <int NI, int NJ, int KC = kSyrkKC>
__device__ __forceinline__ void syrk_chunk(const double* As, const double* Bs, int ld, int lc, double (&acc)[4][4][2]) {
#pragma unroll
    for (int ks = 0; ks < KC / 4; ++ks) {
        double af[NI], bf[NJ];
#pragma unroll
        for (int i = 0; i < NI; ++i) af[i] = As[(ks * 4 + lc) * ld + i * 8];
#pragma unroll
        for (int j = 0; j < NJ; ++j) bf[j] = Bs[(ks * 4 + lc) * ld + j * 8];
#pragma unroll
        for (int i = 0; i < NI; ++i)
#pragma unroll
            for (int j = 0; j < NJ; ++j) dmma(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
    }
}
// diagonal tile: rows == columns, only blocks j <= i
template <int NI, int KC = kSyrkKC>
__device__ __forceinline__ void syrk_chunk_diag(const double* As, int ld, int lc, double (&acc)[4][4][2]) {
#pragma unroll
    for (int ks = 0; ks < KC / 4; ++ks) {
        double af[NI];
#pragma unroll
        for (int i = 0; i < NI; ++i) af[i] = As[(ks * 4 + lc) * ld + i * 8];
#pragma unroll
        for (int i = 0; i < NI; ++i)
#pragma unroll
            for (int j = 0; j <= i; ++j) dmma(acc[i][j][0], acc[i][j][1], af[i], af[j]);
    }
}

// warp w copies rows w, w + nwarp, ...; lanes stride over the 16-byte groups of a row (no index division)
__device__ __forceinline__ void syrk_stage_load(double* st, int lds, const double* W, int ldw, int n8, int k0, int Kd) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    const int gpr = n8 >> 1;                      // 16-byte groups per row
    const unsigned s0 = (unsigned)__cvta_generic_to_shared(st);
    for (int r = warp; r < kSyrkKC; r += nwarp) {
        const bool ok = (k0 + r) < Kd;
        const double* src = ok ? W + (long)(k0 + r) * ldw : W;
        const unsigned dst = s0 + (unsigned)(r * lds) * (unsigned)sizeof(double);
        for (int g = lane; g < gpr; g += 32) cp_async16s(dst + 16u * g, ok ? src + 2 * g : src, ok ? 16 : 0);
    }
}

// bars != nullptr: the chunks are brought in by the TMA unit - lane r of warp 0 issues ONE bulk copy for row r of the chunk
// (16 instructions per chunk instead of ~1600 cp.async with their address arithmetic), completion through one mbarrier per
// stage (phase: the caller's parity bits, one per stage).  Requires Nc % 8 == 0 and rows [Kd, pad16(Kd)) of W zero.
__device__ __forceinline__ void syrk_stage_bulk(double* st, int lds, const double* W, int ldw, int n8, int k0, uint64_t* bar) {
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        if (lane == 0) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // earlier generic-proxy accesses to this stage
            mbar_expect_tx(bar, (unsigned)(kSyrkKC * n8 * sizeof(double)));
        }
        __syncwarp();
        if (lane < kSyrkKC) bulk_g2s(st + lane * lds, W + (long)(k0 + lane) * ldw, (unsigned)(n8 * sizeof(double)), bar);
    }
}

template <class Epi>
__device__ void cta_syrk_dmma(const double* __restrict__ W, int ldw, int Nc, int Kd, double* sm, SyrkPlan& plan,
                              uint64_t* bars, unsigned& phase, Epi epi) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    const int lr = lane >> 2, lc = lane & 3;
    const int nb = (Nc + 7) >> 3, n8 = nb * 8, lds = syrk_lds(Nc);
    const int ng = (nb + 3) >> 2;
    const int nchunk = (Kd + kSyrkKC - 1) / kSyrkKC;
    __syncthreads();
    if (plan.nb != nb) {                         // (re)build the sorted tile list: once per CTA for a uniform batch
        if (threadIdx.x == 0) syrk_build_plan(plan, nb);
        __syncthreads();
        if (threadIdx.x == 0) plan.nb = nb;
    }
    const int nt = plan.nt;
    for (int t0 = 0; t0 < nt; t0 += nwarp) {
        const int t = t0 + warp;
        const bool has = t < nt;
        const int a = has ? plan.ta[t] : 0, b = has ? plan.tb[t] : 0;
        const int rb0 = a * nb / ng, rb1 = (a + 1) * nb / ng, cb0 = b * nb / ng, cb1 = (b + 1) * nb / ng;
        const int nri = rb1 - rb0, nci = cb1 - cb0;
        const int shape = !has ? -1 : (a == b ? 16 + (nri - 1) : (nri - 1) * 4 + (nci - 1));
        double acc[4][4][2];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
        __syncthreads();      // previous users of the stage buffers are done
        if (bars) {
            syrk_stage_bulk(sm, lds, W, ldw, n8, 0, &bars[0]);
            if (1 < nchunk) syrk_stage_bulk(sm + kSyrkKC * lds, lds, W, ldw, n8, kSyrkKC, &bars[1]);
        } else {
            syrk_stage_load(sm, lds, W, ldw, n8, 0, Kd);
            cp_async_commit();
            if (1 < nchunk) syrk_stage_load(sm + kSyrkKC * lds, lds, W, ldw, n8, kSyrkKC, Kd);
            cp_async_commit();
        }
        for (int c = 0; c < nchunk; ++c) {
            const int st_ = c % kGemmStages;
            if (bars) { mbar_wait(&bars[st_], (phase >> st_) & 1u); phase ^= 1u << st_; }
            else cp_async_wait<1>();
            __syncthreads();
            if (c + 2 < nchunk) {
                const int s2 = (c + 2) % kGemmStages;
                if (bars) syrk_stage_bulk(sm + s2 * kSyrkKC * lds, lds, W, ldw, n8, (c + 2) * kSyrkKC, &bars[s2]);
                else syrk_stage_load(sm + s2 * kSyrkKC * lds, lds, W, ldw, n8, (c + 2) * kSyrkKC, Kd);
            }
            if (!bars) cp_async_commit();
            const double* cur = sm + (c % kGemmStages) * kSyrkKC * lds;
            const double* As = cur + rb0 * 8 + lr;
            const double* Bs = cur + cb0 * 8 + lr;
            switch (shape) {
#define MEEM_SYRK_CASE(NI, NJ) case (NI - 1) * 4 + (NJ - 1): syrk_chunk<NI, NJ>(As, Bs, lds, lc, acc); break;
                MEEM_SYRK_CASE(1, 1) MEEM_SYRK_CASE(1, 2) MEEM_SYRK_CASE(1, 3) MEEM_SYRK_CASE(1, 4)
                MEEM_SYRK_CASE(2, 1) MEEM_SYRK_CASE(2, 2) MEEM_SYRK_CASE(2, 3) MEEM_SYRK_CASE(2, 4)
                MEEM_SYRK_CASE(3, 1) MEEM_SYRK_CASE(3, 2) MEEM_SYRK_CASE(3, 3) MEEM_SYRK_CASE(3, 4)
                MEEM_SYRK_CASE(4, 1) MEEM_SYRK_CASE(4, 2) MEEM_SYRK_CASE(4, 3) MEEM_SYRK_CASE(4, 4)
#undef MEEM_SYRK_CASE
                case 16: syrk_chunk_diag<1>(As, lds, lc, acc); break;
                case 17: syrk_chunk_diag<2>(As, lds, lc, acc); break;
                case 18: syrk_chunk_diag<3>(As, lds, lc, acc); break;
                case 19: syrk_chunk_diag<4>(As, lds, lc, acc); break;
                default: break;
            }
        }
        cp_async_wait<0>();
        if (has) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (i < nri && j < nci && (a != b || j <= i))
                        epi((rb0 + i) * 8 + lr, (cb0 + j) * 8 + 2 * lc, acc[i][j][0], acc[i][j][1]);
        }
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------------
// Triangular product of the factorised-H form as ONE chunk stream (round 2, v19):
//   [aug ; W] = LA^T-contraction with B,   aug = the 2R right-hand-side rows (LA columns 0 .. 2R-1, full contraction),
//   W[m][:] = sum_{k >= m} LA[k][2R + m] B[k][:]   (LA columns 2R .. 2R + Mw - 1: row k of the Cholesky factor L)
// The output is processed in row groups of 4 8-row blocks (plus one group for the aug rows); group a contracts rows
// k >= 32 a only.  Per group every warp owns NI x NJ blocks (NJ = nb / 8 column blocks) plus up to 4 "extra" blocks of the
// nb % 8 left-over column blocks, dealt over the warps block by block, so that all 8 warps (and the 4 SM sub-partitions)
// carry the same number of DMMAs (nb = 25: 12 or 13 blocks each, against 12 / 16 for an even column split).
// Operands move in chunks of KC (16, or 8 when three 16-row stages do not fit) contraction rows: the 32-column LA tile of the
// group and the WHOLE-width rows of B.  Both sources are laid out with the leading dimensions of the stage (LAT: the tiles
// re-packed by k_hchol with 36-double rows; B: syrk_lds(K)), so a chunk is TWO contiguous bulk copies (TMA unit) issued by one
// lane of the producer warp - per-row copies (16 UBLKCP per chunk, serialised in one warp) starved the consumers.  The ring
// has a full (transaction count) and an empty (one arrival per warp) mbarrier per stage; there is no CTA barrier inside the
// stream - warps only meet through the ring - and the ring does not drain between groups.  The producer is the CTA's LAST
// warp, which carries the fewest blocks.
// Requires 8 <= nb <= 32, 2R <= 8, N >= 32; LAT and B zero-padded to a multiple of 8 rows.  Columns K .. 8 nb - 1 of B may hold
// anything (finite or not): they only feed output columns that are never stored or, for odd K, the one pad column next to K - 1.
constexpr int kTrKC = 8;                          // smallest chunk (the granularity of LAT's zero padding and of the triangular skip)
constexpr int kTrLdA = 36;
constexpr int kTrMaxStages = 8;
struct TrRing {                                   // static shared memory of the calling kernel
    uint64_t full[kTrMaxStages], empty[kTrMaxStages];
};
template <int KC = kTrKC>
__host__ __device__ inline int trmm_stage_doubles(int K) { return KC * (kTrLdA + syrk_lds(K)); }
template <int KC = kTrKC>
__host__ __device__ inline int trmm_stages(int K) {
    const int s = kGemmSmemDoubles / trmm_stage_doubles<KC>(K);
    return s > kTrMaxStages ? kTrMaxStages : s;
}
__host__ __device__ inline bool trmm_fits(int N, int K, int R) {
    const int nb = (K + 7) >> 3;
    return N >= 32 && nb >= 8 && nb <= 32 && 2 * R <= 8 && trmm_stages<kTrKC>(K) >= 4;
}
// Blocks of one row group (NI row blocks) that warp `warp` of 8 owns: NI x (nb / 8) plus its share of the (nb % 8) NI extra
// blocks, dealt block by block (extra e -> warp e % 8).  Host-visible through meemb200_debug_layout (CPU tests).
__host__ __device__ inline int trmm_warp_extras(int nb, int NI, int warp) {
    const int ext = (nb & 7) * NI;
    return warp < ext ? (ext - warp + 7) / 8 : 0;
}
__host__ __device__ inline int trmm_warp_blocks(int nb, int NI, int warp) { return NI * (nb >> 3) + trmm_warp_extras(nb, NI, warp); }
// Chunks of the whole stream: the aug group over rows 0 .. N-1, then group a over rows 32 a .. N-1, KC rows per chunk.
__host__ __device__ inline int trmm_total_chunks(int N, int Mw, int KC) {
    const int ng = (((Mw + 7) >> 3) + 3) >> 2;
    int total = (N + KC - 1) / KC;
    for (int a = 0; a < ng; ++a) total += (N - 32 * a + KC - 1) / KC;
    return total;
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}

// 8 contraction rows (two k-steps) of a warp tile of NA x NJ blocks
template <int NA, int NJ>
__device__ __forceinline__ void trmm_chunk(const double* As, const double* Bs, int ldb, int lc, double (&acc)[16][2]) {
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
        double af[NA], bf[NJ];
#pragma unroll
        for (int i = 0; i < NA; ++i) af[i] = As[(ks * 4 + lc) * kTrLdA + i * 8];
#pragma unroll
        for (int j = 0; j < NJ; ++j) bf[j] = Bs[(ks * 4 + lc) * ldb + j * 8];
#pragma unroll
        for (int i = 0; i < NA; ++i)
#pragma unroll
            for (int j = 0; j < NJ; ++j) dmma(acc[i * NJ + j][0], acc[i * NJ + j][1], af[i], bf[j]);
    }
}

// NJ = nb / 8 column blocks per warp; KC = contraction rows per chunk (8 or 16; the last chunk of a group may hold 8).
// bits: per-stage use parity of the ring (bit s), kept by the caller across calls.
// LAT: the LA tiles of the groups back to back in stream order, leading dimension kTrLdA (k_hchol); B: leading dimension
// syrk_lds(K) = that of the stage - so a chunk is TWO contiguous bulk copies.
// epi_w(m, n, v0, v1): W row m (0-based), columns n, n + 1;  epi_aug(m, n, v0, v1): aug row m (0 .. 7: check m < 2R)
template <int NJ, int KC, class EpiW, class EpiA>
__device__ void cta_trmm_stream(const double* __restrict__ LAT, const double* __restrict__ B, int Mw,
                                int N, int K, double* sm, TrRing& ring, unsigned& bits, int prod_warp, EpiW epi_w, EpiA epi_aug) {
    constexpr int NX = NJ == 4 ? 0 : 4;                           // extra accumulator blocks 4 NJ .. 4 NJ + NX - 1
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int lr = lane >> 2, lc = lane & 3;
    const int nb = (K + 7) >> 3, ldsb = syrk_lds(K);
    const int S = trmm_stages<KC>(K), stage_doubles = trmm_stage_doubles<KC>(K);
    const int mbw = (Mw + 7) >> 3, ng = (mbw + 3) >> 2;          // W row blocks, row groups
    const int rem = nb - 8 * NJ;
    // chunk stream: group -1 (aug) over rows 0 .. N-1, then groups a = 0 .. ng - 1 over rows 32 a .. N-1, KC rows per chunk
    const int Np = (N + 7) & ~7;                                  // LAT tiles and B are zero-padded to whole 8-row chunks
    const int total = trmm_total_chunks(N, Mw, KC);
    const int look = S > 3 ? S - 2 : S - 1;                       // chunks in flight ahead of the consumer
    const int pw = prod_warp < 0 ? (int)(blockDim.x >> 5) - 1 : prod_warp;   // producer: the last warp carries the fewest blocks
    int pg = -1, pk = 0, pissued = 0, pst = 0;                    // producer state (warp 0): next chunk = rows pk .. of group pg
    unsigned pbits = bits;
    const double* pa = LAT;
    auto issue = [&]() {
        if (pissued >= S) mbar_wait(&ring.empty[pst], ((pbits >> pst) & 1u) ^ 1u);   // the stage's previous chunk is consumed
        const int rows = min(KC, Np - pk);
        double* As = sm + (long)pst * stage_doubles;
        double* Bs = As + KC * kTrLdA;
        if (lane == 0) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_expect_tx(&ring.full[pst], (unsigned)(rows * (kTrLdA + ldsb) * sizeof(double)));
            bulk_g2s(As, pa, (unsigned)(rows * kTrLdA * sizeof(double)), &ring.full[pst]);
            bulk_g2s(Bs, B + (long)pk * ldsb, (unsigned)(rows * ldsb * sizeof(double)), &ring.full[pst]);
        }
        __syncwarp();
        pa += rows * kTrLdA;
        pbits ^= 1u << pst;
        ++pissued;
        if (++pst == S) pst = 0;
        pk += rows;
        if (pk >= Np) { ++pg; pk = 32 * pg; }
    };
    asm volatile("fence.proxy.async;" ::: "memory");              // this thread's generic-proxy stores to B (the table) before the bulk reads
    __syncthreads();                                              // ... everyone's; previous users of the stage memory are done
    if (warp == pw) for (int i = 0; i < look && pissued < total; ++i) issue();
    int st = 0;
    for (int g = -1; g < ng; ++g) {
        const int NI = g < 0 ? 1 : min(4, mbw - 4 * g);
        const int k0g = g < 0 ? 0 : 32 * g;
        // extra blocks of this warp: e = warp, warp + 8, ... < rem * NI -> (row block e % NI, column block 8 NJ + e / NI)
        const int nx = trmm_warp_extras(nb, NI, warp);
        int xi[4], xj[4];
#pragma unroll
        for (int x = 0; x < 4; ++x) {
            const int e = warp + 8 * x;
            xi[x] = 0; xj[x] = 0;
            if (e < rem * NI) { xi[x] = e % NI; xj[x] = 8 * NJ + e / NI; }
        }
        double acc[16][2];
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i][0] = acc[i][1] = 0.0;
        for (int k = k0g; k < N; k += KC) {
            if (warp == pw && pissued < total) issue();
            mbar_wait(&ring.full[st], (bits >> st) & 1u);
            bits ^= 1u << st;
#pragma unroll
            for (int h = 0; h < KC / 8; ++h) {
                if (k + 8 * h < N) {
                    const double* As = sm + (long)st * stage_doubles + h * 8 * kTrLdA + lr;
                    const double* Bs = sm + (long)st * stage_doubles + KC * kTrLdA + h * 8 * ldsb + lr;
                    const double* Bw = Bs + warp * NJ * 8;
                    const int na = g < 0 ? 1 : min(NI, ((k - k0g) >> 3) + h + 1);   // row blocks these contraction rows reach
                    switch (na) {
                        case 1: trmm_chunk<1, NJ>(As, Bw, ldsb, lc, acc); break;
                        case 2: trmm_chunk<2, NJ>(As, Bw, ldsb, lc, acc); break;
                        case 3: trmm_chunk<3, NJ>(As, Bw, ldsb, lc, acc); break;
                        default: trmm_chunk<4, NJ>(As, Bw, ldsb, lc, acc); break;
                    }
                    if (NX > 0) {
#pragma unroll
                        for (int x = 0; x < NX; ++x) {
                            if (x < nx && xi[x] < na) {               // warp-uniform
#pragma unroll
                                for (int ks = 0; ks < 2; ++ks)
                                    dmma(acc[(4 * NJ + x) & 15][0], acc[(4 * NJ + x) & 15][1],
                                         As[(ks * 4 + lc) * kTrLdA + xi[x] * 8], Bs[(ks * 4 + lc) * ldsb + xj[x] * 8]);
                            }
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&ring.empty[st]);
            if (++st == S) st = 0;
        }
        // epilogue of the group (no CTA barrier: the stream goes on)
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                if (i < NI) {
                    const int n = (warp * NJ + j) * 8 + 2 * lc;
                    if (g < 0) epi_aug(lr, n, acc[i * NJ + j][0], acc[i * NJ + j][1]);
                    else epi_w(32 * g + i * 8 + lr, n, acc[i * NJ + j][0], acc[i * NJ + j][1]);
                }
            }
#pragma unroll
        for (int x = 0; x < NX; ++x) {
            if (x < nx) {
                const int n = xj[x] * 8 + 2 * lc;
                if (g < 0) epi_aug(lr, n, acc[(4 * NJ + x) & 15][0], acc[(4 * NJ + x) & 15][1]);
                else epi_w(32 * g + xi[x] * 8 + lr, n, acc[(4 * NJ + x) & 15][0], acc[(4 * NJ + x) & 15][1]);
            }
        }
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------------
// Lower triangle of C = W^T W as a chunk stream (round 2, v19): the tile list of cta_syrk_dmma, but the kTrKC-row chunks of W
// go through the full / empty mbarrier ring of the triangular product - no CTA barrier per chunk, and the ring does not
// drain between passes (the chunk sequence simply wraps around to W's first rows).  Requires Nc % 8 == 0 (whole 16-byte
// groups per row) and rows [Kd, pad8(Kd)) of W zero.
template <int KC>
__host__ __device__ inline int syrk_stream_stages(int Nc) {
    const int s = kGemmSmemDoubles / (KC * syrk_lds(Nc));
    return s > kTrMaxStages ? kTrMaxStages : s;
}
template <int KC>
__host__ __device__ inline bool syrk_stream_fits(int Nc) { return syrk_fits(Nc) && syrk_stream_stages<KC>(Nc) >= 3; }

// W: Kd x Nc row-major with the leading dimension syrk_lds(Nc) of the stages (a chunk is ONE contiguous bulk copy); rows
// [Kd, pad16(Kd)) must be zero.
template <int KC, class Epi>
__device__ void cta_syrk_stream(const double* __restrict__ W, int Nc, int Kd, double* sm, SyrkPlan& plan,
                                TrRing& ring, unsigned& bits, int prod_warp, Epi epi) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    const int lr = lane >> 2, lc = lane & 3;
    const int nb = (Nc + 7) >> 3, lds = syrk_lds(Nc);
    const int ng = (nb + 3) >> 2;
    const int nchunk = (Kd + KC - 1) / KC;
    const int S = syrk_stream_stages<KC>(Nc), stage_doubles = KC * lds;
    asm volatile("fence.proxy.async;" ::: "memory");              // this thread's generic-proxy stores to W before the bulk reads
    __syncthreads();
    if (plan.nb != nb) {                         // (re)build the sorted tile list: once per CTA for a uniform batch
        if (threadIdx.x == 0) syrk_build_plan(plan, nb);
        __syncthreads();
        if (threadIdx.x == 0) plan.nb = nb;
    }
    const int nt = plan.nt;
    const int npass = (nt + nwarp - 1) / nwarp;
    const int total = npass * nchunk;
    const int look = S > 3 ? S - 2 : S - 1;
    const int pw = prod_warp < 0 ? nwarp - 1 : prod_warp;        // producer: the last warp carries the smallest tiles
    int pc = 0, pissued = 0, pst = 0;
    unsigned pbits = bits;
    auto issue = [&]() {
        if (pissued >= S) mbar_wait(&ring.empty[pst], ((pbits >> pst) & 1u) ^ 1u);
        if (lane == 0) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_expect_tx(&ring.full[pst], (unsigned)(stage_doubles * sizeof(double)));
            bulk_g2s(sm + (long)pst * stage_doubles, W + (long)pc * stage_doubles, (unsigned)(stage_doubles * sizeof(double)), &ring.full[pst]);
        }
        __syncwarp();
        pbits ^= 1u << pst;
        ++pissued;
        if (++pst == S) pst = 0;
        if (++pc == nchunk) pc = 0;
    };
    if (warp == pw) for (int i = 0; i < look && pissued < total; ++i) issue();
    int st = 0;
    for (int t0 = 0; t0 < nt; t0 += nwarp) {
        const int t = t0 + warp;
        const bool has = t < nt;
        const int a = has ? plan.ta[t] : 0, b = has ? plan.tb[t] : 0;
        const int rb0 = a * nb / ng, rb1 = (a + 1) * nb / ng, cb0 = b * nb / ng, cb1 = (b + 1) * nb / ng;
        const int nri = rb1 - rb0, nci = cb1 - cb0;
        const int shape = !has ? -1 : (a == b ? 16 + (nri - 1) : (nri - 1) * 4 + (nci - 1));
        double acc[4][4][2];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
        for (int c = 0; c < nchunk; ++c) {
            if (warp == pw && pissued < total) issue();
            mbar_wait(&ring.full[st], (bits >> st) & 1u);
            bits ^= 1u << st;
            const double* cur = sm + (long)st * stage_doubles;
            const double* As = cur + rb0 * 8 + lr;
            const double* Bs = cur + cb0 * 8 + lr;
            switch (shape) {
#define MEEM_SYRK_CASE(NI, NJ) case (NI - 1) * 4 + (NJ - 1): syrk_chunk<NI, NJ, KC>(As, Bs, lds, lc, acc); break;
                MEEM_SYRK_CASE(1, 1) MEEM_SYRK_CASE(1, 2) MEEM_SYRK_CASE(1, 3) MEEM_SYRK_CASE(1, 4)
                MEEM_SYRK_CASE(2, 1) MEEM_SYRK_CASE(2, 2) MEEM_SYRK_CASE(2, 3) MEEM_SYRK_CASE(2, 4)
                MEEM_SYRK_CASE(3, 1) MEEM_SYRK_CASE(3, 2) MEEM_SYRK_CASE(3, 3) MEEM_SYRK_CASE(3, 4)
                MEEM_SYRK_CASE(4, 1) MEEM_SYRK_CASE(4, 2) MEEM_SYRK_CASE(4, 3) MEEM_SYRK_CASE(4, 4)
#undef MEEM_SYRK_CASE
                case 16: syrk_chunk_diag<1, KC>(As, lds, lc, acc); break;
                case 17: syrk_chunk_diag<2, KC>(As, lds, lc, acc); break;
                case 18: syrk_chunk_diag<3, KC>(As, lds, lc, acc); break;
                case 19: syrk_chunk_diag<4, KC>(As, lds, lc, acc); break;
                default: break;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&ring.empty[st]);
            if (++st == S) st = 0;
        }
        if (has) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (i < nri && j < nci && (a != b || j <= i))
                        epi((rb0 + i) * 8 + lr, (cb0 + j) * 8 + 2 * lc, acc[i][j][0], acc[i][j][1]);
        }
    }
    __syncthreads();
}

}  // namespace meem
