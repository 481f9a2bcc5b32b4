/* meemb200 — C ABI of the B200 (sm_100a) MEEM hot path.
 *
 * Drop-in boundary for OpenFLASH's batched (geometry x frequency x radiating mode) solve.  The
 * reference has no FFI: its boundary is the Python class API, and these entry points are what a
 * ctypes binding inside the reference's MEEMEngine would call (INTEGRATION.md shows the stub).
 * Reference file:line for what each call replaces (under /root/reference/package/src/openflash/):
 *
 *   meemb200_plan_create   flattening of domains into h, d[], a[], NMK[], heaving[]
 *                          (meem_engine.py:110-116), mode -> region flags (:350-367, :618-627)
 *   meemb200_solve         MEEMEngine.run_and_store_results numeric loop (meem_engine.py:602-676):
 *                          wavenumber / m_k_entry / N_k_multi (multi_equations.py:26-61, :379-384),
 *                          build_problem_cache + assemble_A_multi / assemble_b_multi
 *                          (meem_engine.py:44-63, :100-305), scipy.linalg.lu_factor / lu_solve
 *                          (:609, :633), compute_hydrodynamic_coefficients (:326-393)
 *   meemb200_coefficients  MEEMEngine.compute_hydrodynamic_coefficients for a caller-supplied X (:326-393)
 *   meemb200_assemble      MEEMEngine.assemble_A_multi / assemble_b_multi (meem_engine.py:44-63),
 *                          assemble_c_multi (:65-98) — for A/b parity tests
 *   meemb200_roots         wavenumber + _ensure_m_k_and_N_k_arrays (meem_engine.py:33-42)
 *   meemb200_bessel        scipy.special.ive / kve / hankel1 call sites (multi_equations.py:4-8)
 *
 * Conventions: plain pointers and sizes, no C++ or torch types.  Functions return 0 on success,
 * a negative MEEMB200_E* code on a bad argument or CUDA failure (never throw).  Numerical failure
 * of a single system (zero pivot, overflow to inf/NaN — the reference's LinAlgError / NaN
 * convention, meem_engine.py:611-614, :673-676) is NOT an error: that system's outputs are NaN
 * and its status word is non-zero; the batch continues.
 *
 * Memory: device buffers are caller-owned (e.g. torch tensors); the library allocates nothing on
 * the device.  All work is enqueued on the caller's stream; calls are asynchronous.
 */
#ifndef MEEMB200_H
#define MEEMB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MEEMB200_OK 0
#define MEEMB200_EINVAL (-1)   /* bad argument (NULL, sizes, non-increasing radii, ...) */
#define MEEMB200_ECUDA (-2)    /* CUDA runtime error, see meemb200_last_error() */
#define MEEMB200_ENOMEM (-3)   /* caller-provided buffer too small */
#define MEEMB200_ENODEV (-4)   /* no sm_100 device */

/* algorithm */
#define MEEMB200_ALGO_SCHUR 0  /* per-geometry real LU of A[:p,:p] + per-frequency KxK Schur system (default) */
#define MEEMB200_ALGO_DENSE 1  /* reference-shaped: assemble N_T x N_T complex A per system, partial-pivot LU */
#define MEEMB200_ALGO_SCHUR_LU 2 /* as SCHUR but the K x K frequency system by pivoted LU (the fallback of SCHUR's Cholesky) */
/* root finder */
#define MEEMB200_ROOTS_REFERENCE 0 /* replay scipy newton / brentq iterations (closest parity) */
#define MEEMB200_ROOTS_CONVERGED 1 /* safeguarded Newton to machine precision */
/* how `freq` is to be read */
#define MEEMB200_FREQ_OMEGA 0  /* angular frequencies; m0 = wavenumber(omega, h) */
#define MEEMB200_FREQ_M0 1     /* wavenumbers m0 given directly (solve_linear_system_multi(problem, m0)); inf allowed */

/* Host-side description of a batch: n_geom geometries x one shared grid of n_freq frequencies.
 * Per-geometry arrays are CSR-packed: geometry g owns entries reg_off[g] .. reg_off[g+1] of
 * a / d / nmk (M_g + 1 entries each, exactly the reference's lists incl. the exterior's a=inf, d=0),
 * and R_g * (M_g + 1) bytes of `heaving` starting at flag_off[g] ([mode][region], 0/1). */
typedef struct MeemBatch {
    int32_t n_geom;
    int32_t n_freq;
    const int32_t* reg_off;   /* [n_geom + 1] */
    const double* h;          /* [n_geom] water depth */
    const double* a;          /* radii, strictly increasing per geometry, last = inf */
    const double* d;          /* drafts, last = 0 */
    const int32_t* nmk;       /* harmonics per region, last = K (exterior) */
    const int32_t* n_modes;   /* [n_geom] R_g >= 1 */
    const int64_t* flag_off;  /* [n_geom + 1] */
    const uint8_t* heaving;   /* region heaving flags per radiating mode */
    const double* freq;       /* [n_freq] omega or m0, see freq_kind */
    int32_t freq_kind;
    int32_t algo;
    int32_t root_mode;
    int32_t want_x;           /* also return the solution vectors X */
    double rho;               /* fluid density used for added mass / damping (reference default 1023) */
} MeemBatch;

/* Sizes the caller needs to allocate buffers; element offsets of geometry g inside the outputs. */
typedef struct MeemPlanInfo {
    size_t input_bytes;       /* packed device input blob */
    size_t workspace_bytes;   /* device scratch */
    int64_t n_rr;             /* total elements of added_mass / damping  = sum_g F * R_g^2 */
    int64_t n_r;              /* total elements of force / phase          = sum_g F * R_g */
    int64_t n_sys;            /* n_geom * n_freq (m0, status) */
    int64_t n_x;              /* total complex elements of X = sum_g F * R_g * NT_g (0 unless want_x) */
    int64_t n_launches;       /* kernel launches one meemb200_solve enqueues */
    double flops;             /* real FP64 flops of the dense-matrix kernels one solve executes (closed form) */
    int64_t n_c;              /* total elements of c_vector = sum_g R_g * p_g (frequency independent) */
    int32_t geom_kernel;      /* geometry-stage kernel the plan chose: 0 dense LU per CTA (k_geom), 1 grid-cooperative dense LU
                               * (k_geom_coop), 2/3/4 window elimination (k_geom_band: small shared-memory window / shared-memory
                               * window / global-slot window), 5 grid-cooperative window elimination with static pivots (k_geom_coop2), -1 none (ALGO_DENSE) */
    int32_t freq_kernel;      /* frequency-stage kernels: 0 k_freq_gemm + k_freq_solve (Cholesky), 1 k_freq_small (one warp per system),
                               * 2 k_freq_lu (pivoted LU), 3 k_dense (ALGO_DENSE) */
} MeemPlanInfo;

/* Device output pointers (caller-owned).  Layouts are the reference's (meem_engine.py:573-580):
 * added_mass/damping [g][f][j][i] (force on mode j from motion of mode i), force/phase [g][f][j]. */
typedef struct MeemOutputs {
    double* added_mass;
    double* damping;
    double* exc_force;
    double* exc_phase;
    double* m0;               /* [g][f] wavenumber used */
    int32_t* status;          /* [g][f] 0 ok, 1 singular pivot, 2 non-finite */
    double* x;                /* optional [g][f][i][NT_g] complex128 (re,im interleaved), may be NULL */
    double* c_vector;         /* optional [g][j][p_g] real: the c vector of mode j (meem_engine.py:65-98, :390, :646-652).  It does
                               * not depend on the frequency and its imaginary part is 0: the host broadcasts it to the
                               * reference's (frequency, mode, p) complex layout (results.py:243-248).  May be NULL */
} MeemOutputs;

typedef struct meemb200_ctx meemb200_ctx;
typedef struct meemb200_plan meemb200_plan;

int meemb200_version(void);
const char* meemb200_last_error(void);

/* One context per process / GPU.  Queries the device (must be sm_100) and sets kernel attributes. */
int meemb200_create(int device, meemb200_ctx** ctx);
int meemb200_destroy(meemb200_ctx* ctx);
int meemb200_sm_count(const meemb200_ctx* ctx);

/* Validate and flatten a batch (host only, no device work).  The plan owns a packed host copy of
 * the inputs (meemb200_plan_blob) which the caller uploads with ONE host->device copy. */
int meemb200_plan_create(meemb200_ctx* ctx, const MeemBatch* batch, meemb200_plan** plan);
int meemb200_plan_destroy(meemb200_plan* plan);
int meemb200_plan_info(const meemb200_plan* plan, MeemPlanInfo* info);
const void* meemb200_plan_blob(const meemb200_plan* plan);
/* element offsets of geometry g: out[0]=rr, out[1]=r, out[2]=sys, out[3]=x, out[4]=NT, out[5]=p, out[6]=c_vector, out[7]=R */
int meemb200_plan_offsets(const meemb200_plan* plan, int32_t g, int64_t out[8]);
/* overwrite the frequency grid inside the packed blob (same n_freq), e.g. for the next sweep */
int meemb200_plan_set_freq(meemb200_plan* plan, const double* freq);

/* Enqueue the whole pipeline on `stream` (a cudaStream_t passed as void*; NULL = default stream):
 * roots -> per-geometry tables (+ Schur precompute) -> per-system solve -> coefficients.
 * d_input must hold the uploaded blob; d_workspace >= workspace_bytes. */
int meemb200_solve(meemb200_plan* plan, const void* d_input, void* d_workspace, size_t workspace_bytes,
                   const MeemOutputs* out, void* stream);

/* Host-buffer call (what the reference-side stub of INTEGRATION.md binds): uploads the packed inputs, solves, downloads
 * into the HOST pointers of `out` (any of exc_force / exc_phase / m0 / x / c_vector may be NULL), synchronises.
 * The one entry point that owns device memory: the first call allocates inputs + workspace + outputs with cudaMalloc
 * and the plan keeps the allocation for later calls (released by meemb200_plan_destroy). */
int meemb200_solve_host(meemb200_plan* plan, const MeemOutputs* host_out, void* stream);

/* Allocation-free variant of meemb200_solve.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The workspace - the scratch memory every kernel of the pipeline works in - is a file-scope
 * `static __device__ __align__(256)` array of MEEMB200_STATIC_WS_BYTES inside the library (each GPU that loads the
 * library holds its copy until the process exits); its device address is resolved once per device with
 * cudaGetSymbolAddress.  The caller still passes the uploaded inputs (d_input, as for meemb200_solve) and the device
 * output pointers.  The call itself makes no cudaMalloc / cudaFree / allocating library call; the first solve on a plan
 * creates the plan's six timing events (cudaEventCreate, host-side handles only), as meemb200_solve does.
 *   - plans whose MeemPlanInfo.workspace_bytes exceeds MEEMB200_STATIC_WS_BYTES are refused with MEEMB200_ENOMEM and a
 *     message naming both sizes (nothing is enqueued): use meemb200_solve with a caller-owned workspace for those;
 *   - ALL calls on a device share the one array, which keeps the last call's contents.  The prologue zeroes the
 *     work-queue counters (the only part that must start at zero) on `stream`; everything else is written before it is
 *     read.  Two calls that overlap on the GPU (different streams, or different plans / threads without ordering) RACE
 *     on the array: order them on one stream or with events, or give the overlapping call its own workspace through
 *     meemb200_solve. */
#define MEEMB200_STATIC_WS_BYTES (32u << 20)
int meemb200_solve_static_ws(meemb200_plan* plan, const void* d_input, const MeemOutputs* out, void* stream);

/* Parity / debugging exports mirroring the reference's intermediate API (device pointers).
 * A: [NT][NT] complex128 row-major, b: [R][NT] complex128, c: [R][p+1] complex128 (the c vector of
 * mode j followed by sum_i heaving_j[i]*int_phi_p_i) for geometry g of the plan at frequency index f.
 * Any of A/b/c may be NULL. */
int meemb200_assemble(meemb200_plan* plan, const void* d_input, void* d_workspace, size_t workspace_bytes,
                      int32_t g, int32_t f, double* d_A, double* d_b, double* d_c, void* stream);
/* The coupling table I_mk (multi_equations.py:96-124; table build meem_engine.py:136-141) of system (g, f) exactly as the
 * product frequency stage computes it (angle-addition form, sinc series near resonances - not the reference-ordered form
 * that meemb200_assemble uses): d_out[N][K] row-major, reference index order.  Parity export. */
int meemb200_imk_table(meemb200_plan* plan, const void* d_input, void* d_workspace, size_t workspace_bytes,
                       int32_t g, int32_t f, double* d_out, void* stream);
/* Reduction of ONE given solution X ([NT] complex128, device) of system (g, f) to what
 * MEEMEngine.compute_hydrodynamic_coefficients returns per mode j (meem_engine.py:326-393):
 * d_out[j][0..3] = real (added mass), imag (damping), excitation_phase, excitation_force, j over the plan's modes. */
int meemb200_coefficients(meemb200_plan* plan, const void* d_input, void* d_workspace, size_t workspace_bytes,
                          int32_t g, int32_t f, const double* d_X, double* d_out, void* stream);
/* Potential and velocity fields of one solution at n_points points (device pointers): replaces the arithmetic of
 * MEEMEngine.calculate_potentials / calculate_velocities (meem_engine.py:395-467, :478-553).  X: [NT] complex128
 * solution of geometry g at frequency index f for radiating mode `mode` (its heaving flags select the particular
 * solution).  out: 9 planes of n_points doubles: Re phiH, Im phiH, Re vrH, Im vrH, Re vzH, Im vzH, phiP, vrP, vzP.
 * The region of a point is chosen by r alone; the reference's inside-the-body NaN masks are the caller's to apply. */
int meemb200_fields(meemb200_plan* plan, const void* d_input, void* d_workspace, size_t workspace_bytes,
                    int32_t g, int32_t f, int32_t mode, const double* d_X, int64_t n_points,
                    const double* d_r, const double* d_z, double* d_out, void* stream);
/* m0[n], mk[n][K], Nk[n][K] for (freq[i], h[i]) pairs; device pointers. */
int meemb200_roots(meemb200_ctx* ctx, int32_t n, int32_t K, const double* d_freq, const double* d_h,
                   int32_t freq_kind, int32_t root_mode, double* d_m0, double* d_mk, double* d_Nk, void* stream);
/* kind 0..7 = I0e, I1e, K0e, K1e, J0, Y0, J1, Y1; device pointers */
int meemb200_bessel(meemb200_ctx* ctx, int32_t kind, int64_t n, const double* d_x, double* d_out, void* stream);

/* FP64 FMA / DMMA throughput microbenchmarks used for the roofline denominator (bench.py):
 * returns achieved FLOP/s in *flops (kind 0 = DFMA register loop, 1 = mma.sync m8n8k4 f64). */
int meemb200_fp64_peak(meemb200_ctx* ctx, int32_t kind, double* flops, void* stream);

/* Duration (ms) of the stages of the most recent meemb200_solve on this plan, measured with CUDA
 * events on the launching stream: ms[0] roots, ms[1] geometry stage, ms[2] frequency stage (all of it),
 * ms[3] its GEMM kernel k_freq_gemm, ms[4] its solve kernel k_freq_solve (ms[3] = ms[2], ms[4] = 0 when the
 * frequency stage ran in more than one wave or not on the Cholesky path).
 * Only valid after the stream has been synchronised. */
int meemb200_last_timing(meemb200_plan* plan, float ms[5]);

/* Per-phase SM-cycle totals (summed over CTAs) of the frequency-stage kernel of the most recent
 * meemb200_solve that used `d_workspace`: [0] coupling table, [1] NOT cycles: number of geometries whose H was
 * Cholesky-factorised (k_hchol) and whose frequency GEMM therefore ran the W = L^T Imk form,
 * [2] T = H Imk, [3] Imk^T T, [4] Cholesky, [5] back-substitution, [6] coefficients, [7] X;
 * [8..13] cooperative geometry stage as seen by CTA 0: [8] panel factorisation (owner CTAs), [9] factor + grid
 * barrier, [10] trailing update, [11] barrier after update, [12] back-substitution, [13] back-substitution + barrier.
 * Synchronises the stream.  Measurement only. */
int meemb200_phase_profile(meemb200_plan* plan, const void* d_workspace, double cycles[16], void* stream);

/* Test support: report every n-th system of the following solves as a failed Cholesky, so that it is re-solved by the
 * pivoted-LU kernel (the fallback of MEEMB200_ALGO_SCHUR, unreachable with physical inputs).  n = 0: off (default). */
int meemb200_plan_set_test_hook(meemb200_plan* plan, int32_t flag_every);

/* Host-side view of two device layouts (pure integer functions, no device needed; used by the CPU tests):
 * kind 0: out[r] = offset (doubles) of row r of the packed lower-triangular system matrix for r = 0..n, out[n+1] = padded
 *         size of the K = n triangle;
 * kind 1: warp-tile list of the symmetric rank update for n columns: out[0] = tiles, then per tile
 *         (group a, group b, first row block, row blocks, first column block, column blocks), largest tile first.
 * kind 2: chunk stream of the triangular product of k_freq_gemm for N = K = n (one zero row of H): out[0..7] = applies (0/1),
 *         rows per chunk, ring stages, row groups, chunks of the whole stream, column blocks, doubles of the re-tiled factor,
 *         leading dimension of the table in the slot; then per row group 0..7 ten ints: row blocks, first row, and the number
 *         of 8x8 blocks each of the 8 warps owns;
 * kind 3: offsets (doubles) of the re-tiled factor's tiles -1 (aug columns), 0 .. (n+31)/32 (out[last] = its size). */
int meemb200_debug_layout(int kind, int n, int* out, int out_len);

#ifdef __cplusplus
}
#endif
#endif /* MEEMB200_H */
