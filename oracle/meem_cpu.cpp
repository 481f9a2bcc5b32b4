// Compiled CPU restatement of the reference's hot path (SURVEY §8(d) "CPU baseline" (iii)).
// TEST / BENCHMARK INFRASTRUCTURE, not product: only tests/ and bench.py's cpu_baseline legs load this library.
//
// It follows the reference's algorithm, not the product's: per frequency the dense N_T x N_T complex system is assembled
// (meem_engine.py:44-63, template copy + the omega-dependent blocks, :100-305) and solved by LAPACK zgetrf / zgetrs
// (scipy.linalg.lu_factor / lu_solve, meem_engine.py:609, :633), then reduced to added mass / damping / force / phase
// (meem_engine.py:326-393).  The closed forms are the g++ build of the same headers the CUDA kernels compile
// (openflash_b200/csrc/meem_forms.cuh, meem_math.cuh: MEEM_HD functions citing multi_equations.py line by line).
// LAPACK is the one scipy itself uses: the caller passes the zgetrf / zgetrs entry points taken from
// scipy.linalg.cython_lapack's capsules (Fortran ABI), so no BLAS is linked here.
// Frequencies are dealt to `nthreads` std::threads (each LAPACK call single-threaded: the caller limits BLAS threads to 1).
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

#include "../openflash_b200/csrc/meem_forms.cuh"

using namespace meem;

typedef void (*zgetrf_fn)(int* m, int* n, void* a, int* lda, int* ipiv, int* info);
typedef void (*zgetrs_fn)(char* trans, int* n, int* nrhs, void* a, int* lda, int* ipiv, void* b, int* ldb, int* info);

namespace {

struct Sweep {
    Geom g;
    Layout L;
    int R;
    const unsigned char* flags;      // [R][M+1]
    std::vector<double> tab;
    std::vector<cplx> A_template;    // static part of A, row-major [NT][NT]
    std::vector<cplx> b_static;      // [R][NT]
    std::vector<double> c;           // [R][p]
    std::vector<double> phi, bend;   // [R]
};

void build_geometry(Sweep& S) {
    make_layout(S.g, S.L);
    const Layout& L = S.L;
    const int NT = L.NT, p = NT - L.K, M = S.g.M;
    S.tab.assign(table_doubles(L), 0.0);
    fill_tables(S.g, L, S.tab.data(), 0, 1);
    S.A_template.assign((size_t)NT * NT, cmake(0.0, 0.0));
    fill_A_static(S.g, L, S.tab.data(), S.A_template.data(), (long)NT, 0, 1);
    S.b_static.assign((size_t)S.R * NT, cmake(0.0, 0.0));
    S.c.assign((size_t)S.R * p, 0.0);
    S.phi.assign(S.R, 0.0);
    S.bend.assign(S.R, 0.0);
    std::vector<cplx> cz(p);
    for (int j = 0; j < S.R; ++j) {
        const unsigned char* hv = S.flags + (size_t)j * (M + 1);
        fill_b_static(S.g, L, hv, S.b_static.data() + (size_t)j * NT, 1L, 0, 1);
        S.bend[j] = b_end_constant(S.g, hv);
        fill_c(S.g, L, S.tab.data(), hv, cz.data(), 0, 1);
        for (int n = 0; n < p; ++n) S.c[(size_t)j * p + n] = cz[n].re;
        S.phi[j] = phi_sum(S.g, L, S.tab.data(), hv);
    }
}

// one frequency: outputs of system f.  Work buffers are the caller's (one set per thread).
int solve_frequency(const Sweep& S, double omega, double rho, zgetrf_fn zgetrf, zgetrs_fn zgetrs, std::vector<cplx>& Ar, std::vector<cplx>& At,
                    std::vector<cplx>& B, std::vector<int>& ipiv, std::vector<double>& mk, std::vector<double>& Nk,
                    double* am, double* dp, double* force, double* phase, double* m0_out) {
    const Layout& L = S.L;
    const Geom& g = S.g;
    const int NT = L.NT, K = L.K, p = NT - K, M = g.M, R = S.R;
    const double m0 = wavenumber_mode(omega, g.h, 0);
    *m0_out = m0;
    for (int k = 0; k < K; ++k) { mk[k] = mk_entry(k, m0, g.h, 0); Nk[k] = nk_entry(k, m0, g.h, mk[k]); }
    Exterior ex;
    make_exterior(g, m0, ex);
    // A = template + omega-dependent blocks (row-major, as numpy holds it), then copied to the column-major order LAPACK
    // factors - the copy scipy.linalg.lu_factor makes of a C-contiguous array - so the row pivots are the reference's
    std::memcpy(Ar.data(), S.A_template.data(), sizeof(cplx) * (size_t)NT * NT);
    fill_A_omega(g, L, S.tab.data(), ex, mk.data(), Nk.data(), Ar.data(), (long)NT, 0, 1);
    for (int r0 = 0; r0 < NT; r0 += 32)
        for (int c0 = 0; c0 < NT; c0 += 32)
            for (int r = r0; r < std::min(NT, r0 + 32); ++r)
                for (int c = c0; c < std::min(NT, c0 + 32); ++c) At[(size_t)c * NT + r] = Ar[(size_t)r * NT + c];
    for (int j = 0; j < R; ++j) {
        std::memcpy(B.data() + (size_t)j * NT, S.b_static.data() + (size_t)j * NT, sizeof(cplx) * NT);
        for (int k = 0; k < K; ++k) set_real(B[(size_t)j * NT + L.vel_row[M - 1] + k], S.bend[j] * imk(0, k, g, ex, mk.data(), Nk.data()));
    }
    int n = NT, info = 0, nrhs = R;
    zgetrf(&n, &n, At.data(), &n, ipiv.data(), &info);
    if (info != 0) return 1;
    char trans = 'N';
    zgetrs(&trans, &n, &nrhs, At.data(), &n, ipiv.data(), B.data(), &n, &info);
    if (info != 0) return 1;
    const bool inf_m0 = std::isinf(m0);
    const double om_m0 = inf_m0 ? 0.0 : omega_of_m0(m0, g.h);
    for (int i = 0; i < R; ++i) {                            // motion of body i
        const cplx* X = B.data() + (size_t)i * NT;
        for (int j = 0; j < R; ++j) {                        // force on body j
            double re = 0.0, im = 0.0;
            const double* cj = S.c.data() + (size_t)j * p;
            for (int q = 0; q < p; ++q) { re += cj[q] * X[q].re; im += cj[q] * X[q].im; }
            const double cre = 2.0 * kPi * (re + S.phi[j]), cim = 2.0 * kPi * im;
            const double dpv = inf_m0 ? 0.0 : cim * om_m0 * rho;
            am[j * R + i] = cre * rho;
            dp[j * R + i] = dpv;
            if (i == j) {
                force[j] = excitation_force(dpv, m0, g.h, kRhoDefault);
                phase[j] = inf_m0 ? -(kPi / 2.0) : -(kPi / 2.0) + atan2(X[p].im, X[p].re) - hankel0_angle(m0 * g.a[M - 1]);
            }
        }
    }
    return 0;
}

}  // namespace

extern "C" {

// One geometry x F frequencies.  a, d: M+1 entries as the reference flattens them (a[M] = inf, d[M] = 0); nmk: M+1;
// flags: [R][M+1] region flags of the radiating modes.  Outputs in the reference's index order [f][j][i] / [f][j].
// status[f]: 0 ok, 1 LAPACK reported a singular factor.  Returns N_T, or -1 for a bad argument.
int meem_cpu_sweep(int M, double h, const double* a, const double* d, const int* nmk, int R, const unsigned char* flags, int F,
                   const double* omega, double rho, int nthreads, void* zgetrf_ptr, void* zgetrs_ptr, double* added_mass,
                   double* damping, double* force, double* phase, double* m0, int* status) {
    if (M < 1 || R < 1 || F < 1 || !zgetrf_ptr || !zgetrs_ptr) return -1;
    Sweep S;
    S.g = Geom{M, h, a, d, nmk};
    S.R = R;
    S.flags = flags;
    build_geometry(S);
    const int NT = S.L.NT, K = S.L.K;
    zgetrf_fn zgetrf = reinterpret_cast<zgetrf_fn>(zgetrf_ptr);
    zgetrs_fn zgetrs = reinterpret_cast<zgetrs_fn>(zgetrs_ptr);
    std::atomic<int> next(0);
    auto worker = [&]() {
        std::vector<cplx> Ar((size_t)NT * NT), At((size_t)NT * NT), B((size_t)R * NT);
        std::vector<int> ipiv(NT);
        std::vector<double> mk(K), Nk(K);
        for (;;) {
            const int f = next.fetch_add(1);
            if (f >= F) break;
            double* am = added_mass + (size_t)f * R * R;
            double* dp = damping + (size_t)f * R * R;
            const int rc = solve_frequency(S, omega[f], rho, zgetrf, zgetrs, Ar, At, B, ipiv, mk, Nk, am, dp, force + (size_t)f * R,
                                           phase + (size_t)f * R, m0 + f);
            status[f] = rc;
            if (rc) {
                const double nan = std::nan("");
                for (int e = 0; e < R * R; ++e) am[e] = dp[e] = nan;
                for (int e = 0; e < R; ++e) force[(size_t)f * R + e] = phase[(size_t)f * R + e] = nan;
            }
        }
    };
    const int nt = std::max(1, std::min(nthreads, F));
    if (nt == 1) {
        worker();
    } else {
        std::vector<std::thread> th;
        for (int t = 0; t < nt; ++t) th.emplace_back(worker);
        for (auto& t : th) t.join();
    }
    return NT;
}

}  // extern "C"
