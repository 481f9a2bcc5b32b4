// g++ build of the device math/forms headers for CPU unit tests.
// TEST INFRASTRUCTURE, not product: the product path runs these same headers as CUDA device code.
#include <vector>
#include <cstring>
#include "../../openflash_b200/csrc/meem_forms.cuh"
using namespace meem;
extern "C" {
void hm_bessel(int kind, int n, const double* x, double* out) {
    for (int i = 0; i < n; ++i) {
        double j0, y0, j1, y1;
        switch (kind) {
            case 0: out[i] = i0e(x[i]); break;
            case 1: out[i] = i1e(x[i]); break;
            case 2: out[i] = k0e(x[i]); break;
            case 3: out[i] = k1e(x[i]); break;
            default:
                bessel_jy01(x[i], &j0, &y0, &j1, &y1);
                out[i] = kind == 4 ? j0 : kind == 5 ? y0 : kind == 6 ? j1 : y1;
        }
    }
}
void hm_wavenumber(int n, const double* omega, const double* h, double* out, int mode) {
    for (int i = 0; i < n; ++i) out[i] = wavenumber_mode(omega[i], h[i], mode);
}
void hm_mk(int K, double m0, double h, double* mk, double* nk, int mode) {
    for (int k = 0; k < K; ++k) { mk[k] = mk_entry(k, m0, h, mode); nk[k] = nk_entry(k, m0, h, mk[k]); }
}
// dense A (NT x NT complex, row-major), b (NT complex), c (p complex), phi for flags hv
int hm_assemble(int M, double h, const double* a, const double* d, const int* nmk, const unsigned char* hv,
                double m0, const double* mk, const double* Nk, double* A, double* b, double* c, double* phi) {
    Geom g{M, h, a, d, nmk};
    Layout L; make_layout(g, L);
    std::vector<double> tab(table_doubles(L));
    fill_tables(g, L, tab.data(), 0, 1);
    const int NT = L.NT;
    cplx* Ac = reinterpret_cast<cplx*>(A);
    cplx* bc = reinterpret_cast<cplx*>(b);
    std::memset(A, 0, sizeof(double) * 2 * NT * NT);
    std::memset(b, 0, sizeof(double) * 2 * NT);
    fill_A_static(g, L, tab.data(), Ac, (long)NT, 0, 1);
    Exterior ex; make_exterior(g, m0, ex);
    fill_A_omega(g, L, tab.data(), ex, mk, Nk, Ac, (long)NT, 0, 1);
    fill_b_static(g, L, hv, bc, 1L, 0, 1);
    const double cst = b_end_constant(g, hv);
    for (int k = 0; k < L.K; ++k) set_real(bc[L.vel_row[M - 1] + k], cst * imk(0, k, g, ex, mk, Nk));
    fill_c(g, L, tab.data(), hv, reinterpret_cast<cplx*>(c), 0, 1);
    *phi = phi_sum(g, L, tab.data(), hv);
    return NT;
}
}
