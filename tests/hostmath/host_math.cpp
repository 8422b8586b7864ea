// host_math.cpp -- TEST ONLY: compiles the product's per-particle device math headers
// (csrc/svd3.cuh, csrc/constitutive.cuh) as plain C++ so that their numerics can be compared with the
// CPU oracle without a GPU (tests/test_hostmath.py).  Never loaded by the product path.
#include "../../gaussian-splatting-mpm_b200/csrc/constitutive.cuh"

extern "C" {
__attribute__((visibility("default"))) void hm_constitutive(int n, const float* Ft, const int* mat, float* mu, float* lam,
                                                            float* ys, float dt, float alpha, float plastic_viscosity, float xi,
                                                            float softening, int hardening, unsigned flags, float* F, float* s6) {
    ConstParams P;
    P.alpha = alpha; P.plastic_viscosity = plastic_viscosity; P.xi = xi; P.softening = softening;
    P.hardening = hardening; P.flags = flags;
    for (int p = 0; p < n; ++p)
        constitutive_update(Ft + 9 * p, mat[p], mu[p], lam[p], ys[p], dt, P, F + 9 * p, s6 + 6 * p);
}
__attribute__((visibility("default"))) void hm_svd3(int n, const float* A, float* U, float* s, float* V) {
    for (int p = 0; p < n; ++p) svd3(A + 9 * p, U + 9 * p, s + 3 * p, V + 9 * p);
}
__attribute__((visibility("default"))) int hm_polar3(int n, const float* A, float* R) {
    int bad = 0;
    for (int p = 0; p < n; ++p) bad += polar3(A + 9 * p, R + 9 * p) ? 0 : 1;
    return bad;
}
}
#ifdef SVD3_COUNT
extern "C" __attribute__((visibility("default"))) long hm_sweeps(void) { long c = svd3_sweep_count; svd3_sweep_count = 0; return c; }
#endif
