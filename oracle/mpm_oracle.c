/*
 * mpm_oracle.c -- CPU restatement of the reference's Taichi MPM kernels.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path may link, import or
 * execute this file: only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker or as
 * the timed CPU baseline.
 *
 * PARITY UNPINNED: the reference (ranrandy/gaussian-splatting-mpm @ b2d1da1e)
 * ships no tests, golden vectors or fixtures for this path, and its arithmetic
 * lives in taichi==1.5.0 (requirements.txt:10), which is neither in the tree
 * nor installable here.  This file restates the reference kernels line by line
 * in fp32; ti.svd is restated from the published algorithm (McAdams, Selle,
 * Tamstorf, Teran, Sifakis 2011: Jacobi on A^T A with approximate Givens
 * quaternions, column sort, QR) under the contract U,V in SO(3), |sigma|
 * descending, sign of det(A) carried by sigma[2].
 *
 * Every function cites the reference file:line it follows (paths relative to
 * the reference root).  All arithmetic is fp32; build the parity library with
 * -O2 -ffp-contract=off -fno-fast-math (see oracle/Makefile).
 *
 * Particle loops are in index order and the 27 stencil nodes in (i,j,k) order,
 * k fastest: with one thread this DEFINES the oracle's summation order.  With
 * ORC_OMP the particle loops run under OpenMP (P2G with `omp atomic`), which is
 * the timed CPU baseline; its sums are order-nondeterministic like the
 * reference's float atomics.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

/* quirk flags */
#define ORC_STRICT_Q1 1 /* utils.py:37: jelly stress is never evaluated -> 0   */
#define ORC_STRICT_Q3 2 /* constitutive_models.py:256: foam F via Hadamard     */

/* material ids: utils.py:5-10 (0..3); 4,5 are this build's additions          */
enum { MAT_JELLY = 0, MAT_METAL = 1, MAT_SAND = 2, MAT_FOAM = 3, MAT_FLUID = 4, MAT_PLASTICINE = 5 };

typedef struct {
    float alpha;             /* model.py:49-51  Drucker-Prager friction coefficient */
    float plastic_viscosity; /* model.py:58 */
    float xi;                /* model.py:57 */
    float softening;         /* model.py:59 */
    int hardening;           /* model.py:56 */
    int flags;               /* ORC_STRICT_* */
} orc_params;

/* ------------------------------------------------------------------------- */
/* 3x3 helpers, row-major m[3*r+c]                                            */

static inline void mat_mul(const float* a, const float* b, float* out) {
    float t[9];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) {
            float s = a[3 * r + 0] * b[0 + c];
            s = s + a[3 * r + 1] * b[3 + c];
            s = s + a[3 * r + 2] * b[6 + c];
            t[3 * r + c] = s;
        }
    memcpy(out, t, sizeof t);
}
static inline void mat_T(const float* a, float* out) {
    float t[9];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) t[3 * r + c] = a[3 * c + r];
    memcpy(out, t, sizeof t);
}
static inline float mat_det(const float* m) {
    return m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) +
           m[2] * (m[3] * m[7] - m[4] * m[6]);
}
/* U diag(d) V^T */
static inline void u_diag_vt(const float* U, const float* d, const float* V, float* out) {
    float ud[9], vt[9];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) ud[3 * r + c] = U[3 * r + c] * d[c];
    mat_T(V, vt);
    mat_mul(ud, vt, out);
}

/* ------------------------------------------------------------------------- */
/* ti.svd restatement (call sites utils.py:33,385; constitutive_models.py:64,107,158,218) */

#define SVD_GAMMA 5.828427124746190f  /* 3 + 2 sqrt 2 */
#define SVD_CSTAR 0.923879532511287f  /* cos(pi/8)    */
#define SVD_SSTAR 0.382683432365090f  /* sin(pi/8)    */
#define SVD_SWEEPS 5

/* one Jacobi conjugation on the cyclic index triple (p,q,r); S symmetric 3x3
 * held as full matrix s[3][3]; quaternion qv = (x,y,z), qw                    */
static inline void jacobi_step(float s[3][3], float qv[3], float* qw, int p, int q, int r) {
    float ch = 2.0f * (s[p][p] - s[q][q]);
    float sh = s[p][q];
    int b = (SVD_GAMMA * sh * sh) < (ch * ch);
    float w = 1.0f / sqrtf(ch * ch + sh * sh);
    ch = b ? w * ch : SVD_CSTAR;
    sh = b ? w * sh : SVD_SSTAR;
    float c = ch * ch - sh * sh;
    float sn = 2.0f * ch * sh;
    float spp = s[p][p], sqq = s[q][q], spq = s[p][q], spr = s[p][r], sqr = s[q][r];
    float npp = c * c * spp + 2.0f * c * sn * spq + sn * sn * sqq;
    float nqq = sn * sn * spp - 2.0f * c * sn * spq + c * c * sqq;
    float npq = (c * c - sn * sn) * spq - c * sn * (spp - sqq);
    float npr = c * spr + sn * sqr;
    float nqr = -sn * spr + c * sqr;
    s[p][p] = npp; s[q][q] = nqq;
    s[p][q] = s[q][p] = npq;
    s[p][r] = s[r][p] = npr;
    s[q][r] = s[r][q] = nqr;
    /* qV <- qV (x) (ch ; sh e_r) */
    float vp = qv[p], vq = qv[q], vr = qv[r], ww = *qw;
    *qw = ww * ch - sh * vr;
    qv[p] = ch * vp + sh * vq;
    qv[q] = ch * vq - sh * vp;
    qv[r] = ch * vr + sh * ww;
}

static inline void cswap_neg(float* B, float* V, float* n, int a, int b) {
    /* new col a = old col b ; new col b = -old col a (keeps det V = +1) */
    for (int r = 0; r < 3; ++r) {
        float t = B[3 * r + a]; B[3 * r + a] = B[3 * r + b]; B[3 * r + b] = -t;
        t = V[3 * r + a]; V[3 * r + a] = V[3 * r + b]; V[3 * r + b] = -t;
    }
    float t = n[a]; n[a] = n[b]; n[b] = t;
}

ORC_API void orc_svd3(const float* A, float* U, float* sig, float* V) {
    float s[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            float t = A[0 + i] * A[0 + j];
            t = t + A[3 + i] * A[3 + j];
            t = t + A[6 + i] * A[6 + j];
            s[i][j] = t;
        }
    float qv[3] = {0.0f, 0.0f, 0.0f}, qw = 1.0f;
    for (int it = 0; it < SVD_SWEEPS; ++it) {
        jacobi_step(s, qv, &qw, 0, 1, 2);
        jacobi_step(s, qv, &qw, 1, 2, 0);
        jacobi_step(s, qv, &qw, 2, 0, 1);
    }
    float nq = 1.0f / sqrtf(qw * qw + qv[0] * qv[0] + qv[1] * qv[1] + qv[2] * qv[2]);
    float x = qv[0] * nq, y = qv[1] * nq, z = qv[2] * nq, w = qw * nq;
    V[0] = 1.0f - 2.0f * (y * y + z * z); V[1] = 2.0f * (x * y - w * z); V[2] = 2.0f * (x * z + w * y);
    V[3] = 2.0f * (x * y + w * z); V[4] = 1.0f - 2.0f * (x * x + z * z); V[5] = 2.0f * (y * z - w * x);
    V[6] = 2.0f * (x * z - w * y); V[7] = 2.0f * (y * z + w * x); V[8] = 1.0f - 2.0f * (x * x + y * y);
    float B[9];
    mat_mul(A, V, B);
    float n[3];
    for (int c = 0; c < 3; ++c) n[c] = B[c] * B[c] + B[3 + c] * B[3 + c] + B[6 + c] * B[6 + c];
    if (n[0] < n[1]) cswap_neg(B, V, n, 0, 1);
    if (n[0] < n[2]) cswap_neg(B, V, n, 0, 2);
    if (n[1] < n[2]) cswap_neg(B, V, n, 1, 2);
    /* Gram-Schmidt QR: U = Q, sigma = diag(R) */
    float u0[3], u1[3], u2[3];
    float s0 = sqrtf(n[0]);
    if (s0 > 1e-30f) { for (int r = 0; r < 3; ++r) u0[r] = B[3 * r] / s0; }
    else { u0[0] = 1.0f; u0[1] = 0.0f; u0[2] = 0.0f; }
    float d = u0[0] * B[1] + u0[1] * B[4] + u0[2] * B[7];
    float t[3] = {B[1] - d * u0[0], B[4] - d * u0[1], B[7] - d * u0[2]};
    float s1 = sqrtf(t[0] * t[0] + t[1] * t[1] + t[2] * t[2]);
    if (s1 > 1e-30f) { for (int r = 0; r < 3; ++r) u1[r] = t[r] / s1; }
    else {
        /* rank <= 1: any unit vector orthogonal to u0 */
        float e[3] = {0.0f, 0.0f, 0.0f};
        if (fabsf(u0[0]) < 0.57f) e[0] = 1.0f; else e[1] = 1.0f;
        float de = u0[0] * e[0] + u0[1] * e[1] + u0[2] * e[2];
        for (int r = 0; r < 3; ++r) t[r] = e[r] - de * u0[r];
        float ne = sqrtf(t[0] * t[0] + t[1] * t[1] + t[2] * t[2]);
        for (int r = 0; r < 3; ++r) u1[r] = t[r] / ne;
        s1 = 0.0f;
    }
    u2[0] = u0[1] * u1[2] - u0[2] * u1[1];
    u2[1] = u0[2] * u1[0] - u0[0] * u1[2];
    u2[2] = u0[0] * u1[1] - u0[1] * u1[0];
    float s2 = u2[0] * B[2] + u2[1] * B[5] + u2[2] * B[8];
    for (int r = 0; r < 3; ++r) { U[3 * r] = u0[r]; U[3 * r + 1] = u1[r]; U[3 * r + 2] = u2[r]; }
    sig[0] = s0; sig[1] = s1; sig[2] = s2;
}

/* ------------------------------------------------------------------------- */
/* constitutive_models.py                                                     */

/* constitutive_models.py:11-20  tau = 2 mu (F - R) F^T + lam J (J-1) I */
static void kirchoff_stress_FCR(const float* F, const float* U, const float* V, float J, float mu,
                                float lam, float* out) {
    float Vt[9], R[9], FmR[9], Ft[9], M[9];
    mat_T(V, Vt);
    mat_mul(U, Vt, R);
    for (int i = 0; i < 9; ++i) FmR[i] = 2.0f * mu * (F[i] - R[i]);
    mat_T(F, Ft);
    mat_mul(FmR, Ft, M);
    float iso = lam * J * (J - 1.0f);
    for (int i = 0; i < 9; ++i) out[i] = M[i];
    out[0] += iso; out[4] += iso; out[8] += iso;
}

/* constitutive_models.py:24-38  U diag(2 mu eps + lam tr eps) V^T F^T */
static void kirchoff_stress_StVK(const float* F, const float* U, const float* V, const float* sig,
                                 float mu, float lam, float* out) {
    float sv[3], eps[3], tau[3], M[9], Ft[9];
    for (int i = 0; i < 3; ++i) sv[i] = fmaxf(sig[i], 0.01f);
    for (int i = 0; i < 3; ++i) eps[i] = logf(sv[i]);
    float lss = logf(sv[0]) + logf(sv[1]) + logf(sv[2]);
    for (int i = 0; i < 3; ++i) tau[i] = 2.0f * mu * eps[i] + lam * lss * 1.0f;
    u_diag_vt(U, tau, V, M);
    mat_T(F, Ft);
    mat_mul(M, Ft, out);
}

/* constitutive_models.py:42-58  no clamp: NaN for sigma <= 0, as in the reference */
static void kirchoff_stress_Drucker_Prager(const float* F, const float* U, const float* V,
                                           const float* sig, float mu, float lam, float* out) {
    float lss = logf(sig[0]) + logf(sig[1]) + logf(sig[2]);
    float c[3], M[9], Ft[9];
    for (int i = 0; i < 3; ++i) c[i] = 2.0f * mu * logf(sig[i]) / sig[i] + lam * lss / sig[i];
    u_diag_vt(U, c, V, M);
    mat_T(F, Ft);
    mat_mul(M, Ft, out);
}

static inline float norm3(const float* v) { return sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }

/* constitutive_models.py:63-103 */
static void von_mises_return_mapping(const float* Ft, float mu, float lam, float* ys,
                                     const orc_params* P, float* F) {
    float U[9], V[9], so[3], sg[3], eps[3], tau[3], cond[3];
    orc_svd3(Ft, U, so, V);
    for (int i = 0; i < 3; ++i) sg[i] = fmaxf(so[i], 0.01f);
    for (int i = 0; i < 3; ++i) eps[i] = logf(sg[i]);
    float temp = (eps[0] + eps[1] + eps[2]) / 3.0f;
    for (int i = 0; i < 3; ++i) tau[i] = 2.0f * mu * eps[i] + lam * (eps[0] + eps[1] + eps[2]) * 1.0f;
    float sum_tau = tau[0] + tau[1] + tau[2];
    for (int i = 0; i < 3; ++i) cond[i] = tau[i] - sum_tau / 3.0f;
    if (norm3(cond) > *ys) {
        float eh[3] = {eps[0] - temp, eps[1] - temp, eps[2] - temp};
        float ehn = norm3(eh) + 1e-6f;
        float dg = ehn - *ys / (2.0f * mu);
        float e2[3];
        for (int i = 0; i < 3; ++i) eps[i] = eps[i] - (dg / ehn) * eh[i];
        for (int i = 0; i < 3; ++i) e2[i] = expf(eps[i]);
        u_diag_vt(U, e2, V, F);
        if (P->hardening == 1) *ys += 2.0f * mu * P->xi * dg;
    } else {
        memcpy(F, Ft, 9 * sizeof(float));
    }
}

/* Build-defined (parity unpinned): von Mises with softening.
 * As :63-103 but the yield stress shrinks by softening*|delta eps|; once it is
 * <= 0 the particle is damaged: mu = lam = 0 and F_trial passes through.       */
static void plasticine_return_mapping(const float* Ft, float* mu, float* lam, float* ys,
                                      const orc_params* P, float* F) {
    if (*ys <= 0.0f) { memcpy(F, Ft, 9 * sizeof(float)); return; }
    float U[9], V[9], so[3], sg[3], eps[3], tau[3], cond[3];
    orc_svd3(Ft, U, so, V);
    for (int i = 0; i < 3; ++i) sg[i] = fmaxf(so[i], 0.01f);
    for (int i = 0; i < 3; ++i) eps[i] = logf(sg[i]);
    float temp = (eps[0] + eps[1] + eps[2]) / 3.0f;
    for (int i = 0; i < 3; ++i) tau[i] = 2.0f * (*mu) * eps[i] + (*lam) * (eps[0] + eps[1] + eps[2]) * 1.0f;
    float sum_tau = tau[0] + tau[1] + tau[2];
    for (int i = 0; i < 3; ++i) cond[i] = tau[i] - sum_tau / 3.0f;
    if (norm3(cond) > *ys) {
        float eh[3] = {eps[0] - temp, eps[1] - temp, eps[2] - temp};
        float ehn = norm3(eh) + 1e-6f;
        float dg = ehn - *ys / (2.0f * (*mu));
        float de[3], e2[3];
        for (int i = 0; i < 3; ++i) de[i] = (dg / ehn) * eh[i];
        for (int i = 0; i < 3; ++i) eps[i] = eps[i] - de[i];
        for (int i = 0; i < 3; ++i) e2[i] = expf(eps[i]);
        u_diag_vt(U, e2, V, F);
        *ys = *ys - P->softening * norm3(de);
        if (*ys <= 0.0f) { *mu = 0.0f; *lam = 0.0f; }
    } else {
        memcpy(F, Ft, 9 * sizeof(float));
    }
}

/* constitutive_models.py:106-140 */
static void sand_return_mapping(const float* Ft, float mu, float lam, const orc_params* P, float* F) {
    float U[9], V[9], sg[3], eps[3], eh[3];
    orc_svd3(Ft, U, sg, V);
    for (int i = 0; i < 3; ++i) eps[i] = logf(fmaxf(fabsf(sg[i]), 1e-14f));
    float tr = eps[0] + eps[1] + eps[2];
    for (int i = 0; i < 3; ++i) eh[i] = eps[i] - tr / 3.0f;
    float ehn = norm3(eh);
    float dg = ehn + (3.0f * lam + 2.0f * mu) / (2.0f * mu) * tr * P->alpha;
    for (int i = 0; i < 9; ++i) F[i] = 0.0f;
    if (dg <= 0.0f) memcpy(F, Ft, 9 * sizeof(float));
    if (dg > 0.0f && tr > 0.0f) {
        float Vt[9];
        mat_T(V, Vt);
        mat_mul(U, Vt, F);
    }
    if (dg > 0.0f && tr <= 0.0f) {
        float H[3], sn[3];
        for (int i = 0; i < 3; ++i) H[i] = eps[i] - eh[i] * (dg / ehn);
        for (int i = 0; i < 3; ++i) sn[i] = expf(H[i]);
        u_diag_vt(U, sn, V, F);
    }
}

/* constitutive_models.py:143-213 (defined, never dispatched by the reference) */
static void fluid_return_mapping(const float* Ft, float mu, float ys, const orc_params* P, float dt,
                                 float* F) {
    float U[9], V[9], sg[3], eps[3], eh[3], st[3];
    orc_svd3(Ft, U, sg, V);
    for (int i = 0; i < 3; ++i) eps[i] = logf(fmaxf(fabsf(sg[i]), 0.01f));
    float tr = eps[0] + eps[1] + eps[2];
    for (int i = 0; i < 3; ++i) eh[i] = eps[i] - tr / 3.0f;
    for (int i = 0; i < 3; ++i) st[i] = 2.0f * mu * eh[i];
    float stn = norm3(st);
    float yv = stn - sqrtf(2.0f / 3.0f) * ys;
    if (yv > 0.0f) {
        float mu_hat = mu * (sg[0] * sg[0] + sg[1] * sg[1] + sg[2] * sg[2]) / 3.0f;
        float pf = 1.0f + P->plastic_viscosity / (2.0f * mu_hat * dt);
        float snn = stn - yv / pf;
        float en[3], e2[3];
        for (int i = 0; i < 3; ++i) en[i] = (1.0f / (2.0f * mu)) * ((snn / stn) * st[i]) + tr / 3.0f;
        for (int i = 0; i < 3; ++i) e2[i] = expf(en[i]);
        u_diag_vt(U, e2, V, F);
    } else {
        memcpy(F, Ft, 9 * sizeof(float));
    }
}

/* constitutive_models.py:217-259 */
static void viscoplasticity_return_mapping_with_StVK(const float* Ft, float mu, float ys,
                                                     const orc_params* P, float dt, float* F) {
    float U[9], V[9], so[3], sg[3], b[3], eps[3], eh[3], st[3];
    orc_svd3(Ft, U, so, V);
    for (int i = 0; i < 3; ++i) sg[i] = fmaxf(so[i], 0.01f);
    for (int i = 0; i < 3; ++i) b[i] = sg[i] * sg[i];
    for (int i = 0; i < 3; ++i) eps[i] = logf(sg[i]);
    float tr = eps[0] + eps[1] + eps[2];
    for (int i = 0; i < 3; ++i) eh[i] = eps[i] - tr / 3.0f;
    for (int i = 0; i < 3; ++i) st[i] = 2.0f * mu * eh[i];
    float stn = norm3(st);
    float y = stn - 0.8f * sqrtf(2.0f / 3.0f) * ys;
    if (y > 0.0f) {
        float mu_hat = mu * (b[0] + b[1] + b[2]) / 3.0f;
        float snn = stn - y / (1.0f + P->plastic_viscosity * 2.0f / (2.0f * mu_hat * dt));
        float en[3], e2[3];
        for (int i = 0; i < 3; ++i) en[i] = 1.0f / (2.0f * mu) * ((snn / stn) * st[i]) + tr / 3.0f;
        for (int i = 0; i < 3; ++i) e2[i] = expf(en[i]);
        if (P->flags & ORC_STRICT_Q3) {
            /* :256 `U * sig_elastic * V.transpose()` is element-wise in Taichi */
            for (int i = 0; i < 9; ++i) F[i] = 0.0f;
            for (int i = 0; i < 3; ++i) F[4 * i] = U[4 * i] * e2[i] * V[4 * i];
        } else {
            u_diag_vt(U, e2, V, F);
        }
    } else {
        memcpy(F, Ft, 9 * sizeof(float));
    }
}

/* utils.py:14-54 compute_stress_from_F_trial */
ORC_API void orc_compute_stress_from_F_trial(int n, const float* material, float* mu, float* lam,
                                             float* yield_stress, const float* F_trial, float* F,
                                             float* stress, float* pressure, float dt,
                                             const orc_params* P) {
#ifdef ORC_OMP
#pragma omp parallel for schedule(static)
#endif
    for (int p = 0; p < n; ++p) {
        const float* Ft = F_trial + 9 * (size_t)p;
        float* Fp = F + 9 * (size_t)p;
        int m = (int)material[p];
        if (m == MAT_METAL) von_mises_return_mapping(Ft, mu[p], lam[p], &yield_stress[p], P, Fp);
        else if (m == MAT_SAND) sand_return_mapping(Ft, mu[p], lam[p], P, Fp);
        else if (m == MAT_FOAM) viscoplasticity_return_mapping_with_StVK(Ft, mu[p], yield_stress[p], P, dt, Fp);
        else if (m == MAT_FLUID) fluid_return_mapping(Ft, mu[p], yield_stress[p], P, dt, Fp);
        else if (m == MAT_PLASTICINE) plasticine_return_mapping(Ft, &mu[p], &lam[p], &yield_stress[p], P, Fp);
        else memcpy(Fp, Ft, 9 * sizeof(float));

        float J = mat_det(Fp);
        float st[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        float U[9], S[3], V[9];
        orc_svd3(Fp, U, S, V);
        if (m == MAT_JELLY) {
            /* utils.py:37 tests the field object, not model.material[p] (quirk Q1) */
            if (!(P->flags & ORC_STRICT_Q1)) kirchoff_stress_FCR(Fp, U, V, J, mu[p], lam[p], st);
        } else if (m == MAT_METAL || m == MAT_FOAM || m == MAT_FLUID || m == MAT_PLASTICINE) {
            kirchoff_stress_StVK(Fp, U, V, S, mu[p], lam[p], st);
        } else if (m == MAT_SAND) {
            kirchoff_stress_Drucker_Prager(Fp, U, V, S, mu[p], lam[p], st);
        }
        float* sp = stress + 9 * (size_t)p;
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) sp[3 * r + c] = (st[3 * r + c] + st[3 * c + r]) / 2.0f;
        if (pressure) pressure[p] = 0.0f;
    }
}

/* ------------------------------------------------------------------------- */
/* utils.py:94-110 (identical at :221-253): base cell, fx, w, dw              */

static inline void stencil(const float* x, float inv_dx, int* base, float* fx, float w[3][3], float dw[3][3]) {
    for (int d = 0; d < 3; ++d) {
        float gp = x[d] * inv_dx;
        base[d] = (int)(gp - 0.5f);
        fx[d] = gp - (float)base[d];
        float wa = 1.5f - fx[d], wb = fx[d] - 1.0f, wc = fx[d] - 0.5f;
        w[d][0] = wa * wa * 0.5f;
        w[d][1] = 0.75f - wb * wb;
        w[d][2] = wc * wc * 0.5f;
        dw[d][0] = fx[d] - 1.5f;
        dw[d][1] = -2.0f * (fx[d] - 1.0f);
        dw[d][2] = fx[d] - 0.5f;
    }
}

/* utils.py:80-86 compute_dweight */
static inline void compute_dweight(float w[3][3], float dw[3][3], int i, int j, int k, float inv_dx, float* out) {
    out[0] = dw[0][i] * w[1][j] * w[2][k] * inv_dx;
    out[1] = w[0][i] * dw[1][j] * w[2][k] * inv_dx;
    out[2] = w[0][i] * w[1][j] * dw[2][k] * inv_dx;
}

/* Sort key of this build (not in the reference, which never sorts): particles
 * are ordered by 4x4x4-cell tile, then by cell inside the tile.
 * key = ((tx*nt + ty)*nt + tz) << 6 | (lx<<4 | ly<<2 | lz), base = 4 t + l,
 * nt = ceil(n_grid/4).  base follows utils.py:94-95 exactly.                  */
ORC_API void orc_cell_keys(int n, int n_grid, float inv_dx, const float* xyz, int32_t* base_out, uint32_t* keys) {
    int nt = (n_grid + 3) / 4;
    for (int p = 0; p < n; ++p) {
        int b[3];
        for (int d = 0; d < 3; ++d) {
            float gp = xyz[3 * (size_t)p + d] * inv_dx;
            b[d] = (int)(gp - 0.5f);
            /* memory-safety clamp of this build (reference: none, quirk Q9) */
            if (b[d] < 0) b[d] = 0;
            if (b[d] > n_grid - 3) b[d] = n_grid - 3;
            if (base_out) base_out[3 * (size_t)p + d] = b[d];
        }
        uint32_t tile = (uint32_t)(((b[0] >> 2) * nt + (b[1] >> 2)) * nt + (b[2] >> 2));
        keys[p] = (tile << 6) | (uint32_t)(((b[0] & 3) << 4) | ((b[1] & 3) << 2) | (b[2] & 3));
    }
}

/* stable LSD radix sort (8-bit digits) of (key, index): the CPU sort the GPU
 * sort is compared against bit for bit                                        */
ORC_API void orc_stable_sort(int n, const uint32_t* keys, uint32_t* sorted_keys, uint32_t* perm) {
    uint32_t* k0 = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)n);
    uint32_t* k1 = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)n);
    uint32_t* v0 = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)n);
    uint32_t* v1 = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)n);
    for (int i = 0; i < n; ++i) { k0[i] = keys[i]; v0[i] = (uint32_t)i; }
    for (int pass = 0; pass < 4; ++pass) {
        size_t cnt[257];
        memset(cnt, 0, sizeof cnt);
        int sh = 8 * pass;
        for (int i = 0; i < n; ++i) cnt[((k0[i] >> sh) & 255u) + 1]++;
        for (int d = 0; d < 256; ++d) cnt[d + 1] += cnt[d];
        for (int i = 0; i < n; ++i) {
            size_t pos = cnt[(k0[i] >> sh) & 255u]++;
            k1[pos] = k0[i]; v1[pos] = v0[i];
        }
        uint32_t* t = k0; k0 = k1; k1 = t;
        t = v0; v0 = v1; v1 = t;
    }
    memcpy(sorted_keys, k0, sizeof(uint32_t) * (size_t)n);
    memcpy(perm, v0, sizeof(uint32_t) * (size_t)n);
    free(k0); free(k1); free(v0); free(v1);
}

/* utils.py:90-134 p2g */
ORC_API void orc_p2g(int n, int n_grid, float dx, float inv_dx, float dt, const float* xyz,
                     const float* vel, const float* C, const float* stress, const float* vol,
                     const float* mass, float* grid_v_in, float* grid_mass) {
#ifdef ORC_OMP
#pragma omp parallel for schedule(static)
#endif
    for (int p = 0; p < n; ++p) {
        const float* st = stress + 9 * (size_t)p;
        const float* Cp = C + 9 * (size_t)p;
        const float* vp = vel + 3 * (size_t)p;
        int base[3];
        float fx[3], w[3][3], dw[3][3];
        stencil(xyz + 3 * (size_t)p, inv_dx, base, fx, w, dw);
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j)
                for (int k = 0; k < 3; ++k) {
                    float dpos[3] = {((float)i - fx[0]) * dx, ((float)j - fx[1]) * dx, ((float)k - fx[2]) * dx};
                    int ix = base[0] + i, iy = base[1] + j, iz = base[2] + k;
                    float weight = w[0][i] * w[1][j] * w[2][k];
                    float dwt[3];
                    compute_dweight(w, dw, i, j, k, inv_dx, dwt);
                    float nv = -vol[p];
                    float ef[3], add[3];
                    for (int r = 0; r < 3; ++r) {
                        float s = (nv * st[3 * r]) * dwt[0];
                        s = s + (nv * st[3 * r + 1]) * dwt[1];
                        s = s + (nv * st[3 * r + 2]) * dwt[2];
                        ef[r] = s;
                    }
                    float wm = weight * mass[p];
                    for (int r = 0; r < 3; ++r) {
                        float cd = Cp[3 * r] * dpos[0];
                        cd = cd + Cp[3 * r + 1] * dpos[1];
                        cd = cd + Cp[3 * r + 2] * dpos[2];
                        add[r] = wm * (vp[r] + cd) + dt * ef[r];
                    }
                    if (ix < 0 || iy < 0 || iz < 0 || ix >= n_grid || iy >= n_grid || iz >= n_grid) continue;
                    size_t node = ((size_t)ix * n_grid + iy) * n_grid + iz;
#ifdef ORC_OMP
                    for (int r = 0; r < 3; ++r) {
#pragma omp atomic
                        grid_v_in[3 * node + r] += add[r];
                    }
#pragma omp atomic
                    grid_mass[node] += wm;
#else
                    for (int r = 0; r < 3; ++r) grid_v_in[3 * node + r] += add[r];
                    grid_mass[node] += wm;
#endif
                }
    }
}

/* utils.py:178-183 grid_normalization_and_gravity */
ORC_API void orc_grid_normalization_and_gravity(int n_grid, float dt, const float* gravity,
                                                const float* grid_mass, const float* grid_v_in,
                                                float* grid_v_out) {
    size_t G = (size_t)n_grid * n_grid * n_grid;
#ifdef ORC_OMP
#pragma omp parallel for schedule(static)
#endif
    for (size_t g = 0; g < G; ++g) {
        if (grid_mass[g] > 1e-15f) {
            for (int r = 0; r < 3; ++r) {
                float v = grid_v_in[3 * g + r] / grid_mass[g];
                v += dt * gravity[r];
                grid_v_out[3 * g + r] = v;
            }
        }
    }
}

/* boundary_conditions.py:23-27 BasicBC.apply */
ORC_API void orc_fixed_cube(int n_grid, float dx, const float* center, const float* size, float* grid_v_out) {
#ifdef ORC_OMP
#pragma omp parallel for schedule(static)
#endif
    for (int ix = 0; ix < n_grid; ++ix)
        for (int iy = 0; iy < n_grid; ++iy)
            for (int iz = 0; iz < n_grid; ++iz) {
                int id[3] = {ix, iy, iz};
                int in = 1;
                for (int d = 0; d < 3; ++d) in = in && (fabsf((float)id[d] * dx - center[d]) < size[d]);
                if (in) {
                    size_t g = ((size_t)ix * n_grid + iy) * n_grid + iz;
                    grid_v_out[3 * g] = 0.0f; grid_v_out[3 * g + 1] = 0.0f; grid_v_out[3 * g + 2] = 0.0f;
                }
            }
}

/* collider.py:13-44 MPM_Collider.collide */
ORC_API void orc_collide(int n_grid, float dx, const float* point, const float* normal, float friction,
                         float* grid_v_out) {
#ifdef ORC_OMP
#pragma omp parallel for schedule(static)
#endif
    for (int ix = 0; ix < n_grid; ++ix)
        for (int iy = 0; iy < n_grid; ++iy)
            for (int iz = 0; iz < n_grid; ++iz) {
                float off[3] = {(float)ix * dx - point[0], (float)iy * dx - point[1], (float)iz * dx - point[2]};
                float dp = off[0] * normal[0] + off[1] * normal[1] + off[2] * normal[2];
                if (dp < 0.0f) {
                    size_t g = ((size_t)ix * n_grid + iy) * n_grid + iz;
                    float v[3] = {grid_v_out[3 * g], grid_v_out[3 * g + 1], grid_v_out[3 * g + 2]};
                    float nc = v[0] * normal[0] + v[1] * normal[1] + v[2] * normal[2];
                    float mn = fminf(nc, 0.0f);
                    for (int r = 0; r < 3; ++r) v[r] = v[r] - mn * normal[r];
                    float len = norm3(v);
                    if (nc < 0.0f && len > 1e-20f) {
                        float sc = fmaxf(0.0f, len + nc * friction);
                        for (int r = 0; r < 3; ++r) v[r] = sc * (v[r] / len);
                    }
                    for (int r = 0; r < 3; ++r) grid_v_out[3 * g + r] = v[r] * 0.99f;
                }
            }
}

/* boundary_conditions.py:41-45 ImpulseBC.apply */
ORC_API void orc_impulse(int n, const float* xyz, float* vel, const float* mass, const float* center,
                         const float* size, const float* force, float substep_dt) {
    for (int p = 0; p < n; ++p) {
        int in = 1;
        for (int d = 0; d < 3; ++d) in = in && (fabsf(xyz[3 * (size_t)p + d] - center[d]) < size[d]);
        if (in)
            for (int d = 0; d < 3; ++d) vel[3 * (size_t)p + d] = vel[3 * (size_t)p + d] + force[d] / mass[p] * substep_dt;
    }
}

/* utils.py:219-282 g2p (+ :187-215 update_cov) */
ORC_API void orc_g2p(int n, int n_grid, float dx, float inv_dx, float dt, float* xyz, float* vel,
                     float* C, const float* F, float* F_trial, float* cov, const float* grid_v_out) {
    (void)dx;
#ifdef ORC_OMP
#pragma omp parallel for schedule(static)
#endif
    for (int p = 0; p < n; ++p) {
        int base[3];
        float fx[3], w[3][3], dw[3][3];
        stencil(xyz + 3 * (size_t)p, inv_dx, base, fx, w, dw);
        float nv[3] = {0, 0, 0}, nC[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, nF[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j)
                for (int k = 0; k < 3; ++k) {
                    float dpos[3] = {(float)i - fx[0], (float)j - fx[1], (float)k - fx[2]};
                    int ix = base[0] + i, iy = base[1] + j, iz = base[2] + k;
                    float weight = w[0][i] * w[1][j] * w[2][k];
                    float gv[3] = {0, 0, 0};
                    if (!(ix < 0 || iy < 0 || iz < 0 || ix >= n_grid || iy >= n_grid || iz >= n_grid)) {
                        size_t node = ((size_t)ix * n_grid + iy) * n_grid + iz;
                        gv[0] = grid_v_out[3 * node]; gv[1] = grid_v_out[3 * node + 1]; gv[2] = grid_v_out[3 * node + 2];
                    }
                    float cs = weight * inv_dx * 4.0f;
                    float dwt[3];
                    compute_dweight(w, dw, i, j, k, inv_dx, dwt);
                    for (int r = 0; r < 3; ++r) {
                        nv[r] += gv[r] * weight;
                        for (int c = 0; c < 3; ++c) {
                            nC[3 * r + c] += (gv[r] * dpos[c]) * cs;
                            nF[3 * r + c] += gv[r] * dwt[c];
                        }
                    }
                }
        for (int r = 0; r < 3; ++r) {
            vel[3 * (size_t)p + r] = nv[r];
            xyz[3 * (size_t)p + r] += dt * nv[r];
        }
        memcpy(C + 9 * (size_t)p, nC, sizeof nC);
        float IpF[9], Ft[9];
        for (int i = 0; i < 9; ++i) IpF[i] = nF[i] * dt;
        IpF[0] += 1.0f; IpF[4] += 1.0f; IpF[8] += 1.0f;
        mat_mul(IpF, F + 9 * (size_t)p, Ft);
        memcpy(F_trial + 9 * (size_t)p, Ft, sizeof Ft);
        if (cov) {
            /* update_cov: cov += dt (grad_v cov + cov grad_v^T) */
            float* cp = cov + 6 * (size_t)p;
            float cn[9] = {cp[0], cp[1], cp[2], cp[1], cp[3], cp[4], cp[2], cp[4], cp[5]};
            float a[9], b[9], gT[9], c1[9];
            mat_mul(nF, cn, a);
            mat_T(nF, gT);
            mat_mul(cn, gT, b);
            for (int i = 0; i < 9; ++i) c1[i] = cn[i] + dt * (a[i] + b[i]);
            cp[0] = c1[0]; cp[1] = c1[1]; cp[2] = c1[2]; cp[3] = c1[4]; cp[4] = c1[5]; cp[5] = c1[8];
        }
    }
}

/* utils.py:402-433 compute_cov_from_F */
ORC_API void orc_compute_cov_from_F(int n, const float* F_trial, const float* init_cov, float* cov) {
    for (int p = 0; p < n; ++p) {
        const float* F = F_trial + 9 * (size_t)p;
        const float* ic = init_cov + 6 * (size_t)p;
        float c0[9] = {ic[0], ic[1], ic[2], ic[1], ic[3], ic[4], ic[2], ic[4], ic[5]};
        float a[9], Ft[9], c[9];
        mat_mul(F, c0, a);
        mat_T(F, Ft);
        mat_mul(a, Ft, c);
        float* cp = cov + 6 * (size_t)p;
        cp[0] = c[0]; cp[1] = c[1]; cp[2] = c[2]; cp[3] = c[4]; cp[4] = c[5]; cp[5] = c[8];
    }
}

/* utils.py:377-398 compute_R_from_F */
ORC_API void orc_compute_R_from_F(int n, const float* F_trial, float* R) {
    for (int p = 0; p < n; ++p) {
        float U[9], s[3], V[9], Vt[9], Rm[9];
        orc_svd3(F_trial + 9 * (size_t)p, U, s, V);
        if (mat_det(U) < 0.0f) { U[2] = -U[2]; U[5] = -U[5]; U[8] = -U[8]; }
        if (mat_det(V) < 0.0f) { V[2] = -V[2]; V[5] = -V[5]; V[8] = -V[8]; }
        mat_T(V, Vt);
        mat_mul(U, Vt, Rm);
        mat_T(Rm, R + 9 * (size_t)p);
    }
}

/* utils.py:350-362 compute_mu_lam_from_E_nu */
ORC_API void orc_compute_mu_lam_from_E_nu(int n, const float* logE, const float* y, float* mu, float* lam) {
    for (int p = 0; p < n; ++p) {
        float E = powf(10.0f, logE[p]);
        float nu = 0.49f / (1.0f + expf(-y[p]));
        mu[p] = E / (2.0f * (1.0f + nu));
        lam[p] = E * nu / ((1.0f + nu) * (1.0f - 2.0f * nu));
    }
}

/* utils.py:366-373 compute_mass_from_vol_density */
ORC_API void orc_compute_mass_from_vol_density(int n, const float* density, const float* vol, float* mass) {
    for (int p = 0; p < n; ++p) mass[p] = density[p] * vol[p];
}

/* internel_filling/filling.py:12-24 assign_particle_to_grid + compute_particle_volume */
ORC_API void orc_particle_volume(int n, const float* pos, int n_grid, float grid_dx, float* vol) {
    size_t G = (size_t)n_grid * n_grid * n_grid;
    int32_t* cnt = (int32_t*)calloc(G, sizeof(int32_t));
    for (int p = 0; p < n; ++p) {
        int c[3];
        for (int d = 0; d < 3; ++d) c[d] = (int)floorf(pos[3 * (size_t)p + d] / grid_dx);
        cnt[((size_t)c[0] * n_grid + c[1]) * n_grid + c[2]] += 1;
    }
    for (int p = 0; p < n; ++p) {
        int c[3];
        for (int d = 0; d < 3; ++d) c[d] = (int)floorf(pos[3 * (size_t)p + d] / grid_dx);
        vol[p] = (grid_dx * grid_dx * grid_dx) / (float)cnt[((size_t)c[0] * n_grid + c[1]) * n_grid + c[2]];
    }
    free(cnt);
}

/* model.py:124-127 reset_grid_state */
ORC_API void orc_reset_grid(int n_grid, float* grid_mass, float* grid_v_in, float* grid_v_out) {
    size_t G = (size_t)n_grid * n_grid * n_grid;
    memset(grid_mass, 0, G * sizeof(float));
    memset(grid_v_in, 0, 3 * G * sizeof(float));
    memset(grid_v_out, 0, 3 * G * sizeof(float));
}

/* bench.py: torchrun exports OMP_NUM_THREADS=1 to its workers; the timed CPU arm asks for the cores it may use */
ORC_API void orc_set_num_threads(int n) {
#if defined(_OPENMP) && defined(ORC_OMP)
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

ORC_API int orc_num_threads(void) {
#if defined(_OPENMP) && defined(ORC_OMP)
    return omp_get_max_threads();
#else
    return 1;
#endif
}
