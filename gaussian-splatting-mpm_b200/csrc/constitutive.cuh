// constitutive.cuh -- return mappings and Kirchhoff stresses of the reference
// (mpm_solver/constitutive_models.py, dispatched by utils.py:14-54), evaluated with ONE SVD per
// particle: a return mapping that yields F = U diag(s') V^T already has (U, s', V) as its SVD, and
// without yielding F = F_trial, so the reference's second ti.svd (utils.py:33) is redundant except
// for the strict-Q3 foam branch.  Stress is returned symmetrised (utils.py:52) as 6 floats
// {00,01,02,11,12,22}.
#pragma once
#include "../../include/mpmk.h"
#include "devmath.cuh"
#include "svd3.cuh"

struct ConstParams {
    float alpha, plastic_viscosity, xi, softening;
    int hardening;
    uint32_t flags;
};

// tau = U diag(t) U^T
MPMK_DEV void sym_from_u(const float* U, const float* t, float* s6) {
    s6[0] = U[0] * t[0] * U[0] + U[1] * t[1] * U[1] + U[2] * t[2] * U[2];
    s6[1] = U[0] * t[0] * U[3] + U[1] * t[1] * U[4] + U[2] * t[2] * U[5];
    s6[2] = U[0] * t[0] * U[6] + U[1] * t[1] * U[7] + U[2] * t[2] * U[8];
    s6[3] = U[3] * t[0] * U[3] + U[4] * t[1] * U[4] + U[5] * t[2] * U[5];
    s6[4] = U[3] * t[0] * U[6] + U[4] * t[1] * U[7] + U[5] * t[2] * U[8];
    s6[5] = U[6] * t[0] * U[6] + U[7] * t[1] * U[7] + U[8] * t[2] * U[8];
}

// F = U diag(d) V^T
MPMK_DEV void u_diag_vt(const float* U, const float* d, const float* V, float* F) {
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c)
            F[3 * r + c] = U[3 * r] * d[0] * V[3 * c] + U[3 * r + 1] * d[1] * V[3 * c + 1] + U[3 * r + 2] * d[2] * V[3 * c + 2];
}

// constitutive_models.py:24-38 with V^T F^T = diag(sigma) U^T folded in: t_i = (2 mu eps_i + lam tr eps) sigma_i
MPMK_DEV void stvk_principal(const float* sig, float mu, float lam, float* t) {
    float e0 = logf(fmaxf(sig[0], 0.01f)), e1 = logf(fmaxf(sig[1], 0.01f)), e2 = logf(fmaxf(sig[2], 0.01f));
    float tr = e0 + e1 + e2;
    t[0] = (2.0f * mu * e0 + lam * tr) * sig[0];
    t[1] = (2.0f * mu * e1 + lam * tr) * sig[1];
    t[2] = (2.0f * mu * e2 + lam * tr) * sig[2];
}

// Ft: trial deformation gradient (row-major 9).  Out: F (post return mapping), stress6.  mu/lam/ys are
// per-particle and may be updated (metal hardening :96-97, plasticine softening).
MPMK_DEV void constitutive_update(const float* Ft, int mat, float& mu, float& lam, float& ys,
                                                    float dt, const ConstParams& P, float* F, float* s6) {
#pragma unroll
    for (int i = 0; i < 9; ++i) F[i] = Ft[i];
    float U[9], V[9], sg[3];
    float t[3] = {0.0f, 0.0f, 0.0f};

    if (mat == MPMK_MAT_JELLY) {
        if (P.flags & MPMK_STRICT_Q1) {
            // utils.py:37 compares the field object with 0: the FCR branch is never taken (quirk Q1)
#pragma unroll
            for (int i = 0; i < 6; ++i) s6[i] = 0.0f;
        } else {
            // constitutive_models.py:11-20: tau = 2 mu (F - R) F^T + lam J (J-1) I, R = U V^T.  Only the
            // rotation is needed: Newton polar iteration when det F > 0, the SVD otherwise (inverted
            // element: sigma_3 < 0 under ti.svd's convention, U V^T stays a proper rotation).
            float R[9];
            const float J = F[0] * (F[4] * F[8] - F[5] * F[7]) - F[1] * (F[3] * F[8] - F[5] * F[6]) +
                            F[2] * (F[3] * F[7] - F[4] * F[6]);
            bool have_R = false;
            if (J > 1e-6f) have_R = polar3(F, R);
            if (!have_R) {
                const float one[3] = {1.0f, 1.0f, 1.0f};
                svd3(Ft, U, sg, V);
                u_diag_vt(U, one, V, R);
            }
            float D[9];
#pragma unroll
            for (int i = 0; i < 9; ++i) D[i] = 2.0f * mu * (F[i] - R[i]);
            float M[9];
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int c = 0; c < 3; ++c) M[3 * r + c] = D[3 * r] * F[3 * c] + D[3 * r + 1] * F[3 * c + 1] + D[3 * r + 2] * F[3 * c + 2];
            const float iso = lam * J * (J - 1.0f);
            s6[0] = M[0] + iso; s6[3] = M[4] + iso; s6[5] = M[8] + iso;
            s6[1] = 0.5f * (M[1] + M[3]); s6[2] = 0.5f * (M[2] + M[6]); s6[4] = 0.5f * (M[5] + M[7]);
        }
        return;
    }

    svd3(Ft, U, sg, V);

    if (mat == MPMK_MAT_METAL || mat == MPMK_MAT_PLASTICINE) {
        // constitutive_models.py:63-103 (metal); plasticine = same with softening (unpinned)
        bool dead = (mat == MPMK_MAT_PLASTICINE) && (ys <= 0.0f);
        float c0 = fmaxf(sg[0], 0.01f), c1 = fmaxf(sg[1], 0.01f), c2 = fmaxf(sg[2], 0.01f);
        float e0 = logf(c0), e1 = logf(c1), e2 = logf(c2);
        float tr = e0 + e1 + e2;
        float temp = tr / 3.0f;
        float t0 = 2.0f * mu * e0 + lam * tr, t1 = 2.0f * mu * e1 + lam * tr, t2 = 2.0f * mu * e2 + lam * tr;
        float st = (t0 + t1 + t2) / 3.0f;
        float d0 = t0 - st, d1 = t1 - st, d2 = t2 - st;
        float cn = sqrtf(d0 * d0 + d1 * d1 + d2 * d2);
        if (!dead && cn > ys) {
            float h0 = e0 - temp, h1 = e1 - temp, h2 = e2 - temp;
            float ehn = sqrtf(h0 * h0 + h1 * h1 + h2 * h2) + 1e-6f;
            float dg = ehn - ys / (2.0f * mu);
            float k = dg / ehn;
            float q0 = k * h0, q1 = k * h1, q2 = k * h2;
            e0 -= q0; e1 -= q1; e2 -= q2;
            sg[0] = expf(e0); sg[1] = expf(e1); sg[2] = expf(e2);
            u_diag_vt(U, sg, V, F);
            if (mat == MPMK_MAT_METAL) {
                if (P.hardening == 1) ys += 2.0f * mu * P.xi * dg;
            } else {
                ys = ys - P.softening * sqrtf(q0 * q0 + q1 * q1 + q2 * q2);
                if (ys <= 0.0f) { mu = 0.0f; lam = 0.0f; }
            }
        }
        stvk_principal(sg, mu, lam, t);
        sym_from_u(U, t, s6);
        return;
    }

    if (mat == MPMK_MAT_SAND) {
        // constitutive_models.py:106-140 + :42-58 (c_i sigma_i = 2 mu ln sigma_i + lam tr ln sigma)
        float e0 = logf(fmaxf(fabsf(sg[0]), 1e-14f)), e1 = logf(fmaxf(fabsf(sg[1]), 1e-14f)),
              e2 = logf(fmaxf(fabsf(sg[2]), 1e-14f));
        float tr = e0 + e1 + e2;
        float h0 = e0 - tr / 3.0f, h1 = e1 - tr / 3.0f, h2 = e2 - tr / 3.0f;
        float ehn = sqrtf(h0 * h0 + h1 * h1 + h2 * h2);
        float dg = ehn + (3.0f * lam + 2.0f * mu) / (2.0f * mu) * tr * P.alpha;
        if (dg > 0.0f) {
            if (tr > 0.0f) {
                sg[0] = 1.0f; sg[1] = 1.0f; sg[2] = 1.0f;
            } else {
                float k = dg / ehn;
                sg[0] = expf(e0 - h0 * k); sg[1] = expf(e1 - h1 * k); sg[2] = expf(e2 - h2 * k);
            }
            u_diag_vt(U, sg, V, F);
        }
        float l0 = logf(sg[0]), l1 = logf(sg[1]), l2 = logf(sg[2]);  // no clamp: NaN for sigma <= 0 as in :50-53
        float ls = l0 + l1 + l2;
        t[0] = 2.0f * mu * l0 + lam * ls; t[1] = 2.0f * mu * l1 + lam * ls; t[2] = 2.0f * mu * l2 + lam * ls;
        sym_from_u(U, t, s6);
        return;
    }

    // foam (constitutive_models.py:217-259) and fluid (:143-213, never dispatched by the reference)
    {
        bool fluid = (mat == MPMK_MAT_FLUID);
        float c0, c1, c2;
        if (fluid) { c0 = fmaxf(fabsf(sg[0]), 0.01f); c1 = fmaxf(fabsf(sg[1]), 0.01f); c2 = fmaxf(fabsf(sg[2]), 0.01f); }
        else { c0 = fmaxf(sg[0], 0.01f); c1 = fmaxf(sg[1], 0.01f); c2 = fmaxf(sg[2], 0.01f); }
        float e0 = logf(c0), e1 = logf(c1), e2 = logf(c2);
        float tr = e0 + e1 + e2;
        float h0 = e0 - tr / 3.0f, h1 = e1 - tr / 3.0f, h2 = e2 - tr / 3.0f;
        float a0 = 2.0f * mu * h0, a1 = 2.0f * mu * h1, a2 = 2.0f * mu * h2;
        float stn = sqrtf(a0 * a0 + a1 * a1 + a2 * a2);
        float y = stn - (fluid ? 1.0f : 0.8f) * sqrtf(2.0f / 3.0f) * ys;
        if (y > 0.0f) {
            float bsum = fluid ? (sg[0] * sg[0] + sg[1] * sg[1] + sg[2] * sg[2]) : (c0 * c0 + c1 * c1 + c2 * c2);
            float mu_hat = mu * bsum / 3.0f;
            float pf = fluid ? (1.0f + P.plastic_viscosity / (2.0f * mu_hat * dt))
                             : (1.0f + P.plastic_viscosity * 2.0f / (2.0f * mu_hat * dt));
            float snn = stn - y / pf;
            float k = snn / stn;
            float inv2mu = 1.0f / (2.0f * mu);
            sg[0] = expf(inv2mu * (k * a0) + tr / 3.0f);
            sg[1] = expf(inv2mu * (k * a1) + tr / 3.0f);
            sg[2] = expf(inv2mu * (k * a2) + tr / 3.0f);
            if (!fluid && (P.flags & MPMK_STRICT_Q3)) {
                // :256 `U * sig_elastic * V.transpose()` is element-wise in Taichi: only the diagonal survives
#pragma unroll
                for (int i = 0; i < 9; ++i) F[i] = 0.0f;
                F[0] = U[0] * sg[0] * V[0]; F[4] = U[4] * sg[1] * V[4]; F[8] = U[8] * sg[2] * V[8];
                svd3(F, U, sg, V);  // the result is no longer U diag V^T: decompose again (utils.py:33)
            } else {
                u_diag_vt(U, sg, V, F);
            }
        }
        stvk_principal(sg, mu, lam, t);
        sym_from_u(U, t, s6);
    }
}
