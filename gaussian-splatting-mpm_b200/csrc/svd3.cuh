// svd3.cuh -- 3x3 fp32 SVD and polar rotation standing in for ti.svd (call sites:
// mpm_solver/utils.py:33,385; constitutive_models.py:64,107,158,218).
//
// svd3: published algorithm (McAdams, Selle, Tamstorf, Teran, Sifakis 2011): cyclic Jacobi on A^T A
// with approximate Givens quaternions, columns of A V sorted by decreasing norm, then QR.  Contract:
// U, V in SO(3); |sigma| descending; sign(det A) carried by sigma[2].  The sweep loop leaves as soon
// as the off-diagonal mass of A^T A is at the fp32 rounding floor (SVD3_TOL of its trace squared, the
// trace being invariant): at most SVD3_SWEEPS sweeps run, 3-4 on typical deformation gradients, with
// the same reconstruction error as the fixed five (tests/test_hostmath.py).
//
// polar3: the rotation R = U V^T alone (all the FCR stress of constitutive_models.py:11-20 needs) by
// Newton's iteration X <- (X + X^-T)/2, valid for det A > 0; ~50 instructions per iteration against
// ~900 for the SVD.  Matrices are row-major m[3*r+c].
#pragma once
#include "devmath.cuh"

#define SVD3_GAMMA 5.828427124746190f
#define SVD3_CSTAR 0.923879532511287f
#define SVD3_SSTAR 0.382683432365090f
#define SVD3_SWEEPS 5
#ifndef SVD3_TOL
#define SVD3_TOL 1e-13f
#endif
#ifdef SVD3_COUNT
static long svd3_sweep_count = 0;
#endif

MPMK_DEV void svd3_jacobi(float& spp, float& sqq, float& spq, float& spr, float& sqr, float& vp, float& vq, float& vr,
                          float& qw) {
    float ch = 2.0f * (spp - sqq);
    float sh = spq;
    bool b = (SVD3_GAMMA * sh * sh) < (ch * ch);
    float w = rsqrtf(ch * ch + sh * sh);
    ch = b ? w * ch : SVD3_CSTAR;
    sh = b ? w * sh : SVD3_SSTAR;
    float c = ch * ch - sh * sh;
    float sn = 2.0f * ch * sh;
    // rows, then columns, of the (p,q) block of Q^T S Q
    float a = c * spp + sn * spq;
    float bq = c * spq + sn * sqq;
    float e = c * spq - sn * spp;
    float d = c * sqq - sn * spq;
    float npr = c * spr + sn * sqr;
    float nqr = c * sqr - sn * spr;
    spp = c * a + sn * bq;
    spq = c * bq - sn * a;
    sqq = c * d - sn * e;
    spr = npr; sqr = nqr;
    float op = vp, oq = vq, orr = vr, ow = qw;
    qw = ow * ch - sh * orr;
    vp = ch * op + sh * oq;
    vq = ch * oq - sh * op;
    vr = ch * orr + sh * ow;
}

MPMK_DEV void svd3_cswap(float* B, float* V, float& na, float& nb, int a, int b) {
    // new col a = old col b ; new col b = -old col a (keeps det V = +1)
    bool sw = na < nb;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        float ba = B[3 * r + a], bb = B[3 * r + b], va = V[3 * r + a], vb = V[3 * r + b];
        B[3 * r + a] = sw ? bb : ba;
        B[3 * r + b] = sw ? -ba : bb;
        V[3 * r + a] = sw ? vb : va;
        V[3 * r + b] = sw ? -va : vb;
    }
    float t = na;
    na = sw ? nb : na;
    nb = sw ? t : nb;
}

MPMK_DEV void svd3(const float* A, float* U, float* sig, float* V) {
    float s00 = A[0] * A[0] + A[3] * A[3] + A[6] * A[6];
    float s01 = A[0] * A[1] + A[3] * A[4] + A[6] * A[7];
    float s02 = A[0] * A[2] + A[3] * A[5] + A[6] * A[8];
    float s11 = A[1] * A[1] + A[4] * A[4] + A[7] * A[7];
    float s12 = A[1] * A[2] + A[4] * A[5] + A[7] * A[8];
    float s22 = A[2] * A[2] + A[5] * A[5] + A[8] * A[8];
    float qx = 0.0f, qy = 0.0f, qz = 0.0f, qw = 1.0f;
    const float tr = s00 + s11 + s22;  // invariant under the conjugations
    const float tol = SVD3_TOL * tr * tr;
#pragma unroll 1
    for (int it = 0; it < SVD3_SWEEPS; ++it) {
        svd3_jacobi(s00, s11, s01, s02, s12, qx, qy, qz, qw);  // (p,q,r) = (0,1,2)
        svd3_jacobi(s11, s22, s12, s01, s02, qy, qz, qx, qw);  // (1,2,0)
        svd3_jacobi(s22, s00, s02, s12, s01, qz, qx, qy, qw);  // (2,0,1)
#ifdef SVD3_COUNT
        ++svd3_sweep_count;
#endif
        if ((s01 * s01 + s02 * s02 + s12 * s12) < tol) break;
    }
    float nq = rsqrtf(qw * qw + qx * qx + qy * qy + qz * qz);
    float x = qx * nq, y = qy * nq, z = qz * nq, w = qw * nq;
    V[0] = 1.0f - 2.0f * (y * y + z * z); V[1] = 2.0f * (x * y - w * z); V[2] = 2.0f * (x * z + w * y);
    V[3] = 2.0f * (x * y + w * z); V[4] = 1.0f - 2.0f * (x * x + z * z); V[5] = 2.0f * (y * z - w * x);
    V[6] = 2.0f * (x * z - w * y); V[7] = 2.0f * (y * z + w * x); V[8] = 1.0f - 2.0f * (x * x + y * y);
    float B[9];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) B[3 * r + c] = A[3 * r] * V[c] + A[3 * r + 1] * V[3 + c] + A[3 * r + 2] * V[6 + c];
    float n0 = B[0] * B[0] + B[3] * B[3] + B[6] * B[6];
    float n1 = B[1] * B[1] + B[4] * B[4] + B[7] * B[7];
    float n2 = B[2] * B[2] + B[5] * B[5] + B[8] * B[8];
    svd3_cswap(B, V, n0, n1, 0, 1);
    svd3_cswap(B, V, n0, n2, 0, 2);
    svd3_cswap(B, V, n1, n2, 1, 2);
    // Gram-Schmidt QR
    float s0 = sqrtf(n0);
    float u0x, u0y, u0z;
    if (s0 > 1e-30f) { float r0 = 1.0f / s0; u0x = B[0] * r0; u0y = B[3] * r0; u0z = B[6] * r0; }
    else { u0x = 1.0f; u0y = 0.0f; u0z = 0.0f; }
    float d = u0x * B[1] + u0y * B[4] + u0z * B[7];
    float tx = B[1] - d * u0x, ty = B[4] - d * u0y, tz = B[7] - d * u0z;
    float s1 = sqrtf(tx * tx + ty * ty + tz * tz);
    float u1x, u1y, u1z;
    if (s1 > 1e-30f) { float r1 = 1.0f / s1; u1x = tx * r1; u1y = ty * r1; u1z = tz * r1; }
    else {
        float ex = fabsf(u0x) < 0.57f ? 1.0f : 0.0f, ey = 1.0f - ex;
        float de = u0x * ex + u0y * ey;
        tx = ex - de * u0x; ty = ey - de * u0y; tz = -de * u0z;
        float re = rsqrtf(tx * tx + ty * ty + tz * tz);
        u1x = tx * re; u1y = ty * re; u1z = tz * re;
        s1 = 0.0f;
    }
    float u2x = u0y * u1z - u0z * u1y;
    float u2y = u0z * u1x - u0x * u1z;
    float u2z = u0x * u1y - u0y * u1x;
    float s2 = u2x * B[2] + u2y * B[5] + u2z * B[8];
    U[0] = u0x; U[1] = u1x; U[2] = u2x;
    U[3] = u0y; U[4] = u1y; U[5] = u2y;
    U[6] = u0z; U[7] = u1z; U[8] = u2z;
    sig[0] = s0; sig[1] = s1; sig[2] = s2;
}

// One Newton step X <- (X + X^-T)/2 = X/2 + cof(X) / (2 det X); returns |X'|_F^2.
MPMK_DEV float polar3_step(float* X) {
    const float c0 = X[4] * X[8] - X[5] * X[7];
    const float c1 = X[5] * X[6] - X[3] * X[8];
    const float c2 = X[3] * X[7] - X[4] * X[6];
    const float c3 = X[2] * X[7] - X[1] * X[8];
    const float c4 = X[0] * X[8] - X[2] * X[6];
    const float c5 = X[1] * X[6] - X[0] * X[7];
    const float c6 = X[1] * X[5] - X[2] * X[4];
    const float c7 = X[2] * X[3] - X[0] * X[5];
    const float c8 = X[0] * X[4] - X[1] * X[3];
    const float det = X[0] * c0 + X[1] * c1 + X[2] * c2;
    const float h = 0.5f * MPMK_FAST_RCP(det);
    X[0] = 0.5f * X[0] + h * c0; X[1] = 0.5f * X[1] + h * c1; X[2] = 0.5f * X[2] + h * c2;
    X[3] = 0.5f * X[3] + h * c3; X[4] = 0.5f * X[4] + h * c4; X[5] = 0.5f * X[5] + h * c5;
    X[6] = 0.5f * X[6] + h * c6; X[7] = 0.5f * X[7] + h * c7; X[8] = 0.5f * X[8] + h * c8;
    return X[0] * X[0] + X[1] * X[1] + X[2] * X[2] + X[3] * X[3] + X[4] * X[4] + X[5] * X[5] + X[6] * X[6] +
           X[7] * X[7] + X[8] * X[8];
}

// Rotation factor of A = R S (S symmetric positive definite), requires det A > 0.  The singular values
// of the iterate map as s -> (s + 1/s)/2 >= 1, so |X|_F^2 - 3 = sum(s_i^2 - 1) >= 0 measures the
// distance to SO(3); once it is below 6e-4 (every s_i within 3e-4 of 1) one more quadratically
// convergent step leaves an error of ~5e-8.  Returns false if the iteration did not converge.
#define POLAR3_MAX_IT 48
MPMK_DEV bool polar3(const float* A, float* R) {
#pragma unroll
    for (int i = 0; i < 9; ++i) R[i] = A[i];
    bool ok = false;
#pragma unroll 1
    for (int it = 0; it < POLAR3_MAX_IT; ++it) {
        const float n2 = polar3_step(R);
        if (fabsf(n2 - 3.0f) < 6e-4f) { ok = true; break; }
    }
    polar3_step(R);
    return ok;
}
