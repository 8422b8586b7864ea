// opt.cu -- the differentiable ("fitting") path of the reference: forward kernels with a substep history and their
// hand-written adjoints, replacing Taichi's autodiff of
//     compute_stress_from_F_opt   mpm_solver/utils.py:56-76      (St. Venant-Kirchhoff on the Green strain, Cauchy stress)
//     p2g_opt                     mpm_solver/utils.py:137-174
//     grid_normalization_and_gravity + grid_postprocess[0].apply  utils.py:178-183, boundary_conditions.py:23-27
//     g2p_opt                     mpm_solver/utils.py:285-347
//     compute_cov_from_F_opt      mpm_solver/utils.py:436-467
//     compute_mu_lam_from_E_nu    mpm_solver/utils.py:350-362    (adjoint only; forward is export.cu k_mu_lam)
//     MPM_Simulator.learn         mpm_solver/solver.py:92-108
// driven by solver.py:54-90 (p2g2p_forward / p2g2p_backward) and extra.py:199-236.
//
// State layout: plain row-major arrays in ORIGINAL particle order, history index first, exactly the reference's
// fields: xyz/vel [S+1][N][3], F/C/stress [S+1][N][9] (model.py:147-151), dense grids [n^3][3] + mass [n^3]
// (model.py:171-174).  This path runs 30 substeps of a few hundred thousand particles per training step
// (extra.py:199), keeps every substep for the backward sweep, and is not the substep hot path: one thread per
// particle, float atomics to the dense grid, no sort -- the reference's own structure.
//
// Adjoint conventions follow what Taichi 1.5's reverse mode does with these kernels, including where that is not the
// textbook gradient (SURVEY.md 9.7, quirk register):
//   * grid_mass, particle_mass, particle_vol have no adjoint (fields without needs_grad, model.py:143-145,172);
//   * a global store of a constant (the boundary condition's grid_v_out[I] = 0) leaves the adjoint of the overwritten
//     entry in place, so the gradient of a stuck node still flows back through the normalisation;
//   * the |J| < 1e-2 clamp of the stress makes J a constant in that branch.
#include "common.cuh"

#define OPT_DEV __device__ __forceinline__

struct OptWeights {
    float w[3][3], dw[3][3];   // [axis][node]
    int base[3];
    float fx[3];
};

OPT_DEV void opt_weights(const float* x, float inv_dx, int n_grid, OptWeights& W) {
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float gp = __fmul_rn(x[d], inv_dx);
        int b = (int)__fsub_rn(gp, 0.5f);
        b = max(0, min(b, n_grid - 3));   // memory-safety clamp of this build (quirk Q9)
        const float f = gp - (float)b;
        W.base[d] = b; W.fx[d] = f;
        const float wa = 1.5f - f, wb = f - 1.0f, wc = f - 0.5f;
        W.w[d][0] = wa * wa * 0.5f; W.w[d][1] = 0.75f - wb * wb; W.w[d][2] = wc * wc * 0.5f;
        W.dw[d][0] = f - 1.5f; W.dw[d][1] = -2.0f * (f - 1.0f); W.dw[d][2] = f - 0.5f;
    }
}

// d w[d][i] / d fx[d] and d dw[d][i] / d fx[d]
OPT_DEV void opt_weight_derivs(const OptWeights& W, float dwf[3][3], float ddwf[3][3]) {
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float f = W.fx[d];
        dwf[d][0] = -(1.5f - f); dwf[d][1] = -2.0f * (f - 1.0f); dwf[d][2] = f - 0.5f;
        ddwf[d][0] = 1.0f; ddwf[d][1] = -2.0f; ddwf[d][2] = 1.0f;
    }
}

OPT_DEV void mat_mul3(const float* A, const float* B, float* C) {
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) C[3 * r + c] = A[3 * r] * B[c] + A[3 * r + 1] * B[3 + c] + A[3 * r + 2] * B[6 + c];
}
OPT_DEV void mat_T3(const float* A, float* T) {
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) T[3 * r + c] = A[3 * c + r];
}
OPT_DEV float det3(const float* F) {
    return F[0] * (F[4] * F[8] - F[5] * F[7]) - F[1] * (F[3] * F[8] - F[5] * F[6]) + F[2] * (F[3] * F[7] - F[4] * F[6]);
}

// ---- utils.py:56-76 ------------------------------------------------------------------------------------
OPT_DEV void opt_stress_parts(const float* F, float mu, float lam, float* E, float* S, float* A, float& J, float& Jc, bool& clamped) {
    J = det3(F);
    Jc = J; clamped = false;
    if (fabsf(J) < 1e-2f) { Jc = 1e-2f * (J > 0.f ? 1.f : (J < 0.f ? -1.f : 0.f)); clamped = true; }
    float Ft[9], FtF[9];
    mat_T3(F, Ft);
    mat_mul3(Ft, F, FtF);
#pragma unroll
    for (int i = 0; i < 9; ++i) E[i] = 0.5f * (FtF[i] - ((i % 4 == 0) ? 1.0f : 0.0f));
    const float tr = E[0] + E[4] + E[8];
#pragma unroll
    for (int i = 0; i < 9; ++i) S[i] = 2.0f * mu * E[i] + ((i % 4 == 0) ? lam * tr : 0.0f);
    float FS[9];
    mat_mul3(F, S, FS);
    mat_mul3(FS, Ft, A);   // A = F S F^T, sigma = A / J
}

__global__ void k_opt_stress(int64_t n, const float* __restrict__ F, const float* __restrict__ mu, const float* __restrict__ lam,
                             float* __restrict__ stress) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    float Fp[9], E[9], S[9], A[9], J, Jc;
    bool cl;
#pragma unroll
    for (int i = 0; i < 9; ++i) Fp[i] = F[9 * p + i];
    opt_stress_parts(Fp, mu[p], lam[p], E, S, A, J, Jc, cl);
#pragma unroll
    for (int i = 0; i < 9; ++i) stress[9 * p + i] = A[i] / Jc;
}

// adjoint: stress.grad[s] -> F.grad[s] (accumulated), mu.grad, lam.grad (accumulated)
__global__ void k_opt_stress_grad(int64_t n, const float* __restrict__ F, const float* __restrict__ mu, const float* __restrict__ lam,
                                  const float* __restrict__ g_stress, float* __restrict__ g_F, float* __restrict__ g_mu,
                                  float* __restrict__ g_lam) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    float Fp[9], E[9], S[9], A[9], J, Jc, G[9];
    bool cl;
#pragma unroll
    for (int i = 0; i < 9; ++i) { Fp[i] = F[9 * p + i]; G[i] = g_stress[9 * p + i]; }
    const float m = mu[p], l = lam[p];
    opt_stress_parts(Fp, m, l, E, S, A, J, Jc, cl);
    // sigma = A / Jc
    float Ab[9], gJ = 0.f;
#pragma unroll
    for (int i = 0; i < 9; ++i) { Ab[i] = G[i] / Jc; gJ -= G[i] * A[i]; }
    gJ = cl ? 0.f : gJ / (Jc * Jc);
    // A = F S F^T:  Fbar += Ab F S^T + Ab^T F S ;  Sbar = F^T Ab F
    float Ft[9], AbT[9], St[9], T1[9], T2[9], Fb[9], Sb[9];
    mat_T3(Fp, Ft); mat_T3(Ab, AbT); mat_T3(S, St);
    mat_mul3(Ab, Fp, T1); mat_mul3(T1, St, Fb);
    mat_mul3(AbT, Fp, T1); mat_mul3(T1, S, T2);
#pragma unroll
    for (int i = 0; i < 9; ++i) Fb[i] += T2[i];
    mat_mul3(Ft, Ab, T1); mat_mul3(T1, Fp, Sb);
    // S = 2 mu E + lam tr(E) I
    const float trS = Sb[0] + Sb[4] + Sb[8], trE = E[0] + E[4] + E[8];
    float Eb[9], gm = 0.f;
#pragma unroll
    for (int i = 0; i < 9; ++i) { Eb[i] = 2.0f * m * Sb[i] + ((i % 4 == 0) ? l * trS : 0.0f); gm += 2.0f * Sb[i] * E[i]; }
    const float gl = trE * trS;
    // E = (F^T F - I) / 2:  Fbar += F (Eb + Eb^T) / 2
    float Es[9];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) Es[3 * r + c] = 0.5f * (Eb[3 * r + c] + Eb[3 * c + r]);
    mat_mul3(Fp, Es, T1);
#pragma unroll
    for (int i = 0; i < 9; ++i) Fb[i] += T1[i];
    // J = det F:  Fbar += gJ * cof(F)
    const float cof[9] = {Fp[4] * Fp[8] - Fp[5] * Fp[7], Fp[5] * Fp[6] - Fp[3] * Fp[8], Fp[3] * Fp[7] - Fp[4] * Fp[6],
                          Fp[2] * Fp[7] - Fp[1] * Fp[8], Fp[0] * Fp[8] - Fp[2] * Fp[6], Fp[1] * Fp[6] - Fp[0] * Fp[7],
                          Fp[1] * Fp[5] - Fp[2] * Fp[4], Fp[2] * Fp[3] - Fp[0] * Fp[5], Fp[0] * Fp[4] - Fp[1] * Fp[3]};
#pragma unroll
    for (int i = 0; i < 9; ++i) g_F[9 * p + i] += Fb[i] + gJ * cof[i];
    g_mu[p] += gm;
    g_lam[p] += gl;
}

// ---- utils.py:137-174 ----------------------------------------------------------------------------------
__global__ void k_opt_p2g(int64_t n, const float* __restrict__ x, const float* __restrict__ v, const float* __restrict__ C,
                          const float* __restrict__ stress, const float* __restrict__ mass, const float* __restrict__ vol,
                          float* __restrict__ gvin, float* __restrict__ gmass, int n_grid, float dx, float inv_dx, float dt) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    OptWeights W;
    const float xp[3] = {x[3 * p], x[3 * p + 1], x[3 * p + 2]};
    opt_weights(xp, inv_dx, n_grid, W);
    float Cp[9], Sp[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) { Cp[i] = C[9 * p + i]; Sp[i] = stress[9 * p + i]; }
    const float m = mass[p], vl = vol[p];
    const float vp[3] = {v[3 * p], v[3 * p + 1], v[3 * p + 2]};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            for (int k = 0; k < 3; ++k) {
                const float dpos[3] = {((float)i - W.fx[0]) * dx, ((float)j - W.fx[1]) * dx, ((float)k - W.fx[2]) * dx};
                const float weight = W.w[0][i] * W.w[1][j] * W.w[2][k];
                const float dwt[3] = {W.dw[0][i] * W.w[1][j] * W.w[2][k] * inv_dx, W.w[0][i] * W.dw[1][j] * W.w[2][k] * inv_dx,
                                      W.w[0][i] * W.w[1][j] * W.dw[2][k] * inv_dx};
                const size_t node = ((size_t)(W.base[0] + i) * n_grid + (W.base[1] + j)) * n_grid + (W.base[2] + k);
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float ef = -vl * (Sp[3 * c] * dwt[0] + Sp[3 * c + 1] * dwt[1] + Sp[3 * c + 2] * dwt[2]);
                    const float add = weight * m * (vp[c] + Cp[3 * c] * dpos[0] + Cp[3 * c + 1] * dpos[1] + Cp[3 * c + 2] * dpos[2]) + dt * ef;
                    atomicAdd(&gvin[3 * node + c], add);
                }
                atomicAdd(&gmass[node], weight * m);
            }
}

// adjoint: grid_v_in.grad -> vel.grad[s], C.grad[s], stress.grad[s], xyz.grad[s] (all accumulated)
__global__ void k_opt_p2g_grad(int64_t n, const float* __restrict__ x, const float* __restrict__ v, const float* __restrict__ C,
                               const float* __restrict__ stress, const float* __restrict__ mass, const float* __restrict__ vol,
                               const float* __restrict__ g_gvin, float* __restrict__ g_x, float* __restrict__ g_v, float* __restrict__ g_C,
                               float* __restrict__ g_stress, int n_grid, float dx, float inv_dx, float dt) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    OptWeights W;
    const float xp[3] = {x[3 * p], x[3 * p + 1], x[3 * p + 2]};
    opt_weights(xp, inv_dx, n_grid, W);
    float dwf[3][3], ddwf[3][3];
    opt_weight_derivs(W, dwf, ddwf);
    float Cp[9], Sp[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) { Cp[i] = C[9 * p + i]; Sp[i] = stress[9 * p + i]; }
    const float m = mass[p], vl = vol[p];
    const float vp[3] = {v[3 * p], v[3 * p + 1], v[3 * p + 2]};
    float gv[3] = {0, 0, 0}, gC[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, gS[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, gfx[3] = {0, 0, 0};
    float gw[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}, gdw[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};   // adjoints of the 1-D weights
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            for (int k = 0; k < 3; ++k) {
                const float dpos[3] = {((float)i - W.fx[0]) * dx, ((float)j - W.fx[1]) * dx, ((float)k - W.fx[2]) * dx};
                const float wx = W.w[0][i], wy = W.w[1][j], wz = W.w[2][k];
                const float weight = wx * wy * wz;
                const size_t node = ((size_t)(W.base[0] + i) * n_grid + (W.base[1] + j)) * n_grid + (W.base[2] + k);
                const float g[3] = {g_gvin[3 * node], g_gvin[3 * node + 1], g_gvin[3 * node + 2]};
                float gweight = 0.f, gdwt[3] = {0, 0, 0};
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float aff = vp[c] + Cp[3 * c] * dpos[0] + Cp[3 * c + 1] * dpos[1] + Cp[3 * c + 2] * dpos[2];
                    gweight += m * g[c] * aff;
                    gv[c] += weight * m * g[c];
#pragma unroll
                    for (int d = 0; d < 3; ++d) {
                        gC[3 * c + d] += weight * m * g[c] * dpos[d];
                        gfx[d] -= dx * weight * m * g[c] * Cp[3 * c + d];        // dpos_d = (off_d - fx_d) dx
                        gdwt[d] -= dt * vl * g[c] * Sp[3 * c + d];              // elastic force = -vol stress dweight
                    }
                }
                // weight = wx wy wz ; dweight = inv_dx (dwx wy wz, wx dwy wz, wx wy dwz)
                const float dwx = W.dw[0][i], dwy = W.dw[1][j], dwz = W.dw[2][k];
                const float a0 = gdwt[0] * inv_dx, a1 = gdwt[1] * inv_dx, a2 = gdwt[2] * inv_dx;
                gw[0][i] += gweight * wy * wz + a1 * dwy * wz + a2 * wy * dwz;
                gw[1][j] += gweight * wx * wz + a0 * dwx * wz + a2 * wx * dwz;
                gw[2][k] += gweight * wx * wy + a0 * dwx * wy + a1 * wx * dwy;
                gdw[0][i] += a0 * wy * wz;
                gdw[1][j] += a1 * wx * wz;
                gdw[2][k] += a2 * wx * wy;
                // stress adjoint: ef_c = -vol sum_d S_cd dwt_d, contribution dt * ef
                const float dwt[3] = {dwx * wy * wz * inv_dx, wx * dwy * wz * inv_dx, wx * wy * dwz * inv_dx};
#pragma unroll
                for (int c = 0; c < 3; ++c)
#pragma unroll
                    for (int d = 0; d < 3; ++d) gS[3 * c + d] -= dt * vl * g[c] * dwt[d];
            }
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        float t = gfx[d];
#pragma unroll
        for (int i = 0; i < 3; ++i) t += gw[d][i] * dwf[d][i] + gdw[d][i] * ddwf[d][i];
        g_x[3 * p + d] += t * inv_dx;    // fx = x inv_dx - base
        g_v[3 * p + d] += gv[d];
    }
#pragma unroll
    for (int i = 0; i < 9; ++i) { g_C[9 * p + i] += gC[i]; g_stress[9 * p + i] += gS[i]; }
}

// ---- utils.py:178-183 + boundary_conditions.py:23-27 ---------------------------------------------------
// v_out = v_in / m + dt g where m > 1e-15 (other nodes keep the 0 of reset_grid_state), then grid_postprocess[0]:
// nodes inside the box |idx dx - center| < size get v_out = 0.  use_box = 0: no box.
__global__ void k_opt_grid(const float* __restrict__ gvin, const float* __restrict__ gmass, float* __restrict__ gvout, int n_grid,
                           float dx, float dt, float gx, float gy, float gz, int use_box, float cx, float cy, float cz, float sx, float sy,
                           float sz) {
    const int64_t total = (int64_t)n_grid * n_grid * n_grid;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const float m = gmass[i];
        float vx = 0.f, vy = 0.f, vz = 0.f;
        if (m > 1e-15f) {
            vx = __fdiv_rn(gvin[3 * i], m) + dt * gx;
            vy = __fdiv_rn(gvin[3 * i + 1], m) + dt * gy;
            vz = __fdiv_rn(gvin[3 * i + 2], m) + dt * gz;
        }
        if (use_box) {
            const int iz = (int)(i % n_grid), iy = (int)((i / n_grid) % n_grid), ix = (int)(i / ((int64_t)n_grid * n_grid));
            const float px = __fmul_rn((float)ix, dx), py = __fmul_rn((float)iy, dx), pz = __fmul_rn((float)iz, dx);
            if (fabsf(__fsub_rn(px, cx)) < sx && fabsf(__fsub_rn(py, cy)) < sy && fabsf(__fsub_rn(pz, cz)) < sz) { vx = 0.f; vy = 0.f; vz = 0.f; }
        }
        gvout[3 * i] = vx; gvout[3 * i + 1] = vy; gvout[3 * i + 2] = vz;
    }
}

// adjoint: grid_v_in.grad += grid_v_out.grad / m where m > 1e-15.  Reference-literal (block_box = 0): the boundary
// condition's store of a constant has no adjoint and leaves grid_v_out.grad in place (see the header), so stuck nodes pass
// their gradient on like any other.  block_box = 1: the textbook adjoint, nothing flows through a node the box zeroed.
__global__ void k_opt_grid_grad(const float* __restrict__ gmass, const float* __restrict__ g_gvout, float* __restrict__ g_gvin, int n_grid,
                                float dx, int block_box, float cx, float cy, float cz, float sx, float sy, float sz) {
    const int64_t total = (int64_t)n_grid * n_grid * n_grid;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const float m = gmass[i];
        if (!(m > 1e-15f)) continue;
        if (block_box) {
            const int iz = (int)(i % n_grid), iy = (int)((i / n_grid) % n_grid), ix = (int)(i / ((int64_t)n_grid * n_grid));
            const float px = __fmul_rn((float)ix, dx), py = __fmul_rn((float)iy, dx), pz = __fmul_rn((float)iz, dx);
            if (fabsf(__fsub_rn(px, cx)) < sx && fabsf(__fsub_rn(py, cy)) < sy && fabsf(__fsub_rn(pz, cz)) < sz) continue;
        }
        g_gvin[3 * i] += g_gvout[3 * i] / m;
        g_gvin[3 * i + 1] += g_gvout[3 * i + 1] / m;
        g_gvin[3 * i + 2] += g_gvout[3 * i + 2] / m;
    }
}

// ---- utils.py:285-347 ----------------------------------------------------------------------------------
__global__ void k_opt_g2p(int64_t n, const float* __restrict__ x, const float* __restrict__ F, const float* __restrict__ gvout,
                          float* __restrict__ x1, float* __restrict__ v1, float* __restrict__ C1, float* __restrict__ F1, int n_grid,
                          float inv_dx, float dt) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    OptWeights W;
    const float xp[3] = {x[3 * p], x[3 * p + 1], x[3 * p + 2]};
    opt_weights(xp, inv_dx, n_grid, W);
    float nv[3] = {0, 0, 0}, nC[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, nF[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            for (int k = 0; k < 3; ++k) {
                const float dpos[3] = {(float)i - W.fx[0], (float)j - W.fx[1], (float)k - W.fx[2]};
                const float weight = W.w[0][i] * W.w[1][j] * W.w[2][k];
                const float dwt[3] = {W.dw[0][i] * W.w[1][j] * W.w[2][k] * inv_dx, W.w[0][i] * W.dw[1][j] * W.w[2][k] * inv_dx,
                                      W.w[0][i] * W.w[1][j] * W.dw[2][k] * inv_dx};
                const size_t node = ((size_t)(W.base[0] + i) * n_grid + (W.base[1] + j)) * n_grid + (W.base[2] + k);
                const float g[3] = {gvout[3 * node], gvout[3 * node + 1], gvout[3 * node + 2]};
                const float wc = weight * inv_dx * 4.0f;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    nv[c] += g[c] * weight;
#pragma unroll
                    for (int d = 0; d < 3; ++d) { nC[3 * c + d] += g[c] * dpos[d] * wc; nF[3 * c + d] += g[c] * dwt[d]; }
                }
            }
    float Fp[9], M[9], Fn[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) { Fp[i] = F[9 * p + i]; M[i] = ((i % 4 == 0) ? 1.0f : 0.0f) + nF[i] * dt; }
    mat_mul3(M, Fp, Fn);
#pragma unroll
    for (int d = 0; d < 3; ++d) { v1[3 * p + d] = nv[d]; x1[3 * p + d] = xp[d] + dt * nv[d]; }
#pragma unroll
    for (int i = 0; i < 9; ++i) { C1[9 * p + i] = nC[i]; F1[9 * p + i] = Fn[i]; }
}

// adjoint: {xyz, vel, C, F}.grad[s+1] -> grid_v_out.grad (atomics), xyz.grad[s], F.grad[s] (accumulated)
__global__ void k_opt_g2p_grad(int64_t n, const float* __restrict__ x, const float* __restrict__ F, const float* __restrict__ gvout,
                               const float* __restrict__ g_x1, const float* __restrict__ g_v1, const float* __restrict__ g_C1,
                               const float* __restrict__ g_F1, float* __restrict__ g_gvout, float* __restrict__ g_x, float* __restrict__ g_F,
                               int n_grid, float inv_dx, float dt) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    OptWeights W;
    const float xp[3] = {x[3 * p], x[3 * p + 1], x[3 * p + 2]};
    opt_weights(xp, inv_dx, n_grid, W);
    float dwf[3][3], ddwf[3][3];
    opt_weight_derivs(W, dwf, ddwf);
    // forward quantity needed again: new_F (for F[s+1] = (I + dt new_F) F[s])
    float nF[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            for (int k = 0; k < 3; ++k) {
                const float dwt[3] = {W.dw[0][i] * W.w[1][j] * W.w[2][k] * inv_dx, W.w[0][i] * W.dw[1][j] * W.w[2][k] * inv_dx,
                                      W.w[0][i] * W.w[1][j] * W.dw[2][k] * inv_dx};
                const size_t node = ((size_t)(W.base[0] + i) * n_grid + (W.base[1] + j)) * n_grid + (W.base[2] + k);
#pragma unroll
                for (int c = 0; c < 3; ++c)
#pragma unroll
                    for (int d = 0; d < 3; ++d) nF[3 * c + d] += gvout[3 * node + c] * dwt[d];
            }
    float Fp[9], GF1[9], GC[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) { Fp[i] = F[9 * p + i]; GF1[i] = g_F1[9 * p + i]; GC[i] = g_C1[9 * p + i]; }
    // x[s+1] = x[s] + dt new_v ; v[s+1] = new_v
    float gnv[3], gx1[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) { gx1[d] = g_x1[3 * p + d]; gnv[d] = g_v1[3 * p + d] + dt * gx1[d]; }
    // F[s+1] = M F[s], M = I + dt new_F:  Fbar[s] += M^T Fbar[s+1] ; new_Fbar = dt Fbar[s+1] F[s]^T
    float M[9], Mt[9], Ft[9], gFs[9], gnF[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) M[i] = ((i % 4 == 0) ? 1.0f : 0.0f) + nF[i] * dt;
    mat_T3(M, Mt); mat_T3(Fp, Ft);
    mat_mul3(Mt, GF1, gFs);
    mat_mul3(GF1, Ft, gnF);
#pragma unroll
    for (int i = 0; i < 9; ++i) gnF[i] *= dt;
    float gfx[3] = {0, 0, 0};
    float gw[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}, gdw[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            for (int k = 0; k < 3; ++k) {
                const float dpos[3] = {(float)i - W.fx[0], (float)j - W.fx[1], (float)k - W.fx[2]};
                const float wx = W.w[0][i], wy = W.w[1][j], wz = W.w[2][k];
                const float dwx = W.dw[0][i], dwy = W.dw[1][j], dwz = W.dw[2][k];
                const float weight = wx * wy * wz, wc = weight * inv_dx * 4.0f;
                const float dwt[3] = {dwx * wy * wz * inv_dx, wx * dwy * wz * inv_dx, wx * wy * dwz * inv_dx};
                const size_t node = ((size_t)(W.base[0] + i) * n_grid + (W.base[1] + j)) * n_grid + (W.base[2] + k);
                const float g[3] = {gvout[3 * node], gvout[3 * node + 1], gvout[3 * node + 2]};
                float gweight = 0.f, gdwt[3] = {0, 0, 0};
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    // new_v_c += g_c weight ; new_C_cd += g_c dpos_d wc ; new_F_cd += g_c dwt_d
                    float gg = gnv[c] * weight;
                    float cd = 0.f;
#pragma unroll
                    for (int d = 0; d < 3; ++d) {
                        gg += GC[3 * c + d] * dpos[d] * wc + gnF[3 * c + d] * dwt[d];
                        cd += GC[3 * c + d] * dpos[d];
                        gfx[d] -= GC[3 * c + d] * g[c] * wc;      // dpos_d = off_d - fx_d
                        gdwt[d] += gnF[3 * c + d] * g[c];
                    }
                    atomicAdd(&g_gvout[3 * node + c], gg);
                    gweight += g[c] * (gnv[c] + cd * inv_dx * 4.0f);
                }
                const float a0 = gdwt[0] * inv_dx, a1 = gdwt[1] * inv_dx, a2 = gdwt[2] * inv_dx;
                gw[0][i] += gweight * wy * wz + a1 * dwy * wz + a2 * wy * dwz;
                gw[1][j] += gweight * wx * wz + a0 * dwx * wz + a2 * wx * dwz;
                gw[2][k] += gweight * wx * wy + a0 * dwx * wy + a1 * wx * dwy;
                gdw[0][i] += a0 * wy * wz;
                gdw[1][j] += a1 * wx * wz;
                gdw[2][k] += a2 * wx * wy;
            }
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        float t = gfx[d];
#pragma unroll
        for (int i = 0; i < 3; ++i) t += gw[d][i] * dwf[d][i] + gdw[d][i] * ddwf[d][i];
        g_x[3 * p + d] += gx1[d] + t * inv_dx;
    }
#pragma unroll
    for (int i = 0; i < 9; ++i) g_F[9 * p + i] += gFs[i];
}

// ---- utils.py:436-467 and its adjoint -----------------------------------------------------------------
__global__ void k_opt_cov(int64_t n, const float* __restrict__ F, const float* __restrict__ init_cov, float* __restrict__ cov) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    float Fp[9], A[9], Ft[9], T[9], O[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) Fp[i] = F[9 * p + i];
    const float* a = init_cov + 6 * p;
    A[0] = a[0]; A[1] = a[1]; A[2] = a[2]; A[3] = a[1]; A[4] = a[3]; A[5] = a[4]; A[6] = a[2]; A[7] = a[4]; A[8] = a[5];
    mat_T3(Fp, Ft); mat_mul3(Fp, A, T); mat_mul3(T, Ft, O);
    float* c = cov + 6 * p;
    c[0] = O[0]; c[1] = O[1]; c[2] = O[2]; c[3] = O[4]; c[4] = O[5]; c[5] = O[8];
}

// cov = F A F^T, only the six stored entries carry a gradient: Fbar += G F A^T + G^T F A with G the upper-triangular image
__global__ void k_opt_cov_grad(int64_t n, const float* __restrict__ F, const float* __restrict__ init_cov, const float* __restrict__ g_cov,
                               float* __restrict__ g_F) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    float Fp[9], A[9], G[9], Gt[9], T[9], O1[9], O2[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) Fp[i] = F[9 * p + i];
    const float* a = init_cov + 6 * p;
    A[0] = a[0]; A[1] = a[1]; A[2] = a[2]; A[3] = a[1]; A[4] = a[3]; A[5] = a[4]; A[6] = a[2]; A[7] = a[4]; A[8] = a[5];
    const float* g = g_cov + 6 * p;
    G[0] = g[0]; G[1] = g[1]; G[2] = g[2]; G[3] = 0.f; G[4] = g[3]; G[5] = g[4]; G[6] = 0.f; G[7] = 0.f; G[8] = g[5];
    mat_T3(G, Gt);
    mat_mul3(G, Fp, T); mat_mul3(T, A, O1);    // A symmetric: A^T = A
    mat_mul3(Gt, Fp, T); mat_mul3(T, A, O2);
#pragma unroll
    for (int i = 0; i < 9; ++i) g_F[9 * p + i] += O1[i] + O2[i];
}

// ---- adjoint of utils.py:350-362: (mu.grad, lam.grad) -> (logE.grad, y.grad), accumulated -----------------
__global__ void k_opt_mu_lam_grad(int64_t n, const float* __restrict__ logE, const float* __restrict__ y, const float* __restrict__ g_mu,
                                  const float* __restrict__ g_lam, float* __restrict__ g_logE, float* __restrict__ g_y) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const float E = powf(10.0f, logE[p]);
    const float ey = expf(-y[p]);
    const float nu = 0.49f / (1.0f + ey);
    const float a = 1.0f + nu, b = 1.0f - 2.0f * nu;
    const float dmu_dE = 1.0f / (2.0f * a), dmu_dnu = -E / (2.0f * a * a);
    const float dlam_dE = nu / (a * b), dlam_dnu = E * (1.0f + 2.0f * nu * nu) / (a * a * b * b);
    const float dE = 2.302585092994046f * E, dnu = 0.49f * ey / ((1.0f + ey) * (1.0f + ey));
    const float gm = g_mu[p], gl = g_lam[p];
    g_logE[p] += (gm * dmu_dE + gl * dlam_dE) * dE;
    g_y[p] += (gm * dmu_dnu + gl * dlam_dnu) * dnu;
}

// ---- solver.py:92-108 ----------------------------------------------------------------------------------
__global__ void k_opt_learn(int64_t n, float* __restrict__ logE, float* __restrict__ y, const float* __restrict__ g_logE,
                            const float* __restrict__ g_y) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    float a = g_logE[p], b = g_y[p];
    if (fabsf(a) > 1.0f) a = (a > 0.f) ? 1.0f : -1.0f;
    if (fabsf(b) > 1.0f) b = (b > 0.f) ? 1.0f : -1.0f;
    logE[p] -= 0.8f * a;
    y[p] -= 1.6f * b;
}

// ---- C ABI ------------------------------------------------------------------------------------------------
#define OPT_GRID(n) div_up((n), 128), 128, 0, (cudaStream_t)stream
#define OPT_REQ(ctx) do { if (!(ctx)) return MPMK_ERR_INVALID; } while (0)

extern "C" int mpmk_opt_stress(mpmk_ctx* ctx, int64_t n, const float* F, const float* mu, const float* lam, float* stress, void* stream) {
    OPT_REQ(ctx);
    if (n > 0) k_opt_stress<<<OPT_GRID(n)>>>(n, F, mu, lam, stress);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}
extern "C" int mpmk_opt_stress_grad(mpmk_ctx* ctx, int64_t n, const float* F, const float* mu, const float* lam, const float* g_stress,
                                    float* g_F, float* g_mu, float* g_lam, void* stream) {
    OPT_REQ(ctx);
    if (n > 0) k_opt_stress_grad<<<OPT_GRID(n)>>>(n, F, mu, lam, g_stress, g_F, g_mu, g_lam);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}
extern "C" int mpmk_opt_p2g(mpmk_ctx* ctx, int64_t n, const float* x, const float* v, const float* C, const float* stress, const float* mass,
                            const float* vol, float* grid_v_in, float* grid_mass, float dt, void* stream) {
    OPT_REQ(ctx);
    if (n > 0) k_opt_p2g<<<OPT_GRID(n)>>>(n, x, v, C, stress, mass, vol, grid_v_in, grid_mass, ctx->n_grid, ctx->dx, ctx->inv_dx, dt);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}
extern "C" int mpmk_opt_p2g_grad(mpmk_ctx* ctx, int64_t n, const float* x, const float* v, const float* C, const float* stress,
                                 const float* mass, const float* vol, const float* g_grid_v_in, float* g_x, float* g_v, float* g_C,
                                 float* g_stress, float dt, void* stream) {
    OPT_REQ(ctx);
    if (n > 0) k_opt_p2g_grad<<<OPT_GRID(n)>>>(n, x, v, C, stress, mass, vol, g_grid_v_in, g_x, g_v, g_C, g_stress, ctx->n_grid, ctx->dx, ctx->inv_dx, dt);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}
extern "C" int mpmk_opt_grid(mpmk_ctx* ctx, const float* grid_v_in, const float* grid_mass, float* grid_v_out, float dt, const float* gravity3,
                             const float* box_center3, const float* box_size3, void* stream) {
    OPT_REQ(ctx);
    if (!gravity3) return MPMK_ERR_INVALID;
    const int use_box = (box_center3 && box_size3) ? 1 : 0;
    const float z3[3] = {0.f, 0.f, 0.f};
    const float* c = use_box ? box_center3 : z3;
    const float* s = use_box ? box_size3 : z3;
    k_opt_grid<<<ctx->sm_count * 8, 256, 0, (cudaStream_t)stream>>>(grid_v_in, grid_mass, grid_v_out, ctx->n_grid, ctx->dx, dt, gravity3[0], gravity3[1],
                                                                    gravity3[2], use_box, c[0], c[1], c[2], s[0], s[1], s[2]);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}
extern "C" int mpmk_opt_grid_grad(mpmk_ctx* ctx, const float* grid_mass, const float* g_grid_v_out, float* g_grid_v_in,
                                  const float* block_center3, const float* block_size3, void* stream) {
    OPT_REQ(ctx);
    const int block = (block_center3 && block_size3) ? 1 : 0;
    const float z3[3] = {0.f, 0.f, 0.f};
    const float* c = block ? block_center3 : z3;
    const float* s = block ? block_size3 : z3;
    k_opt_grid_grad<<<ctx->sm_count * 8, 256, 0, (cudaStream_t)stream>>>(grid_mass, g_grid_v_out, g_grid_v_in, ctx->n_grid, ctx->dx, block, c[0], c[1],
                                                                         c[2], s[0], s[1], s[2]);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}
extern "C" int mpmk_opt_g2p(mpmk_ctx* ctx, int64_t n, const float* x, const float* F, const float* grid_v_out, float* x1, float* v1, float* C1,
                            float* F1, float dt, void* stream) {
    OPT_REQ(ctx);
    if (n > 0) k_opt_g2p<<<OPT_GRID(n)>>>(n, x, F, grid_v_out, x1, v1, C1, F1, ctx->n_grid, ctx->inv_dx, dt);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}
extern "C" int mpmk_opt_g2p_grad(mpmk_ctx* ctx, int64_t n, const float* x, const float* F, const float* grid_v_out, const float* g_x1,
                                 const float* g_v1, const float* g_C1, const float* g_F1, float* g_grid_v_out, float* g_x, float* g_F, float dt,
                                 void* stream) {
    OPT_REQ(ctx);
    if (n > 0) k_opt_g2p_grad<<<OPT_GRID(n)>>>(n, x, F, grid_v_out, g_x1, g_v1, g_C1, g_F1, g_grid_v_out, g_x, g_F, ctx->n_grid, ctx->inv_dx, dt);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}
extern "C" int mpmk_opt_cov(mpmk_ctx* ctx, int64_t n, const float* F, const float* init_cov, float* cov, void* stream) {
    OPT_REQ(ctx);
    if (n > 0) k_opt_cov<<<OPT_GRID(n)>>>(n, F, init_cov, cov);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}
extern "C" int mpmk_opt_cov_grad(mpmk_ctx* ctx, int64_t n, const float* F, const float* init_cov, const float* g_cov, float* g_F, void* stream) {
    OPT_REQ(ctx);
    if (n > 0) k_opt_cov_grad<<<OPT_GRID(n)>>>(n, F, init_cov, g_cov, g_F);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}
extern "C" int mpmk_opt_mu_lam_grad(mpmk_ctx* ctx, int64_t n, const float* logE, const float* y, const float* g_mu, const float* g_lam,
                                    float* g_logE, float* g_y, void* stream) {
    OPT_REQ(ctx);
    if (n > 0) k_opt_mu_lam_grad<<<OPT_GRID(n)>>>(n, logE, y, g_mu, g_lam, g_logE, g_y);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}
extern "C" int mpmk_opt_learn(mpmk_ctx* ctx, int64_t n, float* logE, float* y, const float* g_logE, const float* g_y, void* stream) {
    OPT_REQ(ctx);
    if (n > 0) k_opt_learn<<<OPT_GRID(n)>>>(n, logE, y, g_logE, g_y);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}
