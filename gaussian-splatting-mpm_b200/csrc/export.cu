// export.cu -- init-time kernels and the per-frame export.
//   k_export: solver.py:135-137 postprocess() = utils.py:402-433 compute_cov_from_F +
//     utils.py:377-398 compute_R_from_F, fused with the scatter back to ORIGINAL particle order that
//     replaces the reference's Field.to_torch() host round trip (main.py:311-312).
//   k_mu_lam: utils.py:350-362.   k_vol_count / k_vol: internel_filling/filling.py:12-24.
#include "common.cuh"
#include "svd3.cuh"

__global__ void k_export(const float4* __restrict__ xm, const float4* __restrict__ f0, const float4* __restrict__ f1,
                         const float4* __restrict__ f2, const int4* __restrict__ meta, const float* __restrict__ init_cov,
                         float* __restrict__ oxyz, float* __restrict__ ocov, float* __restrict__ oR, int64_t n) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const int64_t id = meta[r].x;
    if (oxyz) {
        const float4 X = xm[r];
        oxyz[3 * id] = X.x; oxyz[3 * id + 1] = X.y; oxyz[3 * id + 2] = X.z;
    }
    if (!ocov && !oR) return;
    const float4 F0 = f0[r], F1 = f1[r], F2 = f2[r];
    const float F[9] = {F0.x, F0.y, F0.z, F0.w, F1.x, F1.y, F1.z, F1.w, F2.x};
    if (ocov) {
        const float* ic = init_cov + 6 * id;
        const float c0[9] = {ic[0], ic[1], ic[2], ic[1], ic[3], ic[4], ic[2], ic[4], ic[5]};
        float t[9];
#pragma unroll
        for (int rr = 0; rr < 3; ++rr)
#pragma unroll
            for (int c = 0; c < 3; ++c) t[3 * rr + c] = F[3 * rr] * c0[c] + F[3 * rr + 1] * c0[3 + c] + F[3 * rr + 2] * c0[6 + c];
        float o[9];
#pragma unroll
        for (int rr = 0; rr < 3; ++rr)
#pragma unroll
            for (int c = 0; c < 3; ++c) o[3 * rr + c] = t[3 * rr] * F[3 * c] + t[3 * rr + 1] * F[3 * c + 1] + t[3 * rr + 2] * F[3 * c + 2];
        float* oc = ocov + 6 * id;
        oc[0] = o[0]; oc[1] = o[1]; oc[2] = o[2]; oc[3] = o[4]; oc[4] = o[5]; oc[5] = o[8];
    }
    if (oR) {
        float U[9], sg[3], V[9];
        svd3(F, U, sg, V);  // det U = det V = +1 by contract, so the reference's det fix-ups (:387-395) are no-ops
        float* R = oR + 9 * id;
        // R = U V^T, stored transposed (:397-398): R^T = V U^T
#pragma unroll
        for (int rr = 0; rr < 3; ++rr)
#pragma unroll
            for (int c = 0; c < 3; ++c) R[3 * rr + c] = V[3 * rr] * U[3 * c] + V[3 * rr + 1] * U[3 * c + 1] + V[3 * rr + 2] * U[3 * c + 2];
    }
}

// postprocess() + grid2world + the undo chain of render_frame in one pass: position and covariance
// of every particle in ORIGINAL order, in the frame the rasteriser wants:
//     p = (x * alpha + beta) M          (main.py:310-313 grid2world, :139-143 undo shift / scale / rotations)
//     S = M^T (gamma F S0 F^T) M        (utils.py:402-433, main.py:145-146)
// M = product of the pre-processing rotations in inverse order (transform_utils.py:106-123); alpha, beta, gamma are
// composed on the host (mpm_solver: utils/transform_utils.py world_affine).
struct WorldXform {
    float alpha, gamma;
    float beta[3];
    float M[9];
};

__global__ void k_export_world(const float4* __restrict__ xm, const float4* __restrict__ f0, const float4* __restrict__ f1,
                               const float4* __restrict__ f2, const int4* __restrict__ meta, const float* __restrict__ init_cov,
                               float* __restrict__ oxyz, float* __restrict__ ocov, int64_t n, const WorldXform w) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const int64_t id = meta[r].x;
    const float4 X = xm[r];
    const float q[3] = {X.x * w.alpha + w.beta[0], X.y * w.alpha + w.beta[1], X.z * w.alpha + w.beta[2]};
#pragma unroll
    for (int c = 0; c < 3; ++c) oxyz[3 * id + c] = q[0] * w.M[c] + q[1] * w.M[3 + c] + q[2] * w.M[6 + c];
    if (!ocov) return;
    const float4 F0 = f0[r], F1 = f1[r], F2 = f2[r];
    const float F[9] = {F0.x, F0.y, F0.z, F0.w, F1.x, F1.y, F1.z, F1.w, F2.x};
    const float* ic = init_cov + 6 * id;
    const float c0[9] = {ic[0], ic[1], ic[2], ic[1], ic[3], ic[4], ic[2], ic[4], ic[5]};
    float t[9], o[9];
#pragma unroll
    for (int rr = 0; rr < 3; ++rr)
#pragma unroll
        for (int c = 0; c < 3; ++c) t[3 * rr + c] = F[3 * rr] * c0[c] + F[3 * rr + 1] * c0[3 + c] + F[3 * rr + 2] * c0[6 + c];
#pragma unroll
    for (int rr = 0; rr < 3; ++rr)
#pragma unroll
        for (int c = 0; c < 3; ++c)
            o[3 * rr + c] = (t[3 * rr] * F[3 * c] + t[3 * rr + 1] * F[3 * c + 1] + t[3 * rr + 2] * F[3 * c + 2]) * w.gamma;
    // M^T o M, upper triangle
#pragma unroll
    for (int rr = 0; rr < 3; ++rr)
#pragma unroll
        for (int c = 0; c < 3; ++c) t[3 * rr + c] = o[3 * rr] * w.M[c] + o[3 * rr + 1] * w.M[3 + c] + o[3 * rr + 2] * w.M[6 + c];
    float* oc = ocov + 6 * id;
    int k = 0;
#pragma unroll
    for (int rr = 0; rr < 3; ++rr)
#pragma unroll
        for (int c = rr; c < 3; ++c) oc[k++] = w.M[rr] * t[c] + w.M[3 + rr] * t[3 + c] + w.M[6 + rr] * t[6 + c];
}

__global__ void k_mu_lam(const float* __restrict__ logE, const float* __restrict__ y, float* __restrict__ mu,
                         float* __restrict__ lam, int64_t n) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const float E = powf(10.0f, logE[p]);
    const float nu = 0.49f / (1.0f + expf(-y[p]));
    mu[p] = E / (2.0f * (1.0f + nu));
    lam[p] = E * nu / ((1.0f + nu) * (1.0f - 2.0f * nu));
}

__device__ __forceinline__ int64_t vol_cell(const float* pos, int64_t p, float grid_dx, int n) {
    int cx = (int)floorf(__fdiv_rn(pos[3 * p], grid_dx)), cy = (int)floorf(__fdiv_rn(pos[3 * p + 1], grid_dx)),
        cz = (int)floorf(__fdiv_rn(pos[3 * p + 2], grid_dx));
    cx = min(max(cx, 0), n - 1); cy = min(max(cy, 0), n - 1); cz = min(max(cz, 0), n - 1);
    return ((int64_t)cx * n + cy) * n + cz;
}
__global__ void k_vol_count(const float* __restrict__ pos, int64_t n_p, float grid_dx, int n, int* __restrict__ cnt) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n_p) atomicAdd(&cnt[vol_cell(pos, p, grid_dx, n)], 1);
}
__global__ void k_vol(const float* __restrict__ pos, int64_t n_p, float grid_dx, int n, const int* __restrict__ cnt,
                      float* __restrict__ vol) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n_p) vol[p] = __fdiv_rn(grid_dx * grid_dx * grid_dx, (float)cnt[vol_cell(pos, p, grid_dx, n)]);
}

extern "C" int mpmk_export(mpmk_ctx* ctx, const void* init_cov_dev, void* xyz_out, void* cov_out, void* R_out, void* stream) {
    if (!ctx || !ctx->bound) return MPMK_ERR_STATE;
    if (ctx->n == 0) return 0;
    if (cov_out && !init_cov_dev) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "mpmk_export: cov_out needs init_cov");
    const mpmk_particle_set& A = ctx->b.set[ctx->cur];
    const bool ft = ctx->ftrial_valid;  // F_trial of the last g2p if it was stored, else F (identical for jelly)
    k_export<<<div_up(ctx->n, 128), 128, 0, (cudaStream_t)stream>>>(
        (const float4*)A.xm, (const float4*)(ft ? ctx->b.ft0 : A.f0), (const float4*)(ft ? ctx->b.ft1 : A.f1),
        (const float4*)(ft ? ctx->b.ft2 : A.f2), (const int4*)A.meta, (const float*)init_cov_dev, (float*)xyz_out,
        (float*)cov_out, (float*)R_out, ctx->n);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}

extern "C" int mpmk_export_world(mpmk_ctx* ctx, const void* init_cov_dev, const mpmk_world_xform* xf, void* xyz_out,
                                 void* cov_out, void* stream) {
    if (!ctx || !ctx->bound) return MPMK_ERR_STATE;
    if (!xf || !xyz_out) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "mpmk_export_world: xf and xyz_out are required");
    if (ctx->n == 0) return 0;
    if (cov_out && !init_cov_dev) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "mpmk_export_world: cov_out needs init_cov");
    const mpmk_particle_set& A = ctx->b.set[ctx->cur];
    const bool ft = ctx->ftrial_valid;
    WorldXform w;
    w.alpha = xf->alpha; w.gamma = xf->gamma;
    for (int i = 0; i < 3; ++i) w.beta[i] = xf->beta[i];
    for (int i = 0; i < 9; ++i) w.M[i] = xf->M[i];
    k_export_world<<<div_up(ctx->n, 128), 128, 0, (cudaStream_t)stream>>>(
        (const float4*)A.xm, (const float4*)(ft ? ctx->b.ft0 : A.f0), (const float4*)(ft ? ctx->b.ft1 : A.f1),
        (const float4*)(ft ? ctx->b.ft2 : A.f2), (const int4*)A.meta, (const float*)init_cov_dev, (float*)xyz_out,
        (float*)cov_out, ctx->n, w);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}

extern "C" int mpmk_compute_mu_lam(mpmk_ctx* ctx, int64_t n, const void* logE, const void* y, void* mu, void* lam, void* stream) {
    if (!ctx) return MPMK_ERR_STATE;
    if (n == 0) return 0;
    k_mu_lam<<<div_up(n, 256), 256, 0, (cudaStream_t)stream>>>((const float*)logE, (const float*)y, (float*)mu, (float*)lam, n);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}

extern "C" int mpmk_particle_volume(mpmk_ctx* ctx, int64_t n, const void* pos, float grid_dx, void* cell_count, void* vol,
                                    void* stream) {
    if (!ctx) return MPMK_ERR_STATE;
    if (n == 0) return 0;
    k_vol_count<<<div_up(n, 256), 256, 0, (cudaStream_t)stream>>>((const float*)pos, n, grid_dx, ctx->n_grid, (int*)cell_count);
    MPMK_LAUNCH_CHECK(ctx);
    k_vol<<<div_up(n, 256), 256, 0, (cudaStream_t)stream>>>((const float*)pos, n, grid_dx, ctx->n_grid, (const int*)cell_count,
                                                            (float*)vol);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}

// ---- mpmk_particle_volume_noalloc: the same two kernels with the cell-count scratch inside the library ---------------
// No counterpart in the reference; added for callers that cannot allocate (code
// under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The scratch is a file-scope device
// array of a stated maximum size -- every GPU that loads the library holds these 8 MiB until the process exits -- and the
// entry point calls no cudaMalloc / cudaFree / allocating library routine.  All calls share the array and it keeps the
// last call's contents: it is cleared on `stream` before the count, and two calls that overlap on the GPU (different
// streams, no ordering between them) race.
#define MPMK_VOL_NOALLOC_MAX_GRID 128   // cells per edge: 128^3 int32 = 8 MiB (the reference's configs use 50^3 and 100^3)
static __device__ __align__(256) int g_vol_cells[MPMK_VOL_NOALLOC_MAX_GRID * MPMK_VOL_NOALLOC_MAX_GRID * MPMK_VOL_NOALLOC_MAX_GRID];

// In host code `g_vol_cells` is a host-side stand-in, not a device address: the address a kernel can use comes from
// cudaGetSymbolAddress, looked up once per device with that device current.
#define MPMK_NOALLOC_MAX_DEVICES 64
static int* g_vol_cells_addr[MPMK_NOALLOC_MAX_DEVICES];

static int vol_cells_for_device(mpmk_ctx* ctx, int** out) {
    int cur = -1;
    MPMK_CHECK(ctx, cudaGetDevice(&cur));
    if (cur != ctx->device) MPMK_FAIL(ctx, MPMK_ERR_STATE, "mpmk_particle_volume_noalloc: device %d is current, the context lives on device %d", cur, ctx->device);
    if (cur < 0 || cur >= MPMK_NOALLOC_MAX_DEVICES) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "mpmk_particle_volume_noalloc: device ordinal %d beyond %d", cur, MPMK_NOALLOC_MAX_DEVICES);
    if (!g_vol_cells_addr[cur]) {
        int* p = nullptr;
        MPMK_CHECK(ctx, cudaGetSymbolAddress((void**)&p, g_vol_cells));
        if (!p) MPMK_FAIL(ctx, MPMK_ERR_STATE, "mpmk_particle_volume_noalloc: cudaGetSymbolAddress returned no address");
        g_vol_cells_addr[cur] = p;
    }
    *out = g_vol_cells_addr[cur];
    return 0;
}

extern "C" int mpmk_particle_volume_noalloc(mpmk_ctx* ctx, int64_t n, const void* pos, float grid_dx, void* vol, void* stream) {
    if (!ctx) return MPMK_ERR_STATE;
    if (n < 0 || (n > 0 && (!pos || !vol))) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "mpmk_particle_volume_noalloc: bad arguments");
    if (ctx->n_grid > MPMK_VOL_NOALLOC_MAX_GRID)
        MPMK_FAIL(ctx, MPMK_ERR_INVALID, "mpmk_particle_volume_noalloc: n_grid = %d exceeds the built-in scratch (max %d cells per edge); use mpmk_particle_volume with a caller-provided cell_count buffer",
                  ctx->n_grid, MPMK_VOL_NOALLOC_MAX_GRID);
    if (n == 0) return 0;
    int* cells = nullptr;
    const int rc = vol_cells_for_device(ctx, &cells);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const size_t n_cells = (size_t)ctx->n_grid * ctx->n_grid * ctx->n_grid;
    MPMK_CHECK(ctx, cudaMemsetAsync(cells, 0, n_cells * sizeof(int), s));   // the array still holds the last call's counts
    k_vol_count<<<div_up(n, 256), 256, 0, s>>>((const float*)pos, n, grid_dx, ctx->n_grid, cells);
    MPMK_LAUNCH_CHECK(ctx);
    k_vol<<<div_up(n, 256), 256, 0, s>>>((const float*)pos, n, grid_dx, ctx->n_grid, cells, (float*)vol);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}
