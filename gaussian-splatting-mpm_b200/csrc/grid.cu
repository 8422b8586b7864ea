// grid.cu -- K3: one fused pass over the node blocks p2g touched:
//   utils.py:178-183 grid_normalization_and_gravity, then the registered grid operations in
//   registration order (solver.py:41-46): boundary_conditions.py:23-27 BasicBC.apply (fixed_cube) and
//   collider.py:13-44 MPM_Collider.collide.  Nodes p2g never touched keep v_out = 0, on which both
//   operations are no-ops, so visiting only touched blocks is bit-equivalent to the dense passes.
#include "common.cuh"

struct GridOpsArg {
    int n_ops;
    mpmk_grid_op ops[MPMK_MAX_GRID_OPS];
};

__device__ __forceinline__ void grid_node_update(float4* __restrict__ gin, float4* __restrict__ gout, int ix, int iy,
                                                 int iz, int n, float dx, float dt, float gx, float gy, float gz,
                                                 const GridOpsArg& ops, int zero_in) {
    size_t node = ((size_t)ix * n + iy) * n + iz;
    float4 m = gin[node];
    if (zero_in) gin[node] = make_float4(0.f, 0.f, 0.f, 0.f);   // fused batches: ready for the next substep's scatter
    float vx = 0.f, vy = 0.f, vz = 0.f;
    if (m.w > 1e-15f) {
        vx = __fdiv_rn(m.x, m.w) + dt * gx;
        vy = __fdiv_rn(m.y, m.w) + dt * gy;
        vz = __fdiv_rn(m.z, m.w) + dt * gz;
    }
    // geometric predicates are discontinuous: keep every product and difference individually rounded
    // (no FMA contraction) so a node lying exactly on a box face / collider plane falls on the same
    // side as in the reference arithmetic (float(idx) * dx - c)
    const float px = __fmul_rn((float)ix, dx), py = __fmul_rn((float)iy, dx), pz = __fmul_rn((float)iz, dx);
    for (int o = 0; o < ops.n_ops; ++o) {
        const mpmk_grid_op& op = ops.ops[o];
        if (op.type == MPMK_OP_FIXED_CUBE) {
            if (fabsf(__fsub_rn(px, op.a[0])) < op.b[0] && fabsf(__fsub_rn(py, op.a[1])) < op.b[1] &&
                fabsf(__fsub_rn(pz, op.a[2])) < op.b[2]) {
                vx = 0.f; vy = 0.f; vz = 0.f;
            }
        } else {
            const float ox = __fsub_rn(px, op.a[0]), oy = __fsub_rn(py, op.a[1]), oz = __fsub_rn(pz, op.a[2]);
            const float nx = op.b[0], ny = op.b[1], nz = op.b[2];
            const float dp = __fadd_rn(__fadd_rn(__fmul_rn(ox, nx), __fmul_rn(oy, ny)), __fmul_rn(oz, nz));
            if (dp < 0.f) {
                float nc = vx * nx + vy * ny + vz * nz;
                float mn = fminf(nc, 0.f);
                vx -= mn * nx; vy -= mn * ny; vz -= mn * nz;
                float len = sqrtf(vx * vx + vy * vy + vz * vz);
                if (nc < 0.f && len > 1e-20f) {
                    float sc = fmaxf(0.f, len + nc * op.c);
                    vx = sc * __fdiv_rn(vx, len); vy = sc * __fdiv_rn(vy, len); vz = sc * __fdiv_rn(vz, len);
                }
                vx *= 0.99f; vy *= 0.99f; vz *= 0.99f;  // collider.py:44, every node below the plane (Q4)
            }
        }
    }
    gout[node] = make_float4(vx, vy, vz, 0.f);
}

__global__ void __launch_bounds__(256) k_grid(float4* __restrict__ gin, float4* __restrict__ gout,
                                              const uint32_t* __restrict__ list, const uint32_t* __restrict__ count,
                                              int n, int nb, float dx, float dt, float gx, float gy, float gz,
                                              const GridOpsArg ops, int zero_in) {
    pdl_wait();   // programmatic dependent launch: nothing of the previous grid is read or overwritten before this
    const int64_t total = (int64_t)(*count) * 64;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        uint32_t b = list[i >> 6];
        int l = (int)(i & 63);
        int bz = b % nb, by = (b / nb) % nb, bx = b / (nb * nb);
        int ix = 4 * bx + (l >> 4), iy = 4 * by + ((l >> 2) & 3), iz = 4 * bz + (l & 3);
        if (ix >= n || iy >= n || iz >= n) continue;
        grid_node_update(gin, gout, ix, iy, iz, n, dx, dt, gx, gy, gz, ops, zero_in);
    }
}

// multi-GPU: the same update, densely over the (up to two) plane ranges shared with the neighbouring slabs (mg.cu),
// rows [y0, y1) x nodes [z0, z1)
__global__ void __launch_bounds__(256) k_grid_planes(float4* __restrict__ gin, float4* __restrict__ gout, int xa0, int xa1, int xb0, int xb1,
                                                     int y0, int y1, int z0, int z1, int n, float dx, float dt, float gx, float gy, float gz,
                                                     const GridOpsArg ops) {
    const int nz = z1 - z0;
    const int64_t per_plane = (int64_t)(y1 - y0) * nz;
    const int64_t ta = per_plane * (xa1 - xa0), total = ta + per_plane * (xb1 - xb0);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const bool second = i >= ta;
        const int64_t j = second ? i - ta : i;
        const int ix = (second ? xb0 : xa0) + (int)(j / per_plane);
        const int64_t rem = j % per_plane;
        const int iy = y0 + (int)(rem / nz), iz = z0 + (int)(rem % nz);
        // zeroes grid_in behind itself: the shared planes collect sums from the neighbours on nodes no local block list
        // knows about, so they are reset here, densely, instead of by a separate clear before the next scatter
        grid_node_update(gin, gout, ix, iy, iz, n, dx, dt, gx, gy, gz, ops, 1);
    }
}

static int launch_grid(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, cudaStream_t s, bool zero_in) {
    GridOpsArg a;
    a.n_ops = 0;
    if (ops) {
        if (ops->n_ops < 0 || ops->n_ops > MPMK_MAX_GRID_OPS) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "n_ops out of range");
        a.n_ops = ops->n_ops;
        memcpy(a.ops, ops->ops, sizeof(mpmk_grid_op) * ops->n_ops);
    }
    mpmk_launch(k_grid, dim3(ctx->sm_count * 8), dim3(256), 0, s, (float4*)ctx->b.grid_in, (float4*)ctx->b.grid_out,
                                             ctx->block_list[ctx->cur_list], ctx->block_count[ctx->cur_list],
                                             ctx->n_grid, ctx->nb, ctx->dx, dt, ctx->prm.gravity[0], ctx->prm.gravity[1],
                                             ctx->prm.gravity[2], a, (zero_in || ctx->zero_grid_in) ? 1 : 0);
    MPMK_LAUNCH_CHECK(ctx);
    if (ctx->mg.enabled) {
        const MgState& m = ctx->mg;
        if (m.lo1 > m.lo0 || m.hi1 > m.hi0) {
            // (the z window only narrows the peer-memory path; the NCCL path exchanges whole rows)
            const int z0 = m.peer ? m.z_lo : 0, z1 = m.peer ? m.z_hi : ctx->n_grid;
            k_grid_planes<<<ctx->sm_count * 4, 256, 0, s>>>((float4*)ctx->b.grid_in, (float4*)ctx->b.grid_out, m.lo0, m.lo1, m.hi0, m.hi1,
                                                            m.y_lo, m.y_hi, z0, z1, ctx->n_grid, ctx->dx, dt, ctx->prm.gravity[0],
                                                            ctx->prm.gravity[1], ctx->prm.gravity[2], a);
            MPMK_LAUNCH_CHECK(ctx);
        }
    }
    return 0;
}

int mpmk_launch_grid(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, cudaStream_t s) {
    if (!ctx->prepared) MPMK_FAIL(ctx, MPMK_ERR_STATE, "mpmk_grid_update: prepare_grid first");
    return launch_grid(ctx, dt, ops, s, false);
}

// lagged batches: no `prepared` state to check (the lists were built a substep ago)
int mpmk_launch_grid_lagged(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, cudaStream_t s) { return launch_grid(ctx, dt, ops, s, false); }
