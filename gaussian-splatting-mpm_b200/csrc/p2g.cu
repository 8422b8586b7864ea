// p2g.cu -- K2: APIC particle-to-grid scatter (mpm_solver/utils.py:90-134) with the impulse
// pre-process (boundary_conditions.py:41-45) folded into its prologue.
//
// One warp owns one work item = up to MPMK_ITEM_MAX consecutive sorted particles of one 4x4x4-cell
// tile, and a private 6x6x6-node float4 tile {px,py,pz,m} in shared memory.  Per batch of 32 particles:
//   phase A (lane = particle): coalesced float4 loads through the sort permutation, quadratic B-spline
//     weights, and the tensor-product partial sums
//         Y_ij = wy_j (wx_i (q + i A0) + dwx_i S0) + wx_i (wy_j j A1 + dwy_j S1),   R_k = wz_k k A2 + dwz_k S2
//     with A = m C dx, S = -dt vol inv_dx stress, q = m v - A fx, staged in shared memory;
//     (R_k is stored divided by the particle mass and Y_ij carries m wx_i wy_j in its 4th lane);
//   phase B (lane = stencil node (i,j,k)): walks the batch in sorted order; node value
//         wz_k Y_ij + (m wx_i wy_j) R_k / m,  mass (m wx_i wy_j) wz_k
//     accumulates in registers while consecutive particles share a cell -- run heads are the lanes whose cell
//     differs from their predecessor's, collected with one ballot per batch -- and is
//     added to the shared tile once per (cell run, node): no shared-memory atomics, every lane owns a
//     distinct node.
// The tile is flushed with one 16-byte red.global.add.v4.f32 per non-zero node.
//
// LAG = true (lagged batches, api.cu): the work list was built from the sort of the PREVIOUS substep and the physical
// order is that sorted order, so a batch is 32 consecutive physical slots (no permutation) and a particle's cell
// comes from its current key.  A particle that changed cell inside the item's tile simply starts a run of its own; one
// whose cell left the tile ("crosser", ~1 % per substep) is accumulated the same way and its 27 node values go straight
// to the global grid, with the node blocks they touch added to this substep's block list.
#include "common.cuh"
#include <cstdlib>

#ifndef P2G_WARPS
#define P2G_WARPS 4
#endif
#ifndef P2G_MIN_CTAS
#define P2G_MIN_CTAS 5   // caps the kernel at 96 registers (113 without: four CTAs would then leave nothing for the side branch of a lagged batch)
#endif
#define P2G_REC_STRIDE 33  // float4 per record slot: 32 lanes + 1 pad (bank-conflict-free phase B reads)
#define P2G_REC_SLOTS 12   // 9 x {Y_ij.xyz, wx_i wy_j} + 3 x {R_k.xyz, wz_k}

#define P2G_TILE_SX 37     // x stride of the shared tile in float4 (6*6 + 1 pad)
#define P2G_TILE_F4 (6 * P2G_TILE_SX)

// lane -> stencil node (i*9 + j*3 + k).  Shared memory serves a 16-byte access per quarter-warp, 8 float4 bank
// groups wide: the map puts into every quarter-warp nodes whose tile slots (37 i + 6 j + k) and whose Y_ij record
// slots (3 i + j) fall into distinct bank groups, so neither the tile read-modify-write nor the record reads
// replay.  Lanes 27..31 mirror lane 26 (same addresses: a broadcast) and never write.
__constant__ unsigned char P2G_LANE_NODE[32] = {2, 1, 11, 19, 9, 10, 21, 13, 24, 18, 4, 15, 5, 14, 6, 12, 3, 23, 7, 22, 25, 17, 20, 26, 8, 16, 0, 0, 0, 0, 0, 0};

struct __align__(16) P2GWarpSmem {
    float4 rec[P2G_REC_SLOTS * P2G_REC_STRIDE];
    float4 tile[P2G_TILE_F4];
};

#define P2G_CROSS 0x40000000   // cell code of a crosser: P2G_CROSS | bx | by << 10 | bz << 20 (base cell, n_grid <= 1023)

struct P2GArgs {
    const uint32_t* perm;    // LAG: unused
    const uint32_t* skeys;   // LAG: the current keys, in physical order
    // LAG: this substep's block list, for the blocks only crossers reach
    uint32_t* stamp_arr;
    uint32_t* blist;
    uint32_t* bcount;
    int nb;
    const float4 *xm, *vv, *c0, *c1, *cs, *s1;
    const uint4* items;
    DevCounters* cnt;
    float4* grid_in;
    int n_grid, nt;
    float dx, inv_dx, dt;
    int n_imp;
    mpmk_impulse imp[MPMK_MAX_IMPULSES];
};

// one cell run's accumulated node value -> the warp's tile, or (crossers) the global grid
template <bool LAG>
__device__ __forceinline__ void p2g_flush_run(const P2GArgs& a, float4* tile, int cur_cell, int noff, int ni, int nj, int nk,
                                              const float4& acc, const float4& acc2, uint32_t stamp) {
    const float4 v = make_float4(acc.x + acc2.x, acc.y + acc2.y, acc.z + acc2.z, acc.w);
    if (!LAG || cur_cell < 64) {
        const int nd = (cur_cell >> 4) * P2G_TILE_SX + ((cur_cell >> 2) & 3) * 6 + (cur_cell & 3) + noff;
        float4 t = tile[nd];
        t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
        tile[nd] = t;
    } else {
        const int gx = (cur_cell & 1023) + ni, gy = ((cur_cell >> 10) & 1023) + nj, gz = ((cur_cell >> 20) & 1023) + nk;
        const int n = a.n_grid;
        red_add_v4(a.grid_in + (((size_t)gx * n + gy) * n + gz), v);
        const uint32_t b = (uint32_t)(((gx >> 2) * a.nb + (gy >> 2)) * a.nb + (gz >> 2));
        if (atomicExch(&a.stamp_arr[b], stamp) != stamp) a.blist[atomicAdd(a.bcount, 1u)] = b;
    }
}

// The velocity record is read as 8 + 4 bytes: its fourth component is padding, and ptxas hands the register of a dead
// component of a prefetched 16-byte load to a temporary of phase B, whose first write then has to wait for the whole
// prefetch to land (a write-after-write hazard on the loads' scoreboard: 8 % of the kernel's stall samples sat there).
__device__ __forceinline__ float4 p2g_ldg_vel(const float4* p) {
    const float2 xy = __ldg(reinterpret_cast<const float2*>(p));
    const float z = __ldg(reinterpret_cast<const float*>(p) + 2);
    return make_float4(xy.x, xy.y, z, 0.f);
}

template <bool LAG>
__global__ void __launch_bounds__(P2G_WARPS * 32, P2G_MIN_CTAS) k_p2g(const P2GArgs a) {
    pdl_wait();   // programmatic dependent launch: nothing of the previous grid is read or overwritten before this
    extern __shared__ float4 smem_raw[];
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    P2GWarpSmem& S = reinterpret_cast<P2GWarpSmem*>(smem_raw)[warp];
    const bool node_lane = lane < 27;
    const int node = P2G_LANE_NODE[lane];
    const int ni = node / 9, nj = (node / 3) % 3, nk = node % 3;
    const int nij = ni * 3 + nj;
    const int noff = ni * P2G_TILE_SX + nj * 6 + nk;
    for (int i = lane; i < P2G_TILE_F4; i += 32) S.tile[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();
    const int n = a.n_grid;

    // Batches stream across dynamically fetched work items through a three-deep pipeline (see g2p.cu):
    // C = permutation/key loads in flight, B = particle records in flight (cp.async into S.stage, issued as
    // soon as phase A has pulled batch A's records into registers), A = being scattered.
    BatchStream bs;
    bs.init();
    const uint32_t n_items = a.cnt->n_items;
    const uint32_t stamp = LAG ? a.cnt->list_stamp : 0u;
    BatchDesc dA = bs.next(lane, a.items, &a.cnt->p2g_cursor, n_items);
    if (dA.count == 0) return;
    uint32_t pA = 0, kA = 0, pB = 0, kB = 0, pC = 0, kC = 0;
    if (lane < dA.count) { pA = LAG ? dA.base + lane : a.perm[dA.base + lane]; kA = a.skeys[dA.base + lane]; }
    BatchDesc dB = bs.next(lane, a.items, &a.cnt->p2g_cursor, n_items);
    if (lane < dB.count) { pB = LAG ? dB.base + lane : a.perm[dB.base + lane]; kB = a.skeys[dB.base + lane]; }
    // the next batch's particle records travel in registers (six 16-byte loads per lane, issued before phase B of the
    // current batch, which needs few registers and runs for hundreds of cycles): no shared-memory staging, which cost
    // 1.5 of the kernel's 12.4 shared-memory wavefronts per particle and 3 KB of shared memory per warp
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 Xn = zero4, Vn = zero4, C0n = zero4, C1n = zero4, CSn = zero4, S1n = zero4;
    if (lane < dA.count) {
        Xn = __ldg(a.xm + pA); Vn = p2g_ldg_vel(a.vv + pA); C0n = __ldg(a.c0 + pA);
        C1n = __ldg(a.c1 + pA); CSn = __ldg(a.cs + pA); S1n = __ldg(a.s1 + pA);
    }
    int cur_cell = -1;
    float4 acc = zero4, acc2 = zero4;
    int tx = 0, ty = 0, tz = 0;

    while (dA.count > 0) {
        if (dA.first) { tz = dA.tile % a.nt; ty = (dA.tile / a.nt) % a.nt; tx = dA.tile / (a.nt * a.nt); }
        const int nbatch = dA.count;
            int my_cell = 64;  // lanes past the end of the batch never match a real cell
            const float4 X = Xn, V = Vn, C0 = C0n, C1 = C1n, CS = CSn, S1 = S1n;
            const BatchDesc dC = bs.next(lane, a.items, &a.cnt->p2g_cursor, n_items);
            if (lane < dC.count) { pC = LAG ? dC.base + lane : a.perm[dC.base + lane]; kC = a.skeys[dC.base + lane]; }
            // ---------------- phase A: lane = particle ----------------
            if (lane < nbatch) {
                const uint32_t key = kA;
                my_cell = key & 63;
                int bx = 4 * tx + (my_cell >> 4), by = 4 * ty + ((my_cell >> 2) & 3), bz = 4 * tz + (my_cell & 3);   // base cell
                if (LAG && (key >> 6) != dA.tile) {   // the particle's cell left this item's tile since the list was built
                    const uint32_t tt = key >> 6;
                    bx = 4 * (int)(tt / (uint32_t)(a.nt * a.nt)) + (my_cell >> 4);
                    by = 4 * (int)((tt / (uint32_t)a.nt) % (uint32_t)a.nt) + ((my_cell >> 2) & 3);
                    bz = 4 * (int)(tt % (uint32_t)a.nt) + (my_cell & 3);
                    my_cell = P2G_CROSS | bx | (by << 10) | (bz << 20);
                }
                const float m = X.w, vol = S1.w;
                const float inv_m = m > 0.f ? __frcp_rn(m) : 0.f;
                float vx = V.x, vy = V.y, vz = V.z;
                for (int q = 0; q < a.n_imp; ++q) {
                    const mpmk_impulse& im = a.imp[q];
                    if (fabsf(X.x - im.center[0]) < im.size[0] && fabsf(X.y - im.center[1]) < im.size[1] &&
                        fabsf(X.z - im.center[2]) < im.size[2]) {
                        vx = vx + __fdiv_rn(im.force[0], m) * im.substep_dt;
                        vy = vy + __fdiv_rn(im.force[1], m) * im.substep_dt;
                        vz = vz + __fdiv_rn(im.force[2], m) * im.substep_dt;
                    }
                }
                const float fx = __fmul_rn(X.x, a.inv_dx) - (float)bx;
                const float fy = __fmul_rn(X.y, a.inv_dx) - (float)by;
                const float fz = __fmul_rn(X.z, a.inv_dx) - (float)bz;
                float wx[3], wy[3], wz[3], dwx[3], dwy[3], dwz[3];
                {
                    float wa = 1.5f - fx, wb = fx - 1.0f, wc = fx - 0.5f;
                    wx[0] = wa * wa * 0.5f; wx[1] = 0.75f - wb * wb; wx[2] = wc * wc * 0.5f;
                    dwx[0] = fx - 1.5f; dwx[1] = -2.0f * (fx - 1.0f); dwx[2] = fx - 0.5f;
                    wa = 1.5f - fy; wb = fy - 1.0f; wc = fy - 0.5f;
                    wy[0] = wa * wa * 0.5f; wy[1] = 0.75f - wb * wb; wy[2] = wc * wc * 0.5f;
                    dwy[0] = fy - 1.5f; dwy[1] = -2.0f * (fy - 1.0f); dwy[2] = fy - 0.5f;
                    wa = 1.5f - fz; wb = fz - 1.0f; wc = fz - 0.5f;
                    wz[0] = wa * wa * 0.5f; wz[1] = 0.75f - wb * wb; wz[2] = wc * wc * 0.5f;
                    dwz[0] = fz - 1.5f; dwz[1] = -2.0f * (fz - 1.0f); dwz[2] = fz - 0.5f;
                }
                // A = m C dx (columns A0,A1,A2), S = -dt vol inv_dx stress (symmetric), q = m v - A f
                const float mdx = m * a.dx;
                const float A00 = mdx * C0.x, A01 = mdx * C0.y, A02 = mdx * C0.z;
                const float A10 = mdx * C0.w, A11 = mdx * C1.x, A12 = mdx * C1.y;
                const float A20 = mdx * C1.z, A21 = mdx * C1.w, A22 = mdx * CS.x;
                const float sc = -a.dt * vol * a.inv_dx;
                const float S00 = sc * CS.y, S01 = sc * CS.z, S02 = sc * CS.w, S11 = sc * S1.x, S12 = sc * S1.y, S22 = sc * S1.z;
                const float q0 = m * vx - (A00 * fx + A01 * fy + A02 * fz);
                const float q1 = m * vy - (A10 * fx + A11 * fy + A12 * fz);
                const float q2 = m * vz - (A20 * fx + A21 * fy + A22 * fz);
                float T0[3], T1[3], T2[3];
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                    const float wj = wy[j] * (float)j;
                    T0[j] = wj * A01 + dwy[j] * S01;
                    T1[j] = wj * A11 + dwy[j] * S11;
                    T2[j] = wj * A21 + dwy[j] * S12;
                }
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    const float fi = (float)i;
                    const float X0 = wx[i] * (q0 + fi * A00) + dwx[i] * S00;
                    const float X1 = wx[i] * (q1 + fi * A10) + dwx[i] * S01;
                    const float X2 = wx[i] * (q2 + fi * A20) + dwx[i] * S02;
#pragma unroll
                    for (int j = 0; j < 3; ++j) {
                        float4 yw;
                        yw.x = wy[j] * X0 + wx[i] * T0[j];
                        yw.y = wy[j] * X1 + wx[i] * T1[j];
                        yw.z = wy[j] * X2 + wx[i] * T2[j];
                        yw.w = m * (wx[i] * wy[j]);
                        S.rec[(i * 3 + j) * P2G_REC_STRIDE + lane] = yw;
                    }
                }
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const float wk = wz[k] * (float)k;
                    float4 rw;
                    rw.x = (wk * A02 + dwz[k] * S02) * inv_m;
                    rw.y = (wk * A12 + dwz[k] * S12) * inv_m;
                    rw.z = (wk * A22 + dwz[k] * S22) * inv_m;
                    rw.w = wz[k];
                    S.rec[(9 + k) * P2G_REC_STRIDE + lane] = rw;
                }
            }
            // batch B's records: in flight during phase B
            if (lane < dB.count) {
                Xn = __ldg(a.xm + pB); Vn = p2g_ldg_vel(a.vv + pB); C0n = __ldg(a.c0 + pB);
                C1n = __ldg(a.c1 + pB); CSn = __ldg(a.cs + pB); S1n = __ldg(a.s1 + pB);
            }
            __syncwarp();
            // ---------------- phase B: lane = stencil node ----------------
            // run heads: lanes whose cell differs from the lane before them (lane 0: from the run carried over from
            // the previous batch).  Sorted input makes a cell's particles contiguous; in a lagged batch a particle that
            // changed cell sits inside its old cell's run and splits it.
            const int prev_cell = __shfl_up_sync(0xffffffffu, my_cell, 1);
            const bool head = (lane < nbatch) && (my_cell != (lane == 0 ? cur_cell : prev_cell));
            const unsigned heads = __ballot_sync(0xffffffffu, head);
            const float4* recY = S.rec + nij * P2G_REC_STRIDE;
            const float4* recR = S.rec + (9 + nk) * P2G_REC_STRIDE;
            int l = 0;
            while (l < nbatch) {
                if ((heads >> l) & 1u) {
                    if (cur_cell >= 0 && node_lane) p2g_flush_run<LAG>(a, S.tile, cur_cell, noff, ni, nj, nk, acc, acc2, stamp);
                    acc = zero4; acc2 = zero4;
                    cur_cell = __shfl_sync(0xffffffffu, my_cell, l);
                }
                const unsigned rest = heads & ~((2u << l) - 1u);  // heads strictly after l
                const int end = rest ? (__ffs(rest) - 1) : nbatch;
#pragma unroll 4
                for (; l < end; ++l) {
                    // seven FFMAs on two independent accumulator sets (a sum written as rw.w*yw.x + yw.w*rw.x + acc
                    // compiles to FMUL + FFMA + FADD)
                    const float4 yw = recY[l];
                    const float4 rw = recR[l];
                    acc.x = fmaf(rw.w, yw.x, acc.x); acc2.x = fmaf(yw.w, rw.x, acc2.x);
                    acc.y = fmaf(rw.w, yw.y, acc.y); acc2.y = fmaf(yw.w, rw.y, acc2.y);
                    acc.z = fmaf(rw.w, yw.z, acc.z); acc2.z = fmaf(yw.w, rw.z, acc2.z);
                    acc.w = fmaf(yw.w, rw.w, acc.w);
                }
            }
            __syncwarp();
            // end of the item: last cell run -> tile, tile -> global grid (16-byte vector reductions)
            if (dB.count == 0 || dB.first) {
                if (cur_cell >= 0 && node_lane) p2g_flush_run<LAG>(a, S.tile, cur_cell, noff, ni, nj, nk, acc, acc2, stamp);
                __syncwarp();
#ifdef P2G_FLUSH_BULK
                // the tile as 36 bulk reductions (TMA engine, cp.reduce.async.bulk .add.f32): one 96-byte row of 6 nodes each.
                // Tiles that stick out of the grid keep the per-node path.
                if ((4 * tx + 6 <= n) && (4 * ty + 6 <= n) && (4 * tz + 6 <= n)) {
                    fence_proxy_async_smem();   // the tile was written through the generic proxy
                    __syncwarp();
                    for (int r = lane; r < 36; r += 32) {
                        const int gx = 4 * tx + r / 6, gy = 4 * ty + r % 6;
                        bulk_red_add_f32(a.grid_in + (((size_t)gx * n + gy) * n + 4 * tz), &S.tile[(r / 6) * P2G_TILE_SX + (r % 6) * 6], 96);
                    }
                    bulk_commit();
                    bulk_wait_read0();          // the engine has read the rows: the tile may be reset
                    __syncwarp();
                    for (int i = lane; i < P2G_TILE_F4; i += 32) S.tile[i] = zero4;
                } else
#endif
                for (int i = lane; i < MPMK_TILE_NODES; i += 32) {
                    const int ti = i + i / 36;  // x*37 + y*6 + z
                    const float4 t = S.tile[ti];
                    S.tile[ti] = zero4;
                    if (t.x != 0.f || t.y != 0.f || t.z != 0.f || t.w != 0.f) {
                        const int gx = 4 * tx + i / 36, gy = 4 * ty + (i / 6) % 6, gz = 4 * tz + i % 6;
                        if (gx < n && gy < n && gz < n) red_add_v4(a.grid_in + (((size_t)gx * n + gy) * n + gz), t);
                    }
                }
                cur_cell = -1;
                acc = zero4; acc2 = zero4;
                __syncwarp();
            }
            dA = dB; pA = pB; kA = kB;
            dB = dC; pB = pC; kB = kC;
    }
}

static int launch_p2g(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, cudaStream_t s, bool lag) {
    if (ctx->n == 0) return 0;
    const mpmk_particle_set& ps = ctx->b.set[ctx->cur];
    P2GArgs a;
    a.perm = ctx->perm; a.skeys = lag ? ctx->keys_cur : ctx->sorted_keys;
    a.stamp_arr = ctx->block_stamp[ctx->cur_list]; a.blist = ctx->block_list[ctx->cur_list]; a.bcount = ctx->block_count[ctx->cur_list];
    a.nb = ctx->nb;
    a.xm = (const float4*)ps.xm; a.vv = (const float4*)ps.vv; a.c0 = (const float4*)ps.c0;
    a.c1 = (const float4*)ps.c1; a.cs = (const float4*)ps.cs; a.s1 = (const float4*)ps.s1;
    a.items = ctx->items; a.cnt = ctx->counters; a.grid_in = (float4*)ctx->b.grid_in;
    a.n_grid = ctx->n_grid; a.nt = ctx->nt; a.dx = ctx->dx; a.inv_dx = ctx->inv_dx; a.dt = dt;
    a.n_imp = 0;
    if (ops) {
        if (ops->n_impulses < 0 || ops->n_impulses > MPMK_MAX_IMPULSES) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "n_impulses out of range");
        a.n_imp = ops->n_impulses;
        memcpy(a.imp, ops->impulses, sizeof(mpmk_impulse) * ops->n_impulses);
    }
    const int smem = (int)sizeof(P2GWarpSmem) * P2G_WARPS;
    // persistent grid: as many CTAs as fit per SM, warps fetch items dynamically
    int64_t want = (ctx->n / 32 + P2G_WARPS - 1) / P2G_WARPS + 1;
    int64_t cap = (int64_t)ctx->sm_count * (lag ? ctx->p2g_lag_ctas_per_sm : ctx->p2g_ctas_per_sm);
    int grid = (int)(want < cap ? want : cap);
    if (lag) mpmk_launch(k_p2g<true>, dim3(grid), dim3(P2G_WARPS * 32), smem, s, a);
    else mpmk_launch(k_p2g<false>, dim3(grid), dim3(P2G_WARPS * 32), smem, s, a);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}

int mpmk_launch_p2g(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, cudaStream_t s) {
    if (!ctx->prepared) MPMK_FAIL(ctx, MPMK_ERR_STATE, "mpmk_p2g: prepare_grid first");
    return launch_p2g(ctx, dt, ops, s, false);
}

int mpmk_launch_p2g_lagged(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, cudaStream_t s) { return launch_p2g(ctx, dt, ops, s, true); }

int mpmk_p2g_init(mpmk_ctx* ctx) {
    const int smem = (int)sizeof(P2GWarpSmem) * P2G_WARPS;
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_p2g<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_p2g<false>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_p2g<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_p2g<true>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    int per_sm = 0;
    MPMK_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_p2g<false>, P2G_WARPS * 32, smem));
    ctx->p2g_ctas_per_sm = per_sm > 0 ? per_sm : 1;
    MPMK_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_p2g<true>, P2G_WARPS * 32, smem));
    ctx->p2g_lag_ctas_per_sm = per_sm > 0 ? per_sm : 1;
    // the side branch (sort + work lists of the next substep) needs room beside the persistent scatter grid
    // (4 CTAs of 128 threads x 96 registers and 39.5 KB leave 16 K registers and 66 KB per SM: one 256-thread CTA of
    // every side-branch kernel)
    static const int leave = getenv("MPMK_LAG_P2G_CTAS") ? atoi(getenv("MPMK_LAG_P2G_CTAS")) : 4;
    if (leave > 0 && leave < ctx->p2g_lag_ctas_per_sm) ctx->p2g_lag_ctas_per_sm = leave;
    static const int classic = getenv("MPMK_P2G_CTAS") ? atoi(getenv("MPMK_P2G_CTAS")) : 0;
    if (classic > 0 && classic < ctx->p2g_ctas_per_sm) ctx->p2g_ctas_per_sm = classic;
    return 0;
}
