// g2p2g.cu -- K4+K2 fused: the g2p of substep s (utils.py:219-282, with everything g2p.cu folds into it) and the
// p2g of substep s+1 (utils.py:90-134, impulses of boundary_conditions.py:41-45) in one pass over the particles.
//
// Between two substeps of a constant-dt batch the particle state only travels g2p -> (memory) -> p2g.  Here the
// scatter of substep s+1 runs on the registers the gather of substep s just produced, so v, C and the stress never
// reach memory (80 of the 164 bytes g2p writes per particle) and p2g's own pass over the particles (104 bytes read
// per particle) disappears.  The arithmetic per particle is that of k_g2p followed by that of k_p2g on the same fp32
// values; only the order in which a node's contributions are summed differs from the two-kernel path.
//
// The particles are still in the order of the sort of substep s, i.e. grouped by the cell they were in BEFORE the
// move.  The scatter tile (6^3 nodes of the warp's 4^3-cell tile, as in p2g.cu) therefore only takes the particles
// whose new cell is still inside the tile; the few that crossed a tile face ("leavers", ~1 % per substep at CFL 0.01)
// scatter their 27 nodes straight to global memory with red.global.add.v4.f32.  Cell runs are detected on the NEW
// cell (a particle that changed cell merely splits a run).  No assumption on how far a particle moved is needed.
//
// The substep driver (api.cu) pairs this kernel with a grid pass that zeroes grid_in behind itself and a clear that
// leaves grid_in alone: the scatter of s+1 lands before the sort of s+1 knows which blocks it touched.
#include "common.cuh"
#include "constitutive.cuh"

#ifndef F_WARPS
#define F_WARPS 4
#endif
#ifndef F_MIN_CTAS
#define F_MIN_CTAS 3
#endif
#define F_REC_STRIDE 33   // float4 per record slot: 32 lanes + 1 pad (see p2g.cu)
#define F_REC_SLOTS 12    // 9 x {Y_ij.xyz, m wx_i wy_j} + 3 x {R_k.xyz / m, wz_k}
#define F_TILE_SX 37      // x stride of the scatter tile in float4 (6*6 + 1 pad)
#define F_TILE_F4 (6 * F_TILE_SX)
#define F_NREC 5          // xm, f0, f1, f2, meta

// lane -> stencil node map of p2g.cu (conflict-free tile read-modify-write and record reads per quarter-warp)
__constant__ unsigned char F_LANE_NODE[32] = {2, 1, 11, 19, 9, 10, 21, 13, 24, 18, 4, 15, 5, 14, 6, 12, 3, 23, 7, 22, 25, 17, 20, 26, 8, 16, 0, 0, 0, 0, 0, 0};

struct __align__(16) FWarpSmem {
    float4 rec[F_REC_SLOTS * F_REC_STRIDE];   // p2g records of the batch being scattered
    float4 ptile[F_TILE_F4];                  // scatter tile {px, py, pz, m} of substep s+1
    float4 gtile[MPMK_TILE_NODES];            // gather tile {vx, vy, vz, -} of substep s
    float4 stage[2][F_NREC][32];              // particle records of batches A / B (cp.async)
    uint32_t perm[2][32];
    uint32_t key[3][32];
};

struct FArgs {
    const uint32_t* perm;
    const uint32_t* skeys;
    const float4 *xm, *f0, *f1, *f2, *cv0, *cv1;
    const int4* meta;
    float4 *oxm, *of0, *of1, *of2, *ocv0, *ocv1;
    int4* ometa;
    uint32_t* okeys;
    const uint4* items;
    const float4* grid_out;
    float4* grid_in;
    DevCounters* cnt;
    int n_grid, nt;
    float dx, inv_dx, dt;
    ConstParams cp;
    int n_imp;
    mpmk_impulse imp[MPMK_MAX_IMPULSES];
};

__device__ __forceinline__ void f_bspline_w(float f, float* w) {
    float wa = 1.5f - f, wb = f - 1.0f, wc = f - 0.5f;
    w[0] = wa * wa * 0.5f; w[1] = 0.75f - wb * wb; w[2] = wc * wc * 0.5f;
}

__global__ void __launch_bounds__(F_WARPS * 32, F_MIN_CTAS) k_g2p2g(const FArgs a) {
    extern __shared__ float4 f_smem_raw[];
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    FWarpSmem& S = reinterpret_cast<FWarpSmem*>(f_smem_raw)[warp];
    const int n = a.n_grid;
    const float dt = a.dt;
    const float k4 = 4.0f * a.inv_dx;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    // scatter side: this lane's stencil node
    const bool node_lane = lane < 27;
    const int node = F_LANE_NODE[lane];
    const int ni = node / 9, nj = (node / 3) % 3, nk = node % 3;
    const int nij = ni * 3 + nj;
    const int noff = ni * F_TILE_SX + nj * 6 + nk;
    for (int i = lane; i < F_TILE_F4; i += 32) S.ptile[i] = zero4;
    __syncwarp();

    BatchStream bs;
    bs.init();
    const uint32_t n_items = a.cnt->n_items;
    BatchDesc dA = bs.next(lane, a.items, &a.cnt->g2p_cursor, n_items);
    if (dA.count == 0) return;
    int stage = 0, kslot = 0, pslot = 0;
    uint32_t pA = 0;
    if (lane < dA.count) {
        pA = a.perm[dA.base + lane];
        S.key[0][lane] = a.skeys[dA.base + lane];
        cp_async16(&S.stage[0][0][lane], a.xm + pA);
        cp_async16(&S.stage[0][1][lane], a.f0 + pA);
        cp_async16(&S.stage[0][2][lane], a.f1 + pA);
        cp_async16(&S.stage[0][3][lane], a.f2 + pA);
        cp_async16(&S.stage[0][4][lane], a.meta + pA);
    }
    BatchDesc dB = bs.next(lane, a.items, &a.cnt->g2p_cursor, n_items);
    if (lane < dB.count) {
        cp_async4(&S.perm[0][lane], a.perm + dB.base + lane);
        cp_async4(&S.key[1][lane], a.skeys + dB.base + lane);
    }
    cp_async_commit();
    int cur_cell = -1;     // cell (0..63) of the run being accumulated, -1 = none
    float4 acc = zero4, acc2 = zero4;

    while (dA.count > 0) {
        cp_async_wait<0>();
        __syncwarp();
        const int tz = dA.tile % a.nt, ty = (dA.tile / a.nt) % a.nt, tx = dA.tile / (a.nt * a.nt);
        if (dA.first) {
            for (int i = lane; i < MPMK_TILE_NODES; i += 32) {
                const int gx = 4 * tx + i / 36, gy = 4 * ty + (i / 6) % 6, gz = 4 * tz + i % 6;
                if (gx < n && gy < n && gz < n) cp_async16(&S.gtile[i], a.grid_out + (((size_t)gx * n + gy) * n + gz));
                else S.gtile[i] = zero4;
            }
        }
        cp_async_commit();
        uint32_t pB = 0;
        if (lane < dB.count) {
            pB = S.perm[pslot][lane];
            cp_async16(&S.stage[stage ^ 1][0][lane], a.xm + pB);
            cp_async16(&S.stage[stage ^ 1][1][lane], a.f0 + pB);
            cp_async16(&S.stage[stage ^ 1][2][lane], a.f1 + pB);
            cp_async16(&S.stage[stage ^ 1][3][lane], a.f2 + pB);
            cp_async16(&S.stage[stage ^ 1][4][lane], a.meta + pB);
        }
        const BatchDesc dC = bs.next(lane, a.items, &a.cnt->g2p_cursor, n_items);
        const int kslotC = (kslot + 2) % 3;
        if (lane < dC.count) {
            cp_async4(&S.perm[pslot ^ 1][lane], a.perm + dC.base + lane);
            cp_async4(&S.key[kslotC][lane], a.skeys + dC.base + lane);
        }
        cp_async_commit();
        cp_async_wait<1>();
        __syncwarp();

        const int nbatch = dA.count;
        int my_cell = 64;       // new cell inside the tile; 64 = not a scatter-tile particle (past the end, or a leaver)
        bool leaver = false;
        int kx = 0, ky = 0, kz = 0;
        if (lane < nbatch) {
            // ================= g2p of substep s (k_g2p, g2p.cu) =================
            const uint32_t r = dA.base + lane;
            const uint32_t key = S.key[kslot][lane];
            const float4 X = S.stage[stage][0][lane];
            const int lx = (key >> 4) & 3, ly = (key >> 2) & 3, lz = key & 3;
            float v[3] = {0, 0, 0}, Mx[3] = {0, 0, 0}, My[3] = {0, 0, 0}, Mz[3] = {0, 0, 0};
            float Gx[3] = {0, 0, 0}, Gy[3] = {0, 0, 0}, Gz[3] = {0, 0, 0};
            {
                const float fx = __fmul_rn(X.x, a.inv_dx) - (float)(4 * tx + lx);
                const float fy = __fmul_rn(X.y, a.inv_dx) - (float)(4 * ty + ly);
                const float fz = __fmul_rn(X.z, a.inv_dx) - (float)(4 * tz + lz);
                float wx[3], wy[3], wz[3], dwz[3];
                f_bspline_w(fx, wx); f_bspline_w(fy, wy); f_bspline_w(fz, wz);
                dwz[0] = fz - 1.5f; dwz[1] = -2.0f * (fz - 1.0f); dwz[2] = fz - 0.5f;
                const float wzd[3] = {wz[0] * (0.f - fz), wz[1] * (1.f - fz), wz[2] * (2.f - fz)};
                const float4* g = S.gtile + (lx * 36 + ly * 6 + lz);
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    float A[3] = {0, 0, 0}, Ay[3] = {0, 0, 0}, Gyi[3] = {0, 0, 0}, Az[3] = {0, 0, 0}, Gzi[3] = {0, 0, 0};
#pragma unroll
                    for (int j = 0; j < 3; ++j) {
                        const float4* row = g + (i * 36 + j * 6);
                        const float4 g0 = row[0], g1 = row[1], g2 = row[2];
                        const float ax = wz[0] * g0.x + wz[1] * g1.x + wz[2] * g2.x;
                        const float ay = wz[0] * g0.y + wz[1] * g1.y + wz[2] * g2.y;
                        const float az = wz[0] * g0.z + wz[1] * g1.z + wz[2] * g2.z;
                        const float zx = wzd[0] * g0.x + wzd[1] * g1.x + wzd[2] * g2.x;
                        const float zy = wzd[0] * g0.y + wzd[1] * g1.y + wzd[2] * g2.y;
                        const float zz = wzd[0] * g0.z + wzd[1] * g1.z + wzd[2] * g2.z;
                        const float dx_ = dwz[0] * g0.x + dwz[1] * g1.x + dwz[2] * g2.x;
                        const float dy_ = dwz[0] * g0.y + dwz[1] * g1.y + dwz[2] * g2.y;
                        const float dz_ = dwz[0] * g0.z + dwz[1] * g1.z + dwz[2] * g2.z;
                        const float wj = wy[j], wjj = wy[j] * ((float)j - fy);
                        const float dj = (j == 0) ? fy - 1.5f : (j == 1) ? -2.0f * (fy - 1.0f) : fy - 0.5f;
                        A[0] += wj * ax; A[1] += wj * ay; A[2] += wj * az;
                        Ay[0] += wjj * ax; Ay[1] += wjj * ay; Ay[2] += wjj * az;
                        Gyi[0] += dj * ax; Gyi[1] += dj * ay; Gyi[2] += dj * az;
                        Az[0] += wj * zx; Az[1] += wj * zy; Az[2] += wj * zz;
                        Gzi[0] += wj * dx_; Gzi[1] += wj * dy_; Gzi[2] += wj * dz_;
                    }
                    const float wi = wx[i], wii = wx[i] * ((float)i - fx);
                    const float di = (i == 0) ? fx - 1.5f : (i == 1) ? -2.0f * (fx - 1.0f) : fx - 0.5f;
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        v[c] += wi * A[c];
                        Mx[c] += wii * A[c];
                        Gx[c] += di * A[c];
                        My[c] += wi * Ay[c];
                        Gy[c] += wi * Gyi[c];
                        Mz[c] += wi * Az[c];
                        Gz[c] += wi * Gzi[c];
                    }
                }
            }
            const float nx = X.x + dt * v[0], ny = X.y + dt * v[1], nz = X.z + dt * v[2];
            const float m = X.w;
            a.oxm[r] = make_float4(nx, ny, nz, m);
            bool cl = false;
            kx = base_cell(nx, a.inv_dx, n, cl); ky = base_cell(ny, a.inv_dx, n, cl); kz = base_cell(nz, a.inv_dx, n, cl);
            a.okeys[r] = make_key(kx, ky, kz, a.nt);
            if (cl) atomicAdd(&a.cnt->n_clamped, 1u);
            // C = sum w v (x) (off - f) * 4 inv_dx, rows = velocity component (the values k_g2p stores in c0/c1/cs.x)
            const float C00 = Mx[0] * k4, C01 = My[0] * k4, C02 = Mz[0] * k4;
            const float C10 = Mx[1] * k4, C11 = My[1] * k4, C12 = Mz[1] * k4;
            const float C20 = Mx[2] * k4, C21 = My[2] * k4, C22 = Mz[2] * k4;
            float gv[9];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                gv[3 * c + 0] = Gx[c] * a.inv_dx;
                gv[3 * c + 1] = Gy[c] * a.inv_dx;
                gv[3 * c + 2] = Gz[c] * a.inv_dx;
            }
            const float4 F0 = S.stage[stage][1][lane], F1 = S.stage[stage][2][lane], F2 = S.stage[stage][3][lane];
            const float Fo[9] = {F0.x, F0.y, F0.z, F0.w, F1.x, F1.y, F1.z, F1.w, F2.x};
            float Ft[9];
#pragma unroll
            for (int rr = 0; rr < 3; ++rr)
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    float s = (rr == 0 ? 1.0f + dt * gv[0] : dt * gv[3 * rr + 0]) * Fo[c];
                    s += (rr == 1 ? 1.0f + dt * gv[4] : dt * gv[3 * rr + 1]) * Fo[3 + c];
                    s += (rr == 2 ? 1.0f + dt * gv[8] : dt * gv[3 * rr + 2]) * Fo[6 + c];
                    Ft[3 * rr + c] = s;
                }
            if (a.cp.flags & MPMK_TRACK_COV) {
                const float4 q0 = ldg4(a.cv0 + pA), q1 = ldg4(a.cv1 + pA);
                const float cn[9] = {q0.x, q0.y, q0.z, q0.y, q0.w, q1.x, q0.z, q1.x, q1.y};
                float o[9];
#pragma unroll
                for (int rr = 0; rr < 3; ++rr)
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        float s1 = gv[3 * rr] * cn[c] + gv[3 * rr + 1] * cn[3 + c] + gv[3 * rr + 2] * cn[6 + c];
                        float s2 = cn[3 * rr] * gv[3 * c] + cn[3 * rr + 1] * gv[3 * c + 1] + cn[3 * rr + 2] * gv[3 * c + 2];
                        o[3 * rr + c] = cn[3 * rr + c] + dt * (s1 + s2);
                    }
                a.ocv0[r] = make_float4(o[0], o[1], o[2], o[4]);
                a.ocv1[r] = make_float4(o[5], o[8], 0.f, 0.f);
            }
            const int4 meta = *reinterpret_cast<const int4*>(&S.stage[stage][4][lane]);
            a.ometa[r] = meta;
            float mu = F2.y, lam = F2.z, ys = F2.w;
            float Fn[9], s6[6];
            constitutive_update(Ft, meta.y, mu, lam, ys, dt, a.cp, Fn, s6);
            const float vol = __int_as_float(meta.z);
            a.of0[r] = make_float4(Fn[0], Fn[1], Fn[2], Fn[3]);
            a.of1[r] = make_float4(Fn[4], Fn[5], Fn[6], Fn[7]);
            a.of2[r] = make_float4(Fn[8], mu, lam, ys);

            // ================= p2g of substep s+1, phase A (k_p2g, p2g.cu) on the registers above =================
            leaver = ((kx >> 2) != tx) || ((ky >> 2) != ty) || ((kz >> 2) != tz);
            my_cell = leaver ? 64 : (((kx & 3) << 4) | ((ky & 3) << 2) | (kz & 3));
            const float inv_m = m > 0.f ? __frcp_rn(m) : 0.f;
            float vx = v[0], vy = v[1], vz = v[2];
            for (int q = 0; q < a.n_imp; ++q) {   // boundary_conditions.py:41-45 on the new position
                const mpmk_impulse& im = a.imp[q];
                if (fabsf(nx - im.center[0]) < im.size[0] && fabsf(ny - im.center[1]) < im.size[1] &&
                    fabsf(nz - im.center[2]) < im.size[2]) {
                    vx = vx + __fdiv_rn(im.force[0], m) * im.substep_dt;
                    vy = vy + __fdiv_rn(im.force[1], m) * im.substep_dt;
                    vz = vz + __fdiv_rn(im.force[2], m) * im.substep_dt;
                }
            }
            const float fx = __fmul_rn(nx, a.inv_dx) - (float)kx;
            const float fy = __fmul_rn(ny, a.inv_dx) - (float)ky;
            const float fz = __fmul_rn(nz, a.inv_dx) - (float)kz;
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
            const float mdx = m * a.dx;
            const float A00 = mdx * C00, A01 = mdx * C01, A02 = mdx * C02;
            const float A10 = mdx * C10, A11 = mdx * C11, A12 = mdx * C12;
            const float A20 = mdx * C20, A21 = mdx * C21, A22 = mdx * C22;
            const float sc = -a.dt * vol * a.inv_dx;
            const float S00 = sc * s6[0], S01 = sc * s6[1], S02 = sc * s6[2], S11 = sc * s6[3], S12 = sc * s6[4], S22 = sc * s6[5];
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
                    S.rec[(i * 3 + j) * F_REC_STRIDE + lane] = yw;
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
                S.rec[(9 + k) * F_REC_STRIDE + lane] = rw;
            }
        }
        // leavers: the new cell lies in another tile -- their 27 node values go straight to the global grid, and their
        // records are blanked so that the node lanes below add nothing for them
        if (leaver) {
            float4 rw[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) { rw[k] = S.rec[(9 + k) * F_REC_STRIDE + lane]; S.rec[(9 + k) * F_REC_STRIDE + lane] = zero4; }
#pragma unroll
            for (int ij = 0; ij < 9; ++ij) {
                const float4 yw = S.rec[ij * F_REC_STRIDE + lane];
                S.rec[ij * F_REC_STRIDE + lane] = zero4;
                float4* g = a.grid_in + (((size_t)(kx + ij / 3) * n + (ky + ij % 3)) * n + kz);
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    float4 t;
                    t.x = rw[k].w * yw.x + yw.w * rw[k].x;
                    t.y = rw[k].w * yw.y + yw.w * rw[k].y;
                    t.z = rw[k].w * yw.z + yw.w * rw[k].z;
                    t.w = yw.w * rw[k].w;
                    red_add_v4(g + k, t);
                }
            }
        }
        __syncwarp();
        // ---------------- phase B: lane = stencil node (p2g.cu) ----------------
        // a run = consecutive particles with the same NEW cell; its head is a lane whose cell differs from its
        // predecessor's (lane 0: from the run carried over from the previous batch)
        int prev_cell = __shfl_up_sync(0xffffffffu, my_cell, 1);
        if (lane == 0) prev_cell = cur_cell;
        const bool head = (lane < nbatch) && (my_cell != prev_cell);
        const unsigned heads = __ballot_sync(0xffffffffu, head);
        const float4* recY = S.rec + nij * F_REC_STRIDE;
        const float4* recR = S.rec + (9 + nk) * F_REC_STRIDE;
        int l = 0;
        while (l < nbatch) {
            if ((heads >> l) & 1u) {
                if (cur_cell >= 0 && cur_cell < 64 && node_lane) {
                    const int nd = (cur_cell >> 4) * F_TILE_SX + ((cur_cell >> 2) & 3) * 6 + (cur_cell & 3) + noff;
                    float4 t = S.ptile[nd];
                    t.x += acc.x + acc2.x; t.y += acc.y + acc2.y; t.z += acc.z + acc2.z; t.w += acc.w;
                    S.ptile[nd] = t;
                }
                acc = zero4; acc2 = zero4;
                cur_cell = __shfl_sync(0xffffffffu, my_cell, l);
            }
            const unsigned rest = heads & ~((2u << l) - 1u);
            const int end = rest ? (__ffs(rest) - 1) : nbatch;
#pragma unroll 4
            for (; l < end; ++l) {
                const float4 yw = recY[l];
                const float4 rw = recR[l];
                acc.x = fmaf(rw.w, yw.x, acc.x); acc2.x = fmaf(yw.w, rw.x, acc2.x);
                acc.y = fmaf(rw.w, yw.y, acc.y); acc2.y = fmaf(yw.w, rw.y, acc2.y);
                acc.z = fmaf(rw.w, yw.z, acc.z); acc2.z = fmaf(yw.w, rw.z, acc2.z);
                acc.w = fmaf(yw.w, rw.w, acc.w);
            }
        }
        __syncwarp();
        // end of the item: last run -> tile, tile -> global grid
        if (dB.count == 0 || dB.first) {
            if (cur_cell >= 0 && cur_cell < 64 && node_lane) {
                const int nd = (cur_cell >> 4) * F_TILE_SX + ((cur_cell >> 2) & 3) * 6 + (cur_cell & 3) + noff;
                float4 t = S.ptile[nd];
                t.x += acc.x + acc2.x; t.y += acc.y + acc2.y; t.z += acc.z + acc2.z; t.w += acc.w;
                S.ptile[nd] = t;
            }
            __syncwarp();
            for (int i = lane; i < MPMK_TILE_NODES; i += 32) {
                const int ti = i + i / 36;
                const float4 t = S.ptile[ti];
                S.ptile[ti] = zero4;
                if (t.x != 0.f || t.y != 0.f || t.z != 0.f || t.w != 0.f) {
                    const int gx = 4 * tx + i / 36, gy = 4 * ty + (i / 6) % 6, gz = 4 * tz + i % 6;
                    if (gx < n && gy < n && gz < n) red_add_v4(a.grid_in + (((size_t)gx * n + gy) * n + gz), t);
                }
            }
            cur_cell = -1;
            acc = zero4; acc2 = zero4;
        }
        __syncwarp();  // all lanes are done with the tiles, the records and this stage before any is overwritten
        dA = dB; pA = pB;
        dB = dC;
        stage ^= 1;
        pslot ^= 1;
        kslot = (kslot + 1) % 3;
    }
    cp_async_wait<0>();
}

static ConstParams f_make_cp(const mpmk_ctx* ctx) {
    ConstParams cp;
    cp.alpha = ctx->prm.alpha; cp.plastic_viscosity = ctx->prm.plastic_viscosity; cp.xi = ctx->prm.xi;
    cp.softening = ctx->prm.softening; cp.hardening = ctx->prm.hardening; cp.flags = ctx->prm.flags;
    return cp;
}

int mpmk_g2p2g_init(mpmk_ctx* ctx) {
    const int smem = (int)sizeof(FWarpSmem) * F_WARPS;
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_g2p2g, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_g2p2g, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    int per_sm = 0;
    MPMK_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_g2p2g, F_WARPS * 32, smem));
    ctx->g2p2g_ctas_per_sm = per_sm > 0 ? per_sm : 1;
    return 0;
}

// g2p of the current substep + p2g of the next one (same dt, same impulses).  Leaves the context as mpmk_launch_g2p
// does, with grid_in holding the next substep's scatter.
int mpmk_launch_g2p2g(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, cudaStream_t s) {
    if (!ctx->prepared) MPMK_FAIL(ctx, MPMK_ERR_STATE, "mpmk_g2p2g: sort + prepare_grid first");
    const int src = ctx->cur, dst = 1 - ctx->cur;
    if (ctx->n > 0) {
        const mpmk_particle_set& A = ctx->b.set[src];
        const mpmk_particle_set& B = ctx->b.set[dst];
        FArgs a;
        a.perm = ctx->perm; a.skeys = ctx->sorted_keys;
        a.xm = (const float4*)A.xm; a.f0 = (const float4*)A.f0; a.f1 = (const float4*)A.f1; a.f2 = (const float4*)A.f2;
        a.cv0 = (const float4*)A.cv0; a.cv1 = (const float4*)A.cv1; a.meta = (const int4*)A.meta;
        a.oxm = (float4*)B.xm; a.of0 = (float4*)B.f0; a.of1 = (float4*)B.f1; a.of2 = (float4*)B.f2;
        a.ocv0 = (float4*)B.cv0; a.ocv1 = (float4*)B.cv1; a.ometa = (int4*)B.meta;
        a.okeys = (uint32_t*)ctx->b.keys; a.items = ctx->items;
        a.grid_out = (const float4*)ctx->b.grid_out; a.grid_in = (float4*)ctx->b.grid_in; a.cnt = ctx->counters;
        a.n_grid = ctx->n_grid; a.nt = ctx->nt; a.dx = ctx->dx; a.inv_dx = ctx->inv_dx; a.dt = dt;
        a.cp = f_make_cp(ctx);
        a.cp.flags &= ~MPMK_WRITE_FTRIAL;   // F_trial is only stored by the stand-alone g2p that ends a batch
        if ((a.cp.flags & MPMK_TRACK_COV) && (!A.cv0 || !A.cv1 || !B.cv0 || !B.cv1))
            MPMK_FAIL(ctx, MPMK_ERR_STATE, "MPMK_TRACK_COV needs cv0/cv1 bound");
        a.n_imp = 0;
        if (ops) {
            if (ops->n_impulses < 0 || ops->n_impulses > MPMK_MAX_IMPULSES) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "n_impulses out of range");
            a.n_imp = ops->n_impulses;
            memcpy(a.imp, ops->impulses, sizeof(mpmk_impulse) * ops->n_impulses);
        }
        int64_t want = (ctx->n / 32 + F_WARPS - 1) / F_WARPS + 1;
        int64_t cap = (int64_t)ctx->sm_count * ctx->g2p2g_ctas_per_sm;
        k_g2p2g<<<(int)(want < cap ? want : cap), F_WARPS * 32, sizeof(FWarpSmem) * F_WARPS, s>>>(a);
        MPMK_LAUNCH_CHECK(ctx);
    }
    ctx->cur = dst;
    ctx->ftrial_valid = false;
    ctx->sorted_valid = false;
    ctx->coherent = true;
    ctx->prepared = false;
    return 0;
}
