// g2p.cu -- K4: grid-to-particle gather (mpm_solver/utils.py:219-282) fused with
//   * the C / F_trial update (:266-279),
//   * update_cov (:187-215) when MPMK_TRACK_COV is set,
//   * compute_stress_from_F_trial for the NEXT substep (utils.py:14-54 + constitutive_models.py):
//     return mapping and Kirchhoff stress depend only on the F_trial produced here, so they run in
//     this kernel's epilogue instead of as a separate pass at the start of the next p2g2p(),
//   * the next substep's cell key (utils.py:94-95 base cell),
//   * the physical reorder: particle r of the sorted order is read through the permutation from the
//     source set and written to slot r of the destination set.
// The 27-node gather is contracted axis by axis (z, then y, then x) instead of node by node.
#include "common.cuh"
#include "constitutive.cuh"

struct G2PArgs {
    const uint32_t* perm;    // classic: source slot of sorted particle r.  LAG: rank, the destination slot of physical slot r
    const uint32_t* skeys;   // classic: sorted keys.  LAG: the current keys, in physical order
    // source set
    const float4 *xm, *f0, *f1, *f2, *cv0, *cv1;
    const int4* meta;
    // destination set
    float4 *oxm, *ovv, *oc0, *oc1, *ocs, *os1, *of0, *of1, *of2, *ocv0, *ocv1;
    int4* ometa;
    float4 *ft0, *ft1, *ft2;
    uint32_t* okeys;
    const uint4* items;
    const float4* grid_out;
    DevCounters* cnt;
    int64_t n;
    int n_grid, nt;
    float inv_dx, dt;
    ConstParams cp;
};

__device__ __forceinline__ void bspline_w(float f, float* w) {
    float wa = 1.5f - f, wb = f - 1.0f, wc = f - 0.5f;
    w[0] = wa * wa * 0.5f; w[1] = 0.75f - wb * wb; w[2] = wc * wc * 0.5f;
}

__device__ __forceinline__ void bspline(float f, float* w, float* dw) {
    float wa = 1.5f - f, wb = f - 1.0f, wc = f - 0.5f;
    w[0] = wa * wa * 0.5f; w[1] = 0.75f - wb * wb; w[2] = wc * wc * 0.5f;
    dw[0] = f - 1.5f; dw[1] = -2.0f * (f - 1.0f); dw[2] = f - 0.5f;
}

// The sums of utils.py:255-265 over a particle's 3x3x3 stencil: v = sum w g, the APIC moments M_d = sum w (off_d - f_d) g
// and the velocity-gradient columns G_d = sum (dw_d-weighted) g, g = the node velocities.
struct G2PSums {
    float v[3], Mx[3], My[3], Mz[3], Gx[3], Gy[3], Gz[3];
};

// Gather from the 3x3x3 nodes at g (x stride sx, y stride sy float4s).  SH: g points into the warp's shared tile (LDS);
// otherwise a generic pointer -- the global grid for a particle whose cell left the item's tile.
// The three families of sums differ in ONE axis each (the "special" axis carries w (off - f) or dw instead of w), so each
// needs only the plain-weight contraction of the other two axes: per loaded row (i, j) of three k-nodes
//     p = sum_k wz_k g_k            -> hx_i += wy_j p (x-special sums),  hy_j += wx_i p (y-special sums)
//     r_k += wy_j g_k               -> h_k += wx_i r_k at the end of plane i (z-special sums)
// and the three functionals (w, w (off - f), dw) are applied to the three-element vectors hx, hy, h at the end:
// 306 multiply-adds per particle against 441 for a z -> y -> x contraction that carries all seven sums through every level.
template <bool SH>
__device__ __forceinline__ void g2p_gather(const float4* __restrict__ g, const int sx, const int sy, const float fx, const float fy,
                                           const float fz, G2PSums& s) {
    float wx[3], wy[3], wz[3];
    bspline_w(fx, wx); bspline_w(fy, wy); bspline_w(fz, wz);
    float hy[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};   // [j][component]
    float h[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};    // [k][component]
#pragma unroll
    for (int c = 0; c < 3; ++c) { s.v[c] = 0.f; s.Mx[c] = 0.f; s.Gx[c] = 0.f; }
#pragma unroll 1
    for (int i = 0; i < 3; ++i) {
        const float wi = (i == 0) ? wx[0] : (i == 1) ? wx[1] : wx[2];
        float hx[3] = {0, 0, 0}, r[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            const float4* row = SH ? g + (i * 36 + j * 6) : g + (i * sx + j * sy);
            const float4 g0 = row[0], g1 = row[1], g2 = row[2];
            const float px = wz[0] * g0.x + wz[1] * g1.x + wz[2] * g2.x;
            const float py = wz[0] * g0.y + wz[1] * g1.y + wz[2] * g2.y;
            const float pz = wz[0] * g0.z + wz[1] * g1.z + wz[2] * g2.z;
            const float wj = wy[j];
            hx[0] += wj * px; hx[1] += wj * py; hx[2] += wj * pz;
            hy[j][0] += wi * px; hy[j][1] += wi * py; hy[j][2] += wi * pz;
            r[0][0] += wj * g0.x; r[0][1] += wj * g0.y; r[0][2] += wj * g0.z;
            r[1][0] += wj * g1.x; r[1][1] += wj * g1.y; r[1][2] += wj * g1.z;
            r[2][0] += wj * g2.x; r[2][1] += wj * g2.y; r[2][2] += wj * g2.z;
        }
        const float wii = wi * ((float)i - fx);
        const float di = (i == 0) ? fx - 1.5f : (i == 1) ? -2.0f * (fx - 1.0f) : fx - 0.5f;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            s.v[c] += wi * hx[c];
            s.Mx[c] += wii * hx[c];
            s.Gx[c] += di * hx[c];
            h[0][c] += wi * r[0][c];
            h[1][c] += wi * r[1][c];
            h[2][c] += wi * r[2][c];
        }
    }
    const float wyd[3] = {wy[0] * (0.f - fy), wy[1] * (1.f - fy), wy[2] * (2.f - fy)};
    const float dwy[3] = {fy - 1.5f, -2.0f * (fy - 1.0f), fy - 0.5f};
    const float wzd[3] = {wz[0] * (0.f - fz), wz[1] * (1.f - fz), wz[2] * (2.f - fz)};
    const float dwz[3] = {fz - 1.5f, -2.0f * (fz - 1.0f), fz - 0.5f};
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        s.My[c] = wyd[0] * hy[0][c] + wyd[1] * hy[1][c] + wyd[2] * hy[2][c];
        s.Gy[c] = dwy[0] * hy[0][c] + dwy[1] * hy[1][c] + dwy[2] * hy[2][c];
        s.Mz[c] = wzd[0] * h[0][c] + wzd[1] * h[1][c] + wzd[2] * h[2][c];
        s.Gz[c] = dwz[0] * h[0][c] + dwz[1] * h[1][c] + dwz[2] * h[2][c];
    }
}

// One particle after the gather: everything utils.py:266-279 and the next substep's utils.py:14-54 do with the sums.
// src / dst: the particle's slot in the source / destination set; pF0..pMeta: its records (shared stage).
__device__ __forceinline__ void g2p_particle(const G2PArgs& a, const uint32_t src, const uint32_t dst, const float4 X, const G2PSums& sm,
                                             const float4* pF0, const float4* pF1, const float4* pF2, const int4* pMeta) {
    const int n = a.n_grid;
    const float dt = a.dt;
    const float k4 = 4.0f * a.inv_dx;
    const float* v = sm.v;
    const float *Mx = sm.Mx, *My = sm.My, *Mz = sm.Mz, *Gx = sm.Gx, *Gy = sm.Gy, *Gz = sm.Gz;
    // utils.py:266-275: v, x, C = sum w v (x) (off - f) * 4 inv_dx
    const float nx = X.x + dt * v[0], ny = X.y + dt * v[1], nz = X.z + dt * v[2];
    a.oxm[dst] = make_float4(nx, ny, nz, X.w);
    a.ovv[dst] = make_float4(v[0], v[1], v[2], 0.f);
    bool cl = false;
    const int kx = base_cell(nx, a.inv_dx, n, cl), ky = base_cell(ny, a.inv_dx, n, cl), kz = base_cell(nz, a.inv_dx, n, cl);
    a.okeys[dst] = make_key(kx, ky, kz, a.nt);
    if (cl) atomicAdd(&a.cnt->n_clamped, 1u);
    a.oc0[dst] = make_float4(Mx[0] * k4, My[0] * k4, Mz[0] * k4, Mx[1] * k4);
    a.oc1[dst] = make_float4(My[1] * k4, Mz[1] * k4, Mx[2] * k4, My[2] * k4);
    const float C22 = Mz[2] * k4;
    // grad v = sum v (x) dweight  (row = velocity component, column = derivative direction)
    float gv[9];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        gv[3 * c + 0] = Gx[c] * a.inv_dx;
        gv[3 * c + 1] = Gy[c] * a.inv_dx;
        gv[3 * c + 2] = Gz[c] * a.inv_dx;
    }
    // F_trial = (I + dt grad v) F   (utils.py:276-279)
    const float4 F0 = *pF0, F1 = *pF1, F2 = *pF2;
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
    if (a.cp.flags & MPMK_WRITE_FTRIAL) {
        a.ft0[dst] = make_float4(Ft[0], Ft[1], Ft[2], Ft[3]);
        a.ft1[dst] = make_float4(Ft[4], Ft[5], Ft[6], Ft[7]);
        a.ft2[dst] = make_float4(Ft[8], F2.w, 0.f, 0.f);   // F2.w: the yield stress before this substep's return map
    }
    if (a.cp.flags & MPMK_TRACK_COV) {
        // update_cov, utils.py:187-215: cov += dt (grad_v cov + cov grad_v^T)
        const float4 q0 = ldg4(a.cv0 + src), q1 = ldg4(a.cv1 + src);
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
        a.ocv0[dst] = make_float4(o[0], o[1], o[2], o[4]);
        a.ocv1[dst] = make_float4(o[5], o[8], 0.f, 0.f);
    }
    const int4 meta = *pMeta;
    a.ometa[dst] = meta;
    // constitutive update for the next substep
    float mu = F2.y, lam = F2.z, ys = F2.w;
    float Fn[9], s6[6];
    constitutive_update(Ft, meta.y, mu, lam, ys, dt, a.cp, Fn, s6);
    const float vol = __int_as_float(meta.z);
    a.ocs[dst] = make_float4(C22, s6[0], s6[1], s6[2]);
    a.os1[dst] = make_float4(s6[3], s6[4], s6[5], vol);
    a.of0[dst] = make_float4(Fn[0], Fn[1], Fn[2], Fn[3]);
    a.of1[dst] = make_float4(Fn[4], Fn[5], Fn[6], Fn[7]);
    a.of2[dst] = make_float4(Fn[8], mu, lam, ys);
}

// One warp per work item (<= MPMK_ITEM_MAX sorted particles of one 4x4x4-cell tile).  The 6x6x6 nodes the
// tile's stencils reach are staged in the warp's shared-memory tile; lanes = particles gather from it
// in batches of 32.  The kernel is latency-bound unless every warp keeps global loads in flight while it
// computes (Little: ~40 KB per SM against 2.7 KB per batch), so the batches form a three-deep software
// pipeline that also crosses item boundaries:
//     batch C: permutation + key loads in flight (registers)
//     batch B: particle records xm,f0,f1,f2,meta in flight (cp.async, 16 B per lane per array, into
//              the other half of a double-buffered shared-memory stage)
//     batch A: gather + C/F update + constitutive update + stores
// 4 warps x 5 CTAs per SM: 20 warps at <= 102 registers each.  Measured on S-4M (jelly): 8x3 (cap 85, spills)
// 0.288 ms, 4x6 (cap 85) 0.291, 4x4 (cap 128) 0.305, 2x9 0.262, 4x5 = 2x10 0.221.
//
// LAG = true (lagged batches, api.cu): the work list comes from the sort of the previous substep, batches are 32
// consecutive physical slots, the true cell comes from the current key, and the particle is written to the slot the
// sort of THIS substep (done on the side branch by now) ranks it at.  Particles whose cell left the item's tile gather
// from the global grid instead of the warp's tile (same code: the gather pointer is per lane).
#ifndef G2P_WARPS
#define G2P_WARPS 4
#endif
#ifndef G2P_MIN_CTAS
#define G2P_MIN_CTAS 5
#endif
#define G2P_NREC 5

struct __align__(16) G2PWarpSmem {
    float4 rec[2][G2P_NREC][32];
    float4 tile[MPMK_TILE_NODES];
    uint32_t perm[2][32];  // permutation (LAG: rank) entries of batches B/C (ring), kept out of registers
    uint32_t key[3][32];   // keys of batches A/B/C (ring)
};

// The 6x6x6 nodes tile (tx, ty, tz)'s stencils reach -> the warp's shared tile (cp.async; out-of-grid nodes = 0).  Node
// i = lane + 32 t has local coordinates (x, y, z) = the digits of i in base 6, advanced by 32 = (0, 5, 2) per step.
__device__ __forceinline__ void g2p_stage_tile(float4* tile, const float4* __restrict__ grid, const int tx, const int ty, const int tz,
                                               const int n, const int lane) {
    int lx = lane / 36, ly = (lane / 6) % 6, lz = lane % 6;
    const float4* org = grid + (((size_t)(4 * tx) * n + 4 * ty) * n + 4 * tz);
    if ((4 * tx + 6 <= n) && (4 * ty + 6 <= n) && (4 * tz + 6 <= n)) {
#pragma unroll
        for (int t = 0; t < 7; ++t) {
            const int i = lane + 32 * t;
            if (i < MPMK_TILE_NODES) cp_async16(&tile[i], org + ((lx * n + ly) * n + lz));
            lz += 2; if (lz >= 6) { lz -= 6; ly += 1; }
            ly += 5; if (ly >= 6) { ly -= 6; lx += 1; }
        }
    } else {
        for (int i = lane; i < MPMK_TILE_NODES; i += 32) {
            if (4 * tx + lx < n && 4 * ty + ly < n && 4 * tz + lz < n) cp_async16(&tile[i], org + ((lx * n + ly) * n + lz));
            else tile[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            lz += 2; if (lz >= 6) { lz -= 6; ly += 1; }
            ly += 5; if (ly >= 6) { ly -= 6; lx += 1; }
        }
    }
}

template <bool LAG>
__global__ void __launch_bounds__(G2P_WARPS * 32, G2P_MIN_CTAS) k_g2p(const G2PArgs a) {
    pdl_wait();   // programmatic dependent launch: nothing of the previous grid is read or overwritten before this
    extern __shared__ float4 g2p_smem_raw[];
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    G2PWarpSmem& S = reinterpret_cast<G2PWarpSmem*>(g2p_smem_raw)[warp];
    const int n = a.n_grid;

    BatchStream bs;
    bs.init();
    const uint32_t n_items = a.cnt->n_items;
    BatchDesc dA = bs.next(lane, a.items, &a.cnt->g2p_cursor, n_items);
    if (dA.count == 0) return;
    // prologue: batch A's permutation entries synchronously, its records and batch B's entries async
    int tx = 0, ty = 0, tz = 0;   // tile of the current item
    int stage = 0;   // record stage holding batch A
    int kslot = 0;   // key ring slot of batch A (B = +1, C = +2 mod 3)
    int pslot = 0;   // perm ring slot of batch B (C = the other)
    uint32_t pA = 0;
    if (lane < dA.count) {
        pA = a.perm[dA.base + lane];
        S.key[0][lane] = a.skeys[dA.base + lane];
        const uint32_t sA = LAG ? dA.base + lane : pA;
        cp_async16(&S.rec[0][0][lane], a.xm + sA);
        cp_async16(&S.rec[0][1][lane], a.f0 + sA);
        cp_async16(&S.rec[0][2][lane], a.f1 + sA);
        cp_async16(&S.rec[0][3][lane], a.f2 + sA);
        cp_async16(&S.rec[0][4][lane], a.meta + sA);
    }
    BatchDesc dB = bs.next(lane, a.items, &a.cnt->g2p_cursor, n_items);
    if (lane < dB.count) {
        cp_async4(&S.perm[0][lane], a.perm + dB.base + lane);
        cp_async4(&S.key[1][lane], a.skeys + dB.base + lane);
    }
    cp_async_commit();

    while (dA.count > 0) {
        // batch A's records and batch B's permutation entries (issued one iteration ago) have landed
        cp_async_wait<0>();
        __syncwarp();
        if (dA.first) {
            // a new item: decode its tile (once per item, not per batch) and restage the node tile (the previous item's
            // gathers ended with __syncwarp).  Requesting the tile half a batch early -- right after the previous item's last
            // gather -- was measured: the 21 sums then live across the copies' address arithmetic spill (24 B) and the kernel
            // loses 2 % instead of winning the 4 % of its samples that wait here.
            tz = dA.tile % a.nt; ty = (dA.tile / a.nt) % a.nt; tx = dA.tile / (a.nt * a.nt);
            g2p_stage_tile(S.tile, a.grid_out, tx, ty, tz, n, lane);
        }
        cp_async_commit();
        uint32_t pB = 0;
        if (lane < dB.count) {
            pB = S.perm[pslot][lane];
            const uint32_t sB = LAG ? dB.base + lane : pB;
            cp_async16(&S.rec[stage ^ 1][0][lane], a.xm + sB);
            cp_async16(&S.rec[stage ^ 1][1][lane], a.f0 + sB);
            cp_async16(&S.rec[stage ^ 1][2][lane], a.f1 + sB);
            cp_async16(&S.rec[stage ^ 1][3][lane], a.f2 + sB);
            cp_async16(&S.rec[stage ^ 1][4][lane], a.meta + sB);
        }
        const BatchDesc dC = bs.next(lane, a.items, &a.cnt->g2p_cursor, n_items);
        const int kslotC = (kslot + 2) % 3;
        if (lane < dC.count) {
            cp_async4(&S.perm[pslot ^ 1][lane], a.perm + dC.base + lane);
            cp_async4(&S.key[kslotC][lane], a.skeys + dC.base + lane);
        }
        cp_async_commit();
        cp_async_wait<1>();  // the tile has landed; batch B's records / batch C's entries stay in flight
        __syncwarp();

        {
            const bool live = lane < dA.count;
            const uint32_t r = dA.base + lane;
            const uint32_t key = live ? S.key[kslot][lane] : 0u;
            const float4 X = S.rec[stage][0][lane];
            int bx = 4 * tx + (int)((key >> 4) & 3), by = 4 * ty + (int)((key >> 2) & 3), bz = 4 * tz + (int)(key & 3);   // base cell
            const int toff = (int)(((key >> 4) & 3) * 36 + ((key >> 2) & 3) * 6 + (key & 3));
            const float4* g = nullptr;
            int sx = 36, sy = 6;
            bool cross = false;
            if (LAG && live && (key >> 6) != dA.tile) {
                // its cell left this item's tile since the work list was built (~0.2 % of the particles): the same gather, from
                // the global grid
                cross = true;
                const uint32_t tt = key >> 6;
                bx = 4 * (int)(tt / (uint32_t)(a.nt * a.nt)) + (int)((key >> 4) & 3);
                by = 4 * (int)((tt / (uint32_t)a.nt) % (uint32_t)a.nt) + (int)((key >> 2) & 3);
                bz = 4 * (int)(tt % (uint32_t)a.nt) + (int)(key & 3);
                g = a.grid_out + (((size_t)bx * n + by) * n + bz);
                sx = n * n; sy = n;
                atomicAdd(&a.cnt->n_cross, 1u);
            }
            // a batch without such a particle (94 % of them) gathers with shared-memory loads at constant offsets; one with,
            // through generic loads that serve the tile and the global grid alike
            const bool any_cross = LAG && __any_sync(0xffffffffu, cross);
            if (live) {
                const float fx = __fmul_rn(X.x, a.inv_dx) - (float)bx;
                const float fy = __fmul_rn(X.y, a.inv_dx) - (float)by;
                const float fz = __fmul_rn(X.z, a.inv_dx) - (float)bz;
                G2PSums sm;
                if (!any_cross) g2p_gather<true>(S.tile + toff, 36, 6, fx, fy, fz, sm);
                else g2p_gather<false>(cross ? g : S.tile + toff, sx, sy, fx, fy, fz, sm);
                g2p_particle(a, LAG ? r : pA, LAG ? pA : r, X, sm, &S.rec[stage][1][lane], &S.rec[stage][2][lane],
                             &S.rec[stage][3][lane], reinterpret_cast<const int4*>(&S.rec[stage][4][lane]));
            }
        }
        __syncwarp();  // all lanes are done with the tile and this stage before either is overwritten
        dA = dB; pA = pB;
        dB = dC;
        stage ^= 1;
        pslot ^= 1;
        kslot = (kslot + 1) % 3;
    }
    cp_async_wait<0>();
}

// utils.py:14-54 on the current set in place (init, and after from_torch edits of F)
__global__ void k_constitutive(float4* f0, float4* f1, float4* f2, float4* cs, float4* s1, const int4* meta, int64_t n,
                               float dt, const ConstParams cp, const float4* ft0, const float4* ft1, const float4* ft2,
                               int use_ys_before) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    float4 F0 = f0[r], F1 = f1[r];
    const float4 F2 = f2[r];
    const int4 m = meta[r];
    float F22 = F2.x, ys = F2.w;
    if (ft0) {   // F_trial as the last g2p stored it (the current F has already been through the return map)
        F0 = ft0[r]; F1 = ft1[r];
        const float4 T2 = ft2[r];
        F22 = T2.x;
        if (use_ys_before) ys = T2.y;
    }
    const float Ft[9] = {F0.x, F0.y, F0.z, F0.w, F1.x, F1.y, F1.z, F1.w, F22};
    float mu = F2.y, lam = F2.z, Fn[9], s6[6];
    constitutive_update(Ft, m.y, mu, lam, ys, dt, cp, Fn, s6);
    f0[r] = make_float4(Fn[0], Fn[1], Fn[2], Fn[3]);
    f1[r] = make_float4(Fn[4], Fn[5], Fn[6], Fn[7]);
    f2[r] = make_float4(Fn[8], mu, lam, ys);
    float4 c = cs[r];
    cs[r] = make_float4(c.x, s6[0], s6[1], s6[2]);
    float4 s = s1[r];
    s1[r] = make_float4(s6[3], s6[4], s6[5], s.w);
}

static ConstParams make_cp(const mpmk_ctx* ctx) {
    ConstParams cp;
    cp.alpha = ctx->prm.alpha; cp.plastic_viscosity = ctx->prm.plastic_viscosity; cp.xi = ctx->prm.xi;
    cp.softening = ctx->prm.softening; cp.hardening = ctx->prm.hardening; cp.flags = ctx->prm.flags;
    return cp;
}

int mpmk_g2p_init(mpmk_ctx* ctx) {
    const int smem = (int)sizeof(G2PWarpSmem) * G2P_WARPS;
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_g2p<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_g2p<false>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_g2p<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_g2p<true>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    int per_sm = 0;
    MPMK_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_g2p<false>, G2P_WARPS * 32, smem));
    ctx->g2p_ctas_per_sm = per_sm > 0 ? per_sm : 1;
    return 0;
}

static int launch_g2p(mpmk_ctx* ctx, float dt, cudaStream_t s, bool lag) {
    const int src = ctx->cur, dst = 1 - ctx->cur;
    uint32_t* keys_out = lag ? (ctx->keys_cur == (uint32_t*)ctx->b.keys ? ctx->keys_b : (uint32_t*)ctx->b.keys) : (uint32_t*)ctx->b.keys;
    if (ctx->n > 0) {
        const mpmk_particle_set& A = ctx->b.set[src];
        const mpmk_particle_set& B = ctx->b.set[dst];
        G2PArgs a;
        a.perm = lag ? ctx->rank : ctx->perm; a.skeys = lag ? ctx->keys_cur : ctx->sorted_keys;
        a.xm = (const float4*)A.xm; a.f0 = (const float4*)A.f0; a.f1 = (const float4*)A.f1; a.f2 = (const float4*)A.f2;
        a.cv0 = (const float4*)A.cv0; a.cv1 = (const float4*)A.cv1; a.meta = (const int4*)A.meta;
        a.oxm = (float4*)B.xm; a.ovv = (float4*)B.vv; a.oc0 = (float4*)B.c0; a.oc1 = (float4*)B.c1; a.ocs = (float4*)B.cs;
        a.os1 = (float4*)B.s1; a.of0 = (float4*)B.f0; a.of1 = (float4*)B.f1; a.of2 = (float4*)B.f2;
        a.ocv0 = (float4*)B.cv0; a.ocv1 = (float4*)B.cv1; a.ometa = (int4*)B.meta;
        a.ft0 = (float4*)ctx->b.ft0; a.ft1 = (float4*)ctx->b.ft1; a.ft2 = (float4*)ctx->b.ft2;
        a.items = ctx->items;
        a.okeys = keys_out; a.grid_out = (const float4*)ctx->b.grid_out; a.cnt = ctx->counters;
        a.n = ctx->n; a.n_grid = ctx->n_grid; a.nt = ctx->nt; a.inv_dx = ctx->inv_dx; a.dt = dt;
        a.cp = make_cp(ctx);
        if ((a.cp.flags & MPMK_TRACK_COV) && (!A.cv0 || !A.cv1 || !B.cv0 || !B.cv1))
            MPMK_FAIL(ctx, MPMK_ERR_STATE, "MPMK_TRACK_COV needs cv0/cv1 bound");
        if ((a.cp.flags & MPMK_WRITE_FTRIAL) && (!a.ft0 || !a.ft1 || !a.ft2))
            MPMK_FAIL(ctx, MPMK_ERR_STATE, "MPMK_WRITE_FTRIAL needs ft0..ft2 bound");
        // persistent grid, warps fetch work items dynamically (counters->g2p_cursor, zeroed by prepare)
        int64_t want = (ctx->n / 32 + G2P_WARPS - 1) / G2P_WARPS + 1;
        int64_t cap = (int64_t)ctx->sm_count * ctx->g2p_ctas_per_sm;
        const dim3 grid((int)(want < cap ? want : cap));
        if (lag) {
            mpmk_launch(k_g2p<true>, grid, dim3(G2P_WARPS * 32), sizeof(G2PWarpSmem) * G2P_WARPS, s, a);

        } else {
            mpmk_launch(k_g2p<false>, grid, dim3(G2P_WARPS * 32), sizeof(G2PWarpSmem) * G2P_WARPS, s, a);
        }
        MPMK_LAUNCH_CHECK(ctx);
    }
    ctx->cur = dst;
    if (lag) ctx->keys_cur = keys_out;
    ctx->ftrial_valid = (ctx->prm.flags & MPMK_WRITE_FTRIAL) != 0;
    ctx->sorted_valid = false;  // the permutation referred to the old physical order
    ctx->coherent = true;       // ... but sorted_keys[r] is now the old key of physical slot r (sort.cu, coherent sort)
    ctx->prepared = false;
    return 0;
}

int mpmk_launch_g2p(mpmk_ctx* ctx, float dt, cudaStream_t s) {
    if (!ctx->prepared) MPMK_FAIL(ctx, MPMK_ERR_STATE, "mpmk_g2p: sort + prepare_grid first");
    ctx->lag_ready = false;
    return launch_g2p(ctx, dt, s, false);
}

int mpmk_launch_g2p_lagged(mpmk_ctx* ctx, float dt, cudaStream_t s) { return launch_g2p(ctx, dt, s, true); }

extern "C" int mpmk_constitutive_ex(mpmk_ctx* ctx, float dt, int use_ftrial, int use_ys_before, void* stream) {
    if (!ctx || !ctx->bound) return MPMK_ERR_STATE;
    if (ctx->n == 0) return 0;
    const bool ft = use_ftrial && ctx->ftrial_valid && ctx->b.ft0 && ctx->b.ft1 && ctx->b.ft2;
    const mpmk_particle_set& A = ctx->b.set[ctx->cur];
    k_constitutive<<<div_up(ctx->n, 128), 128, 0, (cudaStream_t)stream>>>(
        (float4*)A.f0, (float4*)A.f1, (float4*)A.f2, (float4*)A.cs, (float4*)A.s1, (const int4*)A.meta, ctx->n, dt, make_cp(ctx),
        ft ? (const float4*)ctx->b.ft0 : nullptr, ft ? (const float4*)ctx->b.ft1 : nullptr, ft ? (const float4*)ctx->b.ft2 : nullptr,
        (ft && use_ys_before) ? 1 : 0);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}

extern "C" int mpmk_constitutive(mpmk_ctx* ctx, float dt, void* stream) { return mpmk_constitutive_ex(ctx, dt, 0, 0, stream); }
