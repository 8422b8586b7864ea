// worklist.cu -- sparse grid bookkeeping around the sort:
//   * model.py:124-127 reset_grid_state, restricted to the 4x4x4-node blocks the previous substep
//     touched (untouched nodes are already zero, so this is bit-equivalent to the dense fills);
//   * the p2g/g2p work list: one item per (tile, chunk of <= MPMK_ITEM_MAX particles);
//   * the list of node blocks this substep's p2g can touch (consumed by the grid pass and by the
//     next substep's clear).
#include "common.cuh"
#include <cstdlib>

__global__ void k_clear_blocks(float4* __restrict__ gin, float4* __restrict__ gout, const uint32_t* __restrict__ list,
                               const uint32_t* __restrict__ count, int n, int nb, uint32_t* stamp, DevCounters* cnt,
                               uint32_t* next_count, int clear_in) {
    pdl_wait();   // programmatic dependent launch: nothing of the previous grid is read or overwritten before this
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        *stamp += 1u;  // consumed by k_build_items (a later launch on the stream)
        // this substep's counters / block list start empty (the previous substep's kernels are done with them)
        cnt->n_items = 0u; cnt->n_active_blocks = 0u; cnt->n_tiles = 0u; cnt->n_clamped = 0u;
        cnt->p2g_cursor = 0u; cnt->g2p_cursor = 0u; cnt->n_cross = 0u; cnt->list_stamp = *stamp;
        *next_count = 0u;
    }
    const int64_t total = (int64_t)(*count) * 64;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        uint32_t b = list[i >> 6];
        int l = (int)(i & 63);
        int bz = b % nb, by = (b / nb) % nb, bx = b / (nb * nb);
        int ix = 4 * bx + (l >> 4), iy = 4 * by + ((l >> 2) & 3), iz = 4 * bz + (l & 3);
        if (ix < n && iy < n && iz < n) {
            size_t node = ((size_t)ix * n + iy) * n + iz;
            if (clear_in) gin[node] = z;   // fused batches: grid_in already holds this substep's scatter
            gout[node] = z;
        }
    }
}

// pass 1 over the sorted keys: the first particle of every tile run appends {tile, start} to the tile
// list, the last one records the run's end in the tile-indexed table (no search, no dependent loads)
__global__ void __launch_bounds__(256) k_tile_runs(const uint32_t* __restrict__ skeys, int64_t n, uint2* __restrict__ tile_list,
                                                   uint32_t* __restrict__ tile_end, DevCounters* cnt) {
    pdl_wait();   // programmatic dependent launch: nothing of the previous grid is read or overwritten before this
    // sixteen consecutive keys per thread (four 16-byte loads in flight); the neighbours' edge keys come through
    // shuffles, only the warp's two outer lanes touch memory again; one list append per warp
    constexpr int K = 16;
    const unsigned lane = threadIdx.x & 31;
    for (int64_t r0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * K; r0 - K * (int64_t)lane < n; r0 += (int64_t)gridDim.x * blockDim.x * K) {
        uint32_t t[K];
        if (r0 + K <= n) {
            const uint4* p = reinterpret_cast<const uint4*>(skeys + r0);
            uint4 v[K / 4];
#pragma unroll
            for (int q = 0; q < K / 4; ++q) v[q] = p[q];
#pragma unroll
            for (int q = 0; q < K / 4; ++q) { t[4 * q] = v[q].x >> 6; t[4 * q + 1] = v[q].y >> 6; t[4 * q + 2] = v[q].z >> 6; t[4 * q + 3] = v[q].w >> 6; }
        } else {
#pragma unroll
            for (int q = 0; q < K; ++q) t[q] = (r0 + q < n) ? (skeys[r0 + q] >> 6) : 0xffffffffu;
        }
        uint32_t prev = __shfl_up_sync(0xffffffffu, t[K - 1], 1);
        uint32_t next = __shfl_down_sync(0xffffffffu, t[0], 1);
        if (lane == 0) prev = (r0 > 0 && r0 - 1 < n) ? (skeys[r0 - 1] >> 6) : 0xffffffffu;
        if (lane == 31) next = (r0 + K < n) ? (skeys[r0 + K] >> 6) : 0xffffffffu;
        uint32_t starts = 0u;
#pragma unroll
        for (int q = 0; q < K; ++q) {
            if (r0 + q < n) {
                const uint32_t p = q ? t[q - 1] : prev;
                const uint32_t nx = q < K - 1 ? t[q + 1] : next;
                if (t[q] != p) starts |= 1u << q;
                if (t[q] != nx) tile_end[t[q]] = (uint32_t)(r0 + q) + 1u;
            }
        }
        const uint32_t c = __popc(starts);
        uint32_t incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (unsigned)o) incl += u; }
        const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
        if (total) {
            uint32_t base = 0u;
            if (lane == 31) base = atomicAdd(&cnt->n_tiles, total);
            uint32_t at = __shfl_sync(0xffffffffu, base, 31) + incl - c;
#pragma unroll
            for (int q = 0; q < K; ++q)
                if (starts & (1u << q)) tile_list[at++] = make_uint2(t[q], (uint32_t)(r0 + q));
        }
    }
}

// pass 2 over the non-empty tiles: work items (chunks of <= MPMK_ITEM_MAX particles) and the node blocks
// the tile's stencils reach.  G lanes share a tile: the reference-sized grids (50^3 .. 100^3 cells) have a few
// hundred tiles holding thousands of particles each, and one thread writing a tile's dozens of items was the
// longest of the small kernels there; the big sparse grids keep one thread per tile.
template <int G>
__global__ void __launch_bounds__(256) k_build_items(const uint2* __restrict__ tile_list, const uint32_t* __restrict__ tile_end,
                                                     uint4* __restrict__ items, DevCounters* cnt, uint32_t* __restrict__ stamp_arr,
                                                     const uint32_t* __restrict__ stamp_ptr, uint32_t* __restrict__ blist,
                                                     uint32_t* __restrict__ bcount, int nt, int nb, uint32_t item_max) {
    pdl_wait();   // programmatic dependent launch: nothing of the previous grid is read or overwritten before this
    const uint32_t n_tiles = cnt->n_tiles;
    const uint32_t stamp = *stamp_ptr;
    const uint32_t gl = threadIdx.x % G;   // lane within the tile's group
    const uint32_t groups = gridDim.x * (256 / G);
    // whole warps stay in the loop together (the shuffle below needs every lane)
    for (uint32_t t0 = blockIdx.x * (256 / G) + (threadIdx.x & ~31u) / G; t0 < n_tiles; t0 += groups) {
        const uint32_t t = t0 + (threadIdx.x & 31u) / G;
        const bool live = t < n_tiles;
        uint2 tl = make_uint2(0u, 0u);
        uint32_t np = 0, ni = 0, base = 0;
        if (live) {
            tl = tile_list[t];
            np = tile_end[tl.x] - tl.y;
            ni = (np + item_max - 1) / item_max;
        }
        if (G == 1) {   // one append per warp
            const unsigned lane = threadIdx.x & 31u;
            uint32_t incl = ni;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (unsigned)o) incl += u; }
            if (lane == 31u) base = atomicAdd(&cnt->n_items, incl);
            base = __shfl_sync(0xffffffffu, base, 31) + incl - ni;
        } else {
            if (live && gl == 0) base = atomicAdd(&cnt->n_items, ni);
            base = __shfl_sync(0xffffffffu, base, 0, G);
        }
        const uint32_t tile = tl.x, r = tl.y;
        for (uint32_t j = gl; j < ni; j += G) {
            const uint32_t b0 = j * item_max;
            items[base + j] = make_uint4(r + b0, min(item_max, np - b0), tile, 0u);
        }
        const int tz = tile % nt, ty = (tile / nt) % nt, tx = tile / (nt * nt);
#pragma unroll
        for (int d0 = 0; d0 < 8; d0 += G) {
            const int d = d0 + (int)(gl & 7u);
            bool fresh = false;
            uint32_t b = 0u;
            if (live && (G <= 8 || gl < 8)) {
                const int bx = tx + (d >> 2), by = ty + ((d >> 1) & 1), bz = tz + (d & 1);
                if (bx < nb && by < nb && bz < nb) {
                    b = (uint32_t)((bx * nb + by) * nb + bz);
                    fresh = atomicExch(&stamp_arr[b], stamp) != stamp;
                }
            }
            // newly stamped blocks: one list append per warp
            const unsigned m = __ballot_sync(0xffffffffu, fresh);
            if (m) {
                const unsigned lane = threadIdx.x & 31u;
                const int leader = __ffs(m) - 1;
                uint32_t bbase = 0u;
                if ((int)lane == leader) bbase = atomicAdd(bcount, (uint32_t)__popc(m));
                bbase = __shfl_sync(0xffffffffu, bbase, leader);
                if (fresh) blist[bbase + __popc(m & ((1u << lane) - 1u))] = b;
            }
        }
    }
}

// Particles per work item: MPMK_ITEM_MAX when that still yields a few items per resident warp, else smaller
// (a multiple of the 32-particle batch) so that small scenes -- a few hundred tiles holding hundreds of particles
// per cell, like the reference's 50^3 configs -- still spread over all SMs.
int mpmk_item_size(const mpmk_ctx* ctx) {
    static const int forced = getenv("MPMK_ITEM_SIZE") ? atoi(getenv("MPMK_ITEM_SIZE")) : 0;   // tuning experiments
    if (forced >= MPMK_ITEM_MIN && forced <= MPMK_ITEM_MAX) return forced / 32 * 32;   // (the items array is sized for MPMK_ITEM_MIN)
    const int64_t warps = (int64_t)ctx->sm_count * ctx->g2p_ctas_per_sm * 4;  // resident g2p warps (the larger grid)
    const int64_t want_items = 4 * warps;   // (2 x was tried: lego-like -2 %, cupnoodle-like +4 %)
    int64_t sz = MPMK_ITEM_MAX;
    if (ctx->n / MPMK_ITEM_MAX < want_items) sz = ((ctx->n / want_items + 31) / 32) * 32;
    if (sz < MPMK_ITEM_MIN) sz = MPMK_ITEM_MIN;
    if (sz > MPMK_ITEM_MAX) sz = MPMK_ITEM_MAX;
    return (int)sz;
}

// tile runs + work items + block list of the current sort, into work-list set `w` / block list `l` (whose counters
// and count have been reset, and whose stamp has been advanced, by the kernel that ran before)
static int launch_build(mpmk_ctx* ctx, int w, int l, cudaStream_t s) {
    if (ctx->n <= 0) return 0;
    DevCounters* cnt = ctx->counters_buf[w];
    mpmk_launch(k_tile_runs, dim3(min(div_up(ctx->n, 4096), ctx->sm_count * 8)), dim3(256), 0, s, ctx->sorted_keys, ctx->n, ctx->tile_list, ctx->tile_end, cnt);
    MPMK_LAUNCH_CHECK(ctx);
    // at most nt^3 tiles exist: with few of them a whole warp takes one tile
    const int64_t tile_bound = min((int64_t)ctx->nt * ctx->nt * ctx->nt, ctx->n);
    const uint32_t item_max = (uint32_t)mpmk_item_size(ctx);
    if (tile_bound <= 16384)
        mpmk_launch(k_build_items<32>, dim3((int)min((int64_t)ctx->sm_count * 8, (int64_t)div_up(tile_bound, 8))), dim3(256), 0, s,
            ctx->tile_list, ctx->tile_end, ctx->items_buf[w], cnt, ctx->block_stamp[l], ctx->dev_stamp, ctx->block_list[l],
            ctx->block_count[l], ctx->nt, ctx->nb, item_max);
    else
        mpmk_launch(k_build_items<1>, dim3(ctx->sm_count), dim3(256), 0, s, ctx->tile_list, ctx->tile_end, ctx->items_buf[w], cnt, ctx->block_stamp[l],
                                                       ctx->dev_stamp, ctx->block_list[l], ctx->block_count[l], ctx->nt,
                                                       ctx->nb, item_max);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}

static void use_worklist(mpmk_ctx* ctx, int w) {
    ctx->wl = w;
    ctx->items = ctx->items_buf[w];
    ctx->counters = ctx->counters_buf[w];
}

int mpmk_launch_prepare(mpmk_ctx* ctx, cudaStream_t s) {
    if (!ctx->sorted_valid) MPMK_FAIL(ctx, MPMK_ERR_STATE, "mpmk_prepare_grid: sort first");
    int prev = ctx->cur_list;
    int next = 1 - prev;
    const int w = 1 - ctx->wl;
    // clear what the previous substep touched; reset the per-substep counters
    mpmk_launch(k_clear_blocks, dim3(ctx->sm_count * 8), dim3(256), 0, s, (float4*)ctx->b.grid_in, (float4*)ctx->b.grid_out,
                                                     ctx->block_list[prev], ctx->block_count[prev], ctx->n_grid, ctx->nb,
                                                     ctx->dev_stamp, ctx->counters_buf[w], ctx->block_count[next], ctx->keep_grid_in ? 0 : 1);
    MPMK_LAUNCH_CHECK(ctx);
    // (slab mode: the shared planes are reset by k_grid_planes, which visits every node of them each substep)
    { int rc = launch_build(ctx, w, next, s); if (rc) return rc; }
    use_worklist(ctx, w);
    ctx->cur_list = next;
    ctx->prepared = true;
    ctx->lag_ready = false;
    return 0;
}

// ---- lagged batches (api.cu) ----------------------------------------------------------------------
// From the second substep of a batch on, p2g / grid / g2p of substep s run on the work list and the block list that
// were built from the sort of substep s-1 (the physical order IS that sorted order; a particle's true cell comes
// from its current key), while the sort of substep s and the lists for substep s+1 are built on a side branch.

// Clears grid_out over block list [l] (the blocks of the substep before the last: g2p has read them), then -- last
// CTA out -- resets that list's count and the counters of work-list set [w] for the rebuild that follows on the stream
// (the same buffers: a separate launch would only be there to order the reset after the clear).
__global__ void __launch_bounds__(256) k_lag_clear_out(float4* __restrict__ gout, const uint32_t* __restrict__ list, uint32_t* count,
                                                       int n, int nb, uint32_t* stamp, DevCounters* cnt, uint32_t* done) {
    const int64_t total = (int64_t)(*count) * 64;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        uint32_t b = list[i >> 6];
        int l = (int)(i & 63);
        int bz = b % nb, by = (b / nb) % nb, bx = b / (nb * nb);
        int ix = 4 * bx + (l >> 4), iy = 4 * by + ((l >> 2) & 3), iz = 4 * bz + (l & 3);
        if (ix < n && iy < n && iz < n) gout[((size_t)ix * n + iy) * n + iz] = z;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(done, 1u) == gridDim.x - 1) {
            *done = 0u;
            *stamp += 1u;
            cnt->n_items = 0u; cnt->n_active_blocks = 0u; cnt->n_tiles = 0u; cnt->n_clamped = 0u;
            cnt->p2g_cursor = 0u; cnt->g2p_cursor = 0u; cnt->n_cross = 0u; cnt->list_stamp = *stamp;
            *count = 0u;
        }
    }
}

// classic -> lagged: the last (classic) substep left its scatter in grid_in and its velocities in grid_out over block
// list [l], and used up the cursors of its work list; the lagged substep that follows runs on that same work list
// and block list, from a clean grid
__global__ void __launch_bounds__(256) k_lag_enter(float4* __restrict__ gin, float4* __restrict__ gout, const uint32_t* __restrict__ list,
                                                   const uint32_t* __restrict__ count, int n, int nb, const uint32_t* stamp, DevCounters* cnt) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        cnt->p2g_cursor = 0u; cnt->g2p_cursor = 0u; cnt->n_cross = 0u; cnt->n_clamped = 0u; cnt->list_stamp = *stamp;
    }
    const int64_t total = (int64_t)(*count) * 64;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        uint32_t b = list[i >> 6];
        int l = (int)(i & 63);
        int bz = b % nb, by = (b / nb) % nb, bx = b / (nb * nb);
        int ix = 4 * bx + (l >> 4), iy = 4 * by + ((l >> 2) & 3), iz = 4 * bz + (l & 3);
        if (ix < n && iy < n && iz < n) {
            const size_t node = ((size_t)ix * n + iy) * n + iz;
            gin[node] = z; gout[node] = z;
        }
    }
}

// grid_in over block list [cur_list], zeroed beside g2p (which only reads grid_out) for the next substep's scatter.
// (Zeroing it behind the grid pass, in k_grid, makes that kernel five times slower -- 12 -> 60 us on S-4M.)
__global__ void __launch_bounds__(128) k_lag_clear_in(float4* __restrict__ gin, const uint32_t* __restrict__ list, const uint32_t* __restrict__ count,
                                                      int n, int nb) {
    const int64_t total = (int64_t)(*count) * 64;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        uint32_t b = list[i >> 6];
        int l = (int)(i & 63);
        int bz = b % nb, by = (b / nb) % nb, bx = b / (nb * nb);
        int ix = 4 * bx + (l >> 4), iy = 4 * by + ((l >> 2) & 3), iz = 4 * bz + (l & 3);
        if (ix < n && iy < n && iz < n) gin[((size_t)ix * n + iy) * n + iz] = z;
    }
}

int mpmk_launch_lag_clear_in(mpmk_ctx* ctx, cudaStream_t s) {
    const int l = ctx->cur_list;
    // (128 threads x 32 registers: fits beside five g2p CTAs)
    mpmk_launch(k_lag_clear_in, dim3(ctx->sm_count * 4), dim3(128), 0, s, (float4*)ctx->b.grid_in, ctx->block_list[l], ctx->block_count[l], ctx->n_grid, ctx->nb);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}

int mpmk_worklist_init(mpmk_ctx* ctx) {
    // side-branch kernels of a lagged batch share SMs with the persistent p2g / g2p grids: same shared-memory carve-out
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_lag_clear_in, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_lag_clear_out, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_tile_runs, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_build_items<1>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_build_items<32>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    return 0;
}

// precondition: ctx->coherent after a classic substep (work list [wl] and block list [cur_list] describe the current
// physical order).  Leaves the context as a lagged substep leaves it: the lists to use next are the "other" ones.
int mpmk_launch_lag_enter(mpmk_ctx* ctx, cudaStream_t s) {
    const int l = ctx->cur_list;
    mpmk_launch(k_lag_enter, dim3(ctx->sm_count * 8), dim3(256), 0, s, (float4*)ctx->b.grid_in, (float4*)ctx->b.grid_out, ctx->block_list[l],
                ctx->block_count[l], ctx->n_grid, ctx->nb, ctx->dev_stamp, ctx->counters_buf[ctx->wl]);
    MPMK_LAUNCH_CHECK(ctx);
    use_worklist(ctx, 1 - ctx->wl);
    ctx->cur_list = 1 - l;
    ctx->lag_ready = true;
    return 0;
}

// start of a lagged substep (host side): the prebuilt lists become the current ones
void mpmk_lag_flip(mpmk_ctx* ctx) {
    use_worklist(ctx, 1 - ctx->wl);
    ctx->cur_list = 1 - ctx->cur_list;
}

// side branch, first kernel: grid_out of the substep before this one, and the reset for the rebuild
int mpmk_launch_lag_clear_out(mpmk_ctx* ctx, cudaStream_t s) {
    const int l = 1 - ctx->cur_list, w = 1 - ctx->wl;
    mpmk_launch(k_lag_clear_out, dim3(ctx->sm_count * 2), dim3(256), 0, s, (float4*)ctx->b.grid_out, ctx->block_list[l], ctx->block_count[l], ctx->n_grid,
                ctx->nb, ctx->dev_stamp, ctx->counters_buf[w], ctx->lag_done);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}

// side branch, after the sort: lists for the NEXT substep into the other set
int mpmk_launch_prepare_lagged(mpmk_ctx* ctx, cudaStream_t s) {
    return launch_build(ctx, 1 - ctx->wl, 1 - ctx->cur_list, s);
}
