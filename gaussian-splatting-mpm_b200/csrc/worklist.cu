// worklist.cu -- sparse grid bookkeeping around the sort:
//   * model.py:124-127 reset_grid_state, restricted to the 4x4x4-node blocks the previous substep
//     touched (untouched nodes are already zero, so this is bit-equivalent to the dense fills);
//   * the p2g/g2p work list: one item per (tile, chunk of <= MPMK_ITEM_MAX particles);
//   * the list of node blocks this substep's p2g can touch (consumed by the grid pass and by the
//     next substep's clear).
#include "common.cuh"

__global__ void k_clear_blocks(float4* __restrict__ gin, float4* __restrict__ gout, const uint32_t* __restrict__ list,
                               const uint32_t* __restrict__ count, int n, int nb, uint32_t* stamp) {
    if (blockIdx.x == 0 && threadIdx.x == 0) *stamp += 1u;  // consumed by k_build_items (next launch on the stream)
    const int64_t total = (int64_t)(*count) * 64;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        uint32_t b = list[i >> 6];
        int l = (int)(i & 63);
        int bz = b % nb, by = (b / nb) % nb, bx = b / (nb * nb);
        int ix = 4 * bx + (l >> 4), iy = 4 * by + ((l >> 2) & 3), iz = 4 * bz + (l & 3);
        if (ix < n && iy < n && iz < n) {
            size_t node = ((size_t)ix * n + iy) * n + iz;
            gin[node] = z;
            gout[node] = z;
        }
    }
}

__global__ void k_build_items(const uint32_t* __restrict__ skeys, int64_t n, uint4* __restrict__ items,
                              DevCounters* cnt, uint32_t* __restrict__ stamp_arr, const uint32_t* __restrict__ stamp_ptr,
                              uint32_t* __restrict__ blist, uint32_t* __restrict__ bcount, int nt, int nb) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    uint32_t tile = skeys[r] >> 6;
    if (r > 0 && (skeys[r - 1] >> 6) == tile) return;
    // first index past this tile's run
    int64_t lo = r + 1, hi = n;
    while (lo < hi) {
        int64_t mid = (lo + hi) >> 1;
        if ((skeys[mid] >> 6) > tile) hi = mid; else lo = mid + 1;
    }
    uint32_t np = (uint32_t)(lo - r);
    uint32_t ni = (np + MPMK_ITEM_MAX - 1) / MPMK_ITEM_MAX;
    uint32_t base = atomicAdd(&cnt->n_items, ni);
    atomicAdd(&cnt->n_tiles, 1u);
    for (uint32_t j = 0; j < ni; ++j) {
        uint32_t b0 = j * MPMK_ITEM_MAX;
        items[base + j] = make_uint4((uint32_t)r + b0, min((uint32_t)MPMK_ITEM_MAX, np - b0), tile, 0u);
    }
    int tz = tile % nt, ty = (tile / nt) % nt, tx = tile / (nt * nt);
    const uint32_t stamp = *stamp_ptr;
#pragma unroll
    for (int d = 0; d < 8; ++d) {
        int bx = tx + (d >> 2), by = ty + ((d >> 1) & 1), bz = tz + (d & 1);
        if (bx < nb && by < nb && bz < nb) {
            uint32_t b = (uint32_t)((bx * nb + by) * nb + bz);
            if (atomicExch(&stamp_arr[b], stamp) != stamp) blist[atomicAdd(bcount, 1u)] = b;
        }
    }
}

int mpmk_launch_prepare(mpmk_ctx* ctx, cudaStream_t s) {
    if (!ctx->sorted_valid) MPMK_FAIL(ctx, MPMK_ERR_STATE, "mpmk_prepare_grid: sort first");
    int prev = ctx->cur_list;
    int next = 1 - prev;
    // clear what the previous substep touched
    k_clear_blocks<<<ctx->sm_count * 8, 256, 0, s>>>((float4*)ctx->b.grid_in, (float4*)ctx->b.grid_out,
                                                     ctx->block_list[prev], ctx->block_count[prev], ctx->n_grid, ctx->nb,
                                                     ctx->dev_stamp);
    MPMK_LAUNCH_CHECK(ctx);
    if (ctx->mg.enabled) { int rc = mpmk_mg_clear_shared(ctx, s); if (rc) return rc; }
    MPMK_CHECK(ctx, cudaMemsetAsync(ctx->counters, 0, sizeof(DevCounters), s));
    MPMK_CHECK(ctx, cudaMemsetAsync(ctx->block_count[next], 0, sizeof(uint32_t), s));
    if (ctx->n > 0) {
        k_build_items<<<div_up(ctx->n, 256), 256, 0, s>>>(ctx->sorted_keys, ctx->n, ctx->items, ctx->counters,
                                                          ctx->block_stamp, ctx->dev_stamp, ctx->block_list[next],
                                                          ctx->block_count[next], ctx->nt, ctx->nb);
        MPMK_LAUNCH_CHECK(ctx);
    }
    ctx->cur_list = next;
    ctx->prepared = true;
    return 0;
}
