// sort.cu -- K1: cell keys and the per-substep stable LSD radix sort of (key, slot) pairs.
// The reference never sorts; the sort exists so that p2g can aggregate same-cell
// particles and stage tiles in shared memory.  Contract: bit-exact against a CPU stable sort of the
// same keys (oracle/mpm_oracle.c orc_stable_sort).
#include "common.cuh"

// ---- keys -------------------------------------------------------------------------------
__global__ void k_keys(const float4* __restrict__ xm, uint32_t* __restrict__ keys, int64_t n, float inv_dx, int n_grid,
                       int nt, DevCounters* cnt) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 x = xm[i];
    bool cl = false;
    int bx = base_cell(x.x, inv_dx, n_grid, cl), by = base_cell(x.y, inv_dx, n_grid, cl), bz = base_cell(x.z, inv_dx, n_grid, cl);
    keys[i] = make_key(bx, by, bz, nt);
    if (cl) atomicAdd(&cnt->n_clamped, 1u);
}

int mpmk_launch_keys(mpmk_ctx* ctx, cudaStream_t s) {
    if (ctx->n == 0) return 0;
    k_keys<<<div_up(ctx->n, 256), 256, 0, s>>>((const float4*)ctx->b.set[ctx->cur].xm, (uint32_t*)ctx->b.keys, ctx->n,
                                               ctx->inv_dx, ctx->n_grid, ctx->nt, ctx->counters);
    MPMK_LAUNCH_CHECK(ctx);
    ctx->sorted_valid = false;
    ctx->prepared = false;
    return 0;
}

// ---- radix sort -------------------------------------------------------------------------
template <int BITS>
__global__ void __launch_bounds__(MPMK_SORT_CTA) k_sort_hist(const uint32_t* __restrict__ keys, int64_t n, int shift,
                                                             uint32_t* __restrict__ hist, int nctas) {
    constexpr int D = 1 << BITS;
    __shared__ uint32_t h[D];
    for (int d = threadIdx.x; d < D; d += MPMK_SORT_CTA) h[d] = 0;
    __syncthreads();
    const int64_t chunk = (int64_t)blockIdx.x * MPMK_SORT_CHUNK;
    const unsigned lane = threadIdx.x & 31;
    // all of this thread's keys first (independent loads in flight together), then the ranking
    uint32_t dg[MPMK_SORT_ROUNDS];
#pragma unroll
    for (int r = 0; r < MPMK_SORT_ROUNDS; ++r) {
        const int64_t idx = chunk + r * MPMK_SORT_CTA + threadIdx.x;
        dg[r] = (idx < n) ? ((keys[idx] >> shift) & (D - 1)) : (uint32_t)D;
    }
#pragma unroll
    for (int r = 0; r < MPMK_SORT_ROUNDS; ++r) {
        const uint32_t d = dg[r];
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        if (d < (uint32_t)D && (peers & ((1u << lane) - 1u)) == 0) atomicAdd(&h[d], (uint32_t)__popc(peers));
    }
    __syncthreads();
    for (int d = threadIdx.x; d < D; d += MPMK_SORT_CTA) hist[(size_t)d * nctas + blockIdx.x] = h[d];
}

// one CTA per digit row: exclusive scan of the row over sort CTAs; the last CTA to finish turns the row
// totals into digit bases (threadfence + counter, no spin-wait)
template <int BITS>
__global__ void __launch_bounds__(256) k_sort_rowscan(uint32_t* __restrict__ hist, uint32_t* __restrict__ tot,
                                                      uint32_t* __restrict__ done_counter, int nctas) {
    constexpr int D = 1 << BITS;
    __shared__ uint32_t part[256];
    __shared__ bool last;
    uint32_t* row = hist + (size_t)blockIdx.x * nctas;
    int seg = (nctas + 255) / 256;
    int lo = threadIdx.x * seg, hi = min(lo + seg, nctas);
    uint32_t s = 0;
    for (int i = lo; i < hi; ++i) s += row[i];
    part[threadIdx.x] = s;
    __syncthreads();
    // Hillis-Steele inclusive scan over 256 partials
    for (int o = 1; o < 256; o <<= 1) {
        uint32_t v = (threadIdx.x >= o) ? part[threadIdx.x - o] : 0u;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    uint32_t run = part[threadIdx.x] - s;
    for (int i = lo; i < hi; ++i) { uint32_t v = row[i]; row[i] = run; run += v; }
    if (threadIdx.x == 255) {
        tot[blockIdx.x] = part[255];
        __threadfence();
        last = (atomicAdd(done_counter, 1u) == (uint32_t)(D - 1));
    }
    __syncthreads();
    if (last) {
        __threadfence();
        // exclusive scan of D (<= 512) totals in place -> digit bases
        uint32_t a = (2 * threadIdx.x < D) ? ((volatile uint32_t*)tot)[2 * threadIdx.x] : 0u;
        uint32_t b = (2 * threadIdx.x + 1 < D) ? ((volatile uint32_t*)tot)[2 * threadIdx.x + 1] : 0u;
        part[threadIdx.x] = a + b;
        __syncthreads();
        for (int o = 1; o < 256; o <<= 1) {
            uint32_t v = (threadIdx.x >= o) ? part[threadIdx.x - o] : 0u;
            __syncthreads();
            part[threadIdx.x] += v;
            __syncthreads();
        }
        uint32_t ex = part[threadIdx.x] - (a + b);
        if (2 * threadIdx.x < D) tot[2 * threadIdx.x] = ex;
        if (2 * threadIdx.x + 1 < D) tot[2 * threadIdx.x + 1] = ex + a;
        if (threadIdx.x == 0) *done_counter = 0;
    }
}

template <int BITS>
__global__ void __launch_bounds__(MPMK_SORT_CTA) k_sort_scatter(const uint32_t* __restrict__ keys_in,
                                                                const uint32_t* __restrict__ vals_in,
                                                                uint32_t* __restrict__ keys_out,
                                                                uint32_t* __restrict__ vals_out, int64_t n, int shift,
                                                                const uint32_t* __restrict__ hist_scanned,
                                                                const uint32_t* __restrict__ base, int nctas) {
    constexpr int D = 1 << BITS;
    constexpr int NW = MPMK_SORT_CTA / 32;
    __shared__ uint32_t wh[NW][D];
    for (int i = threadIdx.x; i < NW * D; i += MPMK_SORT_CTA) (&wh[0][0])[i] = 0;
    __syncthreads();
    const int w = threadIdx.x >> 5;
    const unsigned lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const int64_t start = (int64_t)blockIdx.x * MPMK_SORT_CHUNK + (int64_t)w * (32 * MPMK_SORT_ROUNDS);
    uint32_t key[MPMK_SORT_ROUNDS];
    uint32_t val[MPMK_SORT_ROUNDS];
    uint32_t off[MPMK_SORT_ROUNDS];
    // every key and value of this thread is requested before the first use (the ranking below is a
    // serial chain through shared memory; loads issued inside it would each expose a DRAM latency)
#pragma unroll
    for (int r = 0; r < MPMK_SORT_ROUNDS; ++r) {
        const int64_t idx = start + r * 32 + lane;
        key[r] = (idx < n) ? keys_in[idx] : 0xffffffffu;
    }
#pragma unroll
    for (int r = 0; r < MPMK_SORT_ROUNDS; ++r) {
        const int64_t idx = start + r * 32 + lane;
        val[r] = (vals_in && idx < n) ? vals_in[idx] : (uint32_t)idx;
    }
#pragma unroll
    for (int r = 0; r < MPMK_SORT_ROUNDS; ++r) {
        const bool valid = start + r * 32 + lane < n;
        uint32_t d = valid ? ((key[r] >> shift) & (D - 1)) : (uint32_t)D;
        unsigned peers = __match_any_sync(0xffffffffu, d);
        uint32_t rank = __popc(peers & lt);
        uint32_t prev = valid ? wh[w][d] : 0u;
        __syncwarp();
        if (valid && rank == 0) wh[w][d] = prev + __popc(peers);
        __syncwarp();
        off[r] = prev + rank;
    }
    __syncthreads();
    for (int d = threadIdx.x; d < D; d += MPMK_SORT_CTA) {
        uint32_t run = base[d] + hist_scanned[(size_t)d * nctas + blockIdx.x];
#pragma unroll
        for (int ww = 0; ww < NW; ++ww) { uint32_t t = wh[ww][d]; wh[ww][d] = run; run += t; }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < MPMK_SORT_ROUNDS; ++r) {
        int64_t idx = start + r * 32 + lane;
        if (idx < n) {
            uint32_t d = (key[r] >> shift) & (D - 1);
            uint32_t pos = wh[w][d] + off[r];
            keys_out[pos] = key[r];
            vals_out[pos] = val[r];
        }
    }
}

template <int BITS>
static int sort_pass(mpmk_ctx* ctx, const uint32_t* kin, const uint32_t* vin, uint32_t* kout, uint32_t* vout, int shift,
                     cudaStream_t s) {
    int nctas = div_up(ctx->n, MPMK_SORT_CHUNK);
    k_sort_hist<BITS><<<nctas, MPMK_SORT_CTA, 0, s>>>(kin, ctx->n, shift, ctx->hist, nctas);
    MPMK_LAUNCH_CHECK(ctx);
    k_sort_rowscan<BITS><<<1 << BITS, 256, 0, s>>>(ctx->hist, ctx->hist_tot, ctx->hist_tot + 512, nctas);
    MPMK_LAUNCH_CHECK(ctx);
    k_sort_scatter<BITS><<<nctas, MPMK_SORT_CTA, 0, s>>>(kin, vin, kout, vout, ctx->n, shift, ctx->hist, ctx->hist_tot, nctas);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}

int mpmk_launch_sort(mpmk_ctx* ctx, cudaStream_t s) {
    if (ctx->n == 0) { ctx->sorted_valid = true; return 0; }
    int passes = (ctx->key_bits + MPMK_MAX_RADIX_BITS - 1) / MPMK_MAX_RADIX_BITS;
    int bits = (ctx->key_bits + passes - 1) / passes;
    if (bits < 5) bits = 5;
    uint32_t* kb[2] = {ctx->sorted_keys, ctx->keys_alt};
    uint32_t* vb[2] = {ctx->perm, ctx->perm_alt};
    const uint32_t* kin = (const uint32_t*)ctx->b.keys;
    const uint32_t* vin = nullptr;
    for (int p = 0; p < passes; ++p) {
        int o = (passes - 1 - p) & 1;
        int rc;
        switch (bits) {
            case 5: rc = sort_pass<5>(ctx, kin, vin, kb[o], vb[o], p * bits, s); break;
            case 6: rc = sort_pass<6>(ctx, kin, vin, kb[o], vb[o], p * bits, s); break;
            case 7: rc = sort_pass<7>(ctx, kin, vin, kb[o], vb[o], p * bits, s); break;
            case 8: rc = sort_pass<8>(ctx, kin, vin, kb[o], vb[o], p * bits, s); break;
            default: rc = sort_pass<9>(ctx, kin, vin, kb[o], vb[o], p * bits, s); break;
        }
        if (rc) return rc;
        kin = kb[o];
        vin = vb[o];
    }
    ctx->sorted_valid = true;
    ctx->prepared = false;
    return 0;
}
