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
// One-sweep LSD radix sort: a single pass over the keys builds the global digit histograms of every pass
// (k_sort_digit_hist); each pass is then ONE kernel (k_sort_onesweep) in which a CTA ranks its tile of
// MPMK_SORT_CHUNK keys, publishes its per-digit counts, obtains the counts of all earlier tiles by decoupled
// look-back (status words carry {flag, count}; tiles are handed out by an atomic counter, so every tile a CTA
// waits for is already running), stages the tile in shared memory in sorted order and writes it out in
// contiguous runs.  Stable: tiles, warps, rounds and lanes are all visited in key order.
#define OS_FLAG_AGG 0x40000000u   // status word holds this tile's own count
#define OS_FLAG_INC 0x80000000u   // status word holds the count of this tile and all earlier ones
#define OS_MASK 0x3fffffffu
#define OS_MAX_PASSES 4
// control block layout (uint32 words, zeroed once per sort): [0..3] tile counters per pass, [4] blocks done,
// [8 ..] global digit histograms [OS_MAX_PASSES][512], then the status words [passes][ntiles][D]
#define OS_CTRL_HIST 8
#define OS_CTRL_STATUS (OS_CTRL_HIST + OS_MAX_PASSES * 512)

template <int BITS>
__global__ void __launch_bounds__(256) k_sort_digit_hist(const uint32_t* __restrict__ keys, int64_t n, int passes,
                                                         uint32_t* __restrict__ ctrl, uint32_t* __restrict__ digit_base) {
    constexpr int D = 1 << BITS;
    __shared__ uint32_t h[OS_MAX_PASSES][D];
    __shared__ bool last;
    for (int i = threadIdx.x; i < OS_MAX_PASSES * D; i += 256) (&h[0][0])[i] = 0;
    __syncthreads();
    // every thread owns 16 CONSECUTIVE keys: the keys are nearly sorted (physical order = last substep's order),
    // so a thread sees a handful of digit runs and issues one shared atomic per run, not per key
    const int64_t nchunks = (n + 4095) / 4096;
    for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
        const int64_t base = c * 4096 + (int64_t)threadIdx.x * 16;
        uint32_t k[16];
        if (base + 16 <= n) {
            const uint4* p = reinterpret_cast<const uint4*>(keys + base);
#pragma unroll
            for (int q = 0; q < 4; ++q) { const uint4 v = p[q]; k[4 * q] = v.x; k[4 * q + 1] = v.y; k[4 * q + 2] = v.z; k[4 * q + 3] = v.w; }
        } else {
#pragma unroll
            for (int q = 0; q < 16; ++q) k[q] = (base + q < n) ? keys[base + q] : 0xffffffffu;
        }
        const int cnt = (int)max((int64_t)0, min((int64_t)16, n - base));
        for (int p = 0; p < passes; ++p) {
            const int shift = p * BITS;
            uint32_t run_d = 0, run_n = 0;
#pragma unroll
            for (int q = 0; q < 16; ++q) {
                if (q < cnt) {
                    const uint32_t d = (k[q] >> shift) & (D - 1);
                    if (run_n && d != run_d) { atomicAdd(&h[p][run_d], run_n); run_n = 0; }
                    run_d = d; ++run_n;
                }
            }
            if (run_n) atomicAdd(&h[p][run_d], run_n);
        }
    }
    __syncthreads();
    uint32_t* ghist = ctrl + OS_CTRL_HIST;
    for (int i = threadIdx.x; i < passes * D; i += 256) {
        const int p = i / D, d = i % D;
        const uint32_t v = h[p][d];
        if (v) atomicAdd(&ghist[p * 512 + d], v);
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = (atomicAdd(&ctrl[4], 1u) == gridDim.x - 1);
    __syncthreads();
    if (!last) return;
    __threadfence();
    // the last block turns the histograms into exclusive digit bases (D <= 512, two digits per thread)
    __shared__ uint32_t part[256];
    for (int p = 0; p < passes; ++p) {
        const volatile uint32_t* g = ghist + p * 512;
        const uint32_t a = (2 * threadIdx.x < D) ? g[2 * threadIdx.x] : 0u;
        const uint32_t b = (2 * threadIdx.x + 1 < D) ? g[2 * threadIdx.x + 1] : 0u;
        part[threadIdx.x] = a + b;
        __syncthreads();
        for (int o = 1; o < 256; o <<= 1) {
            const uint32_t v = (threadIdx.x >= o) ? part[threadIdx.x - o] : 0u;
            __syncthreads();
            part[threadIdx.x] += v;
            __syncthreads();
        }
        const uint32_t ex = part[threadIdx.x] - (a + b);
        if (2 * threadIdx.x < D) digit_base[p * 512 + 2 * threadIdx.x] = ex;
        if (2 * threadIdx.x + 1 < D) digit_base[p * 512 + 2 * threadIdx.x + 1] = ex + a;
        __syncthreads();
    }
}

template <int BITS>
__global__ void __launch_bounds__(MPMK_SORT_CTA) k_sort_onesweep(const uint32_t* __restrict__ keys_in,
                                                                 const uint32_t* __restrict__ vals_in,
                                                                 uint32_t* __restrict__ keys_out,
                                                                 uint32_t* __restrict__ vals_out, int64_t n, int shift,
                                                                 const uint32_t* __restrict__ digit_base,
                                                                 uint32_t* __restrict__ status, uint32_t* __restrict__ tile_counter) {
    constexpr int D = 1 << BITS;
    constexpr int NW = MPMK_SORT_CTA / 32;
    constexpr int DPT = (D + MPMK_SORT_CTA - 1) / MPMK_SORT_CTA;  // digits per thread (1 or 2)
    __shared__ uint32_t wh[NW][D];
    __shared__ uint32_t lstart[D];   // tile-local first position of digit d
    __shared__ uint32_t gdelta[D];   // global position = local position + gdelta[d] (mod 2^32)
    extern __shared__ uint32_t os_stage[];   // [2][MPMK_SORT_CHUNK]: the tile's keys and values in sorted order
    uint32_t* skey = os_stage;
    uint32_t* sval = os_stage + MPMK_SORT_CHUNK;
    __shared__ uint32_t part[MPMK_SORT_CTA];
    __shared__ uint32_t tile_s;
    if (threadIdx.x == 0) tile_s = atomicAdd(tile_counter, 1u);
    for (int i = threadIdx.x; i < NW * D; i += MPMK_SORT_CTA) (&wh[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = tile_s;
    const int w = threadIdx.x >> 5;
    const unsigned lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const int64_t tile_base = (int64_t)tile * MPMK_SORT_CHUNK;
    const int tile_count = (int)min((int64_t)MPMK_SORT_CHUNK, n - tile_base);
    uint32_t off[MPMK_SORT_ROUNDS];
    // the tile's keys and values arrive in the staging area by asynchronous 16-byte copies (no registers held
    // while they are in flight; the input tile is contiguous)
    {
        const bool full = (tile_count == MPMK_SORT_CHUNK) && ((tile_base & 3) == 0);
        if (full) {
#pragma unroll
            for (int q = 0; q < MPMK_SORT_CHUNK / (4 * MPMK_SORT_CTA); ++q) {
                const int i4 = q * MPMK_SORT_CTA + threadIdx.x;  // index of a group of 4 words
                cp_async16(skey + 4 * i4, keys_in + tile_base + 4 * i4);
                if (vals_in) cp_async16(sval + 4 * i4, vals_in + tile_base + 4 * i4);
            }
            cp_async_commit();
        } else {
            for (int i = threadIdx.x; i < MPMK_SORT_CHUNK; i += MPMK_SORT_CTA) {
                skey[i] = (i < tile_count) ? keys_in[tile_base + i] : 0xffffffffu;
                if (vals_in) sval[i] = (i < tile_count) ? vals_in[tile_base + i] : 0u;
            }
        }
        if (!vals_in)
            for (int i = threadIdx.x; i < MPMK_SORT_CHUNK; i += MPMK_SORT_CTA) sval[i] = (uint32_t)(tile_base + i);
        cp_async_wait<0>();
        __syncthreads();
    }
    const int seg = w * (32 * MPMK_SORT_ROUNDS);
    // rank of every key among the same-digit keys of its warp's segment (segment order = key order)
#pragma unroll
    for (int r = 0; r < MPMK_SORT_ROUNDS; ++r) {
        const int li = seg + r * 32 + (int)lane;
        const bool valid = li < tile_count;
        const uint32_t d = valid ? ((skey[li] >> shift) & (D - 1)) : (uint32_t)D;
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        const uint32_t rank = __popc(peers & lt);
        const uint32_t prev = valid ? wh[w][d] : 0u;
        __syncwarp();
        if (valid && rank == 0) wh[w][d] = prev + __popc(peers);
        __syncwarp();
        off[r] = prev + rank;
    }
    __syncthreads();
    // per digit: warp counts -> warp-exclusive offsets, tile total published for the tiles behind us
    uint32_t total[DPT];
#pragma unroll
    for (int q = 0; q < DPT; ++q) {
        const int d = threadIdx.x + q * MPMK_SORT_CTA;
        total[q] = 0;
        if (d < D) {
            uint32_t run = 0;
#pragma unroll
            for (int ww = 0; ww < NW; ++ww) { const uint32_t t = wh[ww][d]; wh[ww][d] = run; run += t; }
            total[q] = run;
            atomicExch(&status[(size_t)tile * D + d], run | (tile == 0 ? OS_FLAG_INC : OS_FLAG_AGG));
        }
    }
    // tile-local exclusive scan over the digits
    {
        uint32_t s = 0;
#pragma unroll
        for (int q = 0; q < DPT; ++q) s += total[q];
        // thread t holds digits t and t+CTA: scan in digit order needs two sweeps when DPT == 2
        part[threadIdx.x] = total[0];
        __syncthreads();
        for (int o = 1; o < MPMK_SORT_CTA; o <<= 1) {
            const uint32_t v = (threadIdx.x >= o) ? part[threadIdx.x - o] : 0u;
            __syncthreads();
            part[threadIdx.x] += v;
            __syncthreads();
        }
        const uint32_t lo_total = part[MPMK_SORT_CTA - 1];  // count of digits [0, CTA)
        if (threadIdx.x < D) lstart[threadIdx.x] = part[threadIdx.x] - total[0];
        __syncthreads();
        if (DPT == 2) {
            part[threadIdx.x] = total[DPT - 1];
            __syncthreads();
            for (int o = 1; o < MPMK_SORT_CTA; o <<= 1) {
                const uint32_t v = (threadIdx.x >= o) ? part[threadIdx.x - o] : 0u;
                __syncthreads();
                part[threadIdx.x] += v;
                __syncthreads();
            }
            if (threadIdx.x + MPMK_SORT_CTA < D) lstart[threadIdx.x + MPMK_SORT_CTA] = lo_total + part[threadIdx.x] - total[DPT - 1];
            __syncthreads();
        }
        (void)s;
    }
    // decoupled look-back: count of digit d in all earlier tiles
#pragma unroll
    for (int q = 0; q < DPT; ++q) {
        const int d = threadIdx.x + q * MPMK_SORT_CTA;
        if (d < D) {
            uint32_t excl = 0;
            if (tile > 0) {
                int64_t t = (int64_t)tile - 1;
                while (true) {
                    const uint32_t v = *((volatile const uint32_t*)&status[(size_t)t * D + d]);
                    if ((v & (OS_FLAG_AGG | OS_FLAG_INC)) == 0u) continue;  // that tile is still counting
                    excl += v & OS_MASK;
                    if (v & OS_FLAG_INC) break;
                    --t;
                }
                atomicExch(&status[(size_t)tile * D + d], (excl + total[q]) | OS_FLAG_INC);
            }
            gdelta[d] = digit_base[d] + excl - lstart[d];
        }
    }
    __syncthreads();
    // re-stage the tile in sorted order (through registers: the staging area is rewritten in place)
    {
        uint32_t key[MPMK_SORT_ROUNDS], val[MPMK_SORT_ROUNDS];
#pragma unroll
        for (int r = 0; r < MPMK_SORT_ROUNDS; ++r) {
            const int li = seg + r * 32 + (int)lane;
            key[r] = skey[li];
            val[r] = sval[li];
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < MPMK_SORT_ROUNDS; ++r) {
            if (seg + r * 32 + (int)lane < tile_count) {
                const uint32_t d = (key[r] >> shift) & (D - 1);
                const uint32_t lp = lstart[d] + wh[w][d] + off[r];
                skey[lp] = key[r];
                sval[lp] = val[r];
            }
        }
    }
    __syncthreads();
    // write out: consecutive threads hold consecutive positions of the same digit run
#pragma unroll
    for (int r = 0; r < MPMK_SORT_ROUNDS; ++r) {
        const int i = r * MPMK_SORT_CTA + threadIdx.x;
        if (i < tile_count) {
            const uint32_t k = skey[i];
            const uint32_t pos = (uint32_t)i + gdelta[(k >> shift) & (D - 1)];
            keys_out[pos] = k;
            vals_out[pos] = sval[i];
        }
    }
}

template <int BITS>
static int sort_all(mpmk_ctx* ctx, int passes, cudaStream_t s) {
    constexpr int D = 1 << BITS;
    const int ntiles = div_up(ctx->n, MPMK_SORT_CHUNK);
    const size_t ctrl_words = OS_CTRL_STATUS + (size_t)passes * ntiles * D;
    MPMK_CHECK(ctx, cudaMemsetAsync(ctx->sort_ctrl, 0, ctrl_words * 4, s));
    int hgrid = ntiles < ctx->sm_count * 4 ? ntiles : ctx->sm_count * 4;
    k_sort_digit_hist<BITS><<<hgrid, 256, 0, s>>>((const uint32_t*)ctx->b.keys, ctx->n, passes, ctx->sort_ctrl, ctx->hist_tot);
    MPMK_LAUNCH_CHECK(ctx);
    uint32_t* kb[2] = {ctx->sorted_keys, ctx->keys_alt};
    uint32_t* vb[2] = {ctx->perm, ctx->perm_alt};
    const uint32_t* kin = (const uint32_t*)ctx->b.keys;
    const uint32_t* vin = nullptr;
    constexpr int stage_bytes = 2 * MPMK_SORT_CHUNK * 4;
    static bool attr_set = false;  // static + dynamic shared memory exceeds 48 KB for 9-bit digits: opt in once
    if (!attr_set) {
        MPMK_CHECK(ctx, cudaFuncSetAttribute(k_sort_onesweep<BITS>, cudaFuncAttributeMaxDynamicSharedMemorySize, stage_bytes));
        attr_set = true;
    }
    for (int p = 0; p < passes; ++p) {
        const int o = (passes - 1 - p) & 1;
        k_sort_onesweep<BITS><<<ntiles, MPMK_SORT_CTA, stage_bytes, s>>>(kin, vin, kb[o], vb[o], ctx->n, p * BITS, ctx->hist_tot + p * 512,
                                                               ctx->sort_ctrl + OS_CTRL_STATUS + (size_t)p * ntiles * D,
                                                               ctx->sort_ctrl + p);
        MPMK_LAUNCH_CHECK(ctx);
        kin = kb[o];
        vin = vb[o];
    }
    return 0;
}

int mpmk_launch_sort(mpmk_ctx* ctx, cudaStream_t s) {
    if (ctx->n == 0) { ctx->sorted_valid = true; return 0; }
    int passes = (ctx->key_bits + MPMK_MAX_RADIX_BITS - 1) / MPMK_MAX_RADIX_BITS;
    int bits = (ctx->key_bits + passes - 1) / passes;
    if (bits < 5) bits = 5;
    int rc;
    switch (bits) {
        case 5: rc = sort_all<5>(ctx, passes, s); break;
        case 6: rc = sort_all<6>(ctx, passes, s); break;
        case 7: rc = sort_all<7>(ctx, passes, s); break;
        case 8: rc = sort_all<8>(ctx, passes, s); break;
        default: rc = sort_all<9>(ctx, passes, s); break;
    }
    if (rc) return rc;
    ctx->sorted_valid = true;
    ctx->prepared = false;
    return 0;
}
