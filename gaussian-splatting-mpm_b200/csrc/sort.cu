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
    ctx->lag_ready = false;
    ctx->sorted_valid = false;
    ctx->coherent = false;   // the host rewrote positions: the old sorted keys say nothing about the new ones
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
// [5],[6] chunk tickets of the two coherent-sort scans, [7] mover count, [8 ..] global digit histograms
// [OS_MAX_PASSES][512], then the status words [passes][ntiles][D], then [2][ntiles] scan status words
#define OS_CTRL_HIST 8
#define OS_CTRL_STATUS (OS_CTRL_HIST + OS_MAX_PASSES * 512)

template <int BITS>
__global__ void __launch_bounds__(256) k_sort_digit_hist(const uint32_t* __restrict__ keys, int64_t n, const uint32_t* n_dev,
                                                         int passes, uint32_t* __restrict__ ctrl) {
    pdl_wait();   // programmatic dependent launch: nothing of the previous grid is read or overwritten before this
    constexpr int D = 1 << BITS;
    if (n_dev) n = (int64_t)*n_dev;   // coherent sort: the mover count is only known on the device
    __shared__ uint32_t h[OS_MAX_PASSES][D];
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
}

// inclusive scan of two values across the CTA: shuffles inside the warps, one hop through shared memory for the warp
// totals (three barriers; the log-step shared-memory scan this replaces cost two per step)
template <int CTA>
__device__ __forceinline__ void os_block_scan2(uint32_t& a, uint32_t& b, uint32_t& tot_a, uint32_t& tot_b, uint32_t* wsum) {
    constexpr int NW = CTA / 32;
    const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t ta = __shfl_up_sync(0xffffffffu, a, o), tb = __shfl_up_sync(0xffffffffu, b, o);
        if (lane >= (unsigned)o) { a += ta; b += tb; }
    }
    if (lane == 31) { wsum[w] = a; wsum[32 + w] = b; }
    __syncthreads();
    if (w == 0) {
        uint32_t va = lane < NW ? wsum[lane] : 0u, vb = lane < NW ? wsum[32 + lane] : 0u;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t ta = __shfl_up_sync(0xffffffffu, va, o), tb = __shfl_up_sync(0xffffffffu, vb, o);
            if (lane >= (unsigned)o) { va += ta; vb += tb; }
        }
        wsum[lane] = va; wsum[32 + lane] = vb;
    }
    __syncthreads();
    if (w > 0) { a += wsum[w - 1]; b += wsum[32 + w - 1]; }
    tot_a = wsum[NW - 1]; tot_b = wsum[32 + NW - 1];
    __syncthreads();   // wsum is free again
}

// one batch of the look-back: B status words per unfinished digit, all loads issued before any is looked at
template <int D, int CTA, int DPT, int B>
__device__ __forceinline__ void os_lookback(const uint32_t* status, uint32_t (&excl)[DPT], int64_t (&tb)[DPT], bool (&done)[DPT]) {
    uint32_t v[DPT][B];
#pragma unroll
    for (int q = 0; q < DPT; ++q) {
        const int d = threadIdx.x + q * CTA;
#pragma unroll
        for (int b = 0; b < B; ++b)
            v[q][b] = (!done[q] && tb[q] - b >= 0) ? *((volatile const uint32_t*)&status[(size_t)(tb[q] - b) * D + d]) : OS_FLAG_INC;
    }
#pragma unroll
    for (int q = 0; q < DPT; ++q) {
        if (done[q]) continue;
        int b = 0;
#pragma unroll
        for (; b < B; ++b) {
            if ((v[q][b] & (OS_FLAG_AGG | OS_FLAG_INC)) == 0u) break;   // that tile is still counting: reload from it
            excl[q] += v[q][b] & OS_MASK;
            if (v[q][b] & OS_FLAG_INC) { done[q] = true; break; }
        }
        tb[q] -= b;
    }
}

// CTA threads rank ROUNDS keys each: 256 x 16 for the full sort of a large cloud; 512 x 8 for the coherent sort's few
// movers and for small clouds, where the latency of one tile, not throughput, is what a pass costs (same tile size,
// hence the same status layout; 1024 x 4 measured slower from ~0.6 M keys on: one CTA per SM)
template <int BITS, int CTA, int ROUNDS>
__global__ void __launch_bounds__(CTA) k_sort_onesweep(const uint32_t* __restrict__ keys_in,
                                                                 const uint32_t* __restrict__ vals_in,
                                                                 uint32_t* __restrict__ keys_out,
                                                                 uint32_t* __restrict__ vals_out, int64_t n,
                                                                 const uint32_t* n_dev, int shift,
                                                                 const uint32_t* __restrict__ ghist,
                                                                 uint32_t* __restrict__ status, uint32_t* __restrict__ tile_counter) {
    pdl_wait();   // programmatic dependent launch: nothing of the previous grid is read or overwritten before this
    constexpr int D = 1 << BITS;
    constexpr int CHUNK = CTA * ROUNDS;   // (the host sizes the status words of a pass by the tile count of this CHUNK)
    constexpr int NW = CTA / 32;
    constexpr int DPT = (D + CTA - 1) / CTA;  // digits per thread (1 or 2)
    __shared__ uint32_t lstart[D];   // tile-local first position of digit d
    __shared__ uint32_t gdelta[D];   // global position = local position + gdelta[d] (mod 2^32)
    extern __shared__ uint32_t os_stage[];   // [2][CHUNK]: the tile's keys and values in sorted order, then wh[NW][D]
    uint32_t* skey = os_stage;
    uint32_t* sval = os_stage + CHUNK;
    uint32_t (*wh)[D] = reinterpret_cast<uint32_t (*)[D]>(os_stage + 2 * CHUNK);   // per-warp digit counters
    __shared__ uint32_t wsum[64];     // warp totals of the two block scans
    __shared__ uint32_t tile_s;
    uint32_t hcount[DPT], hbase[DPT];
#pragma unroll
    for (int q = 0; q < DPT; ++q) { const int d = threadIdx.x + q * CTA; hcount[q] = d < D ? ghist[d] : 0u; hbase[q] = 0u; }
    if (n_dev) n = (int64_t)*n_dev;   // coherent sort: the mover count is only known on the device
    // tiles are handed out by ticket (so every tile a CTA waits for in the look-back is running or done); a CTA
    // keeps taking tiles until none are left, which lets the coherent sort launch a small fixed grid
    for (;;) {
    __syncthreads();   // the previous tile's write-out is done with the staging area and tile_s
    if (threadIdx.x == 0) tile_s = atomicAdd(tile_counter, 1u);
    __syncthreads();
    const uint32_t tile = tile_s;
    if ((int64_t)tile * CHUNK >= n) return;
    for (int i = threadIdx.x; i < NW * D; i += CTA) (&wh[0][0])[i] = 0;
    __syncthreads();
    const int w = threadIdx.x >> 5;
    const unsigned lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const int64_t tile_base = (int64_t)tile * CHUNK;
    const int tile_count = (int)min((int64_t)CHUNK, n - tile_base);
    uint32_t off[ROUNDS];
    // the tile's keys and values arrive in the staging area by asynchronous 16-byte copies (no registers held
    // while they are in flight; the input tile is contiguous)
    {
        const bool full = (tile_count == CHUNK) && ((tile_base & 3) == 0);
        if (full) {
#pragma unroll
            for (int q = 0; q < CHUNK / (4 * CTA); ++q) {
                const int i4 = q * CTA + threadIdx.x;  // index of a group of 4 words
                cp_async16(skey + 4 * i4, keys_in + tile_base + 4 * i4);
                if (vals_in) cp_async16(sval + 4 * i4, vals_in + tile_base + 4 * i4);
            }
            cp_async_commit();
        } else {
            for (int i = threadIdx.x; i < CHUNK; i += CTA) {
                skey[i] = (i < tile_count) ? keys_in[tile_base + i] : 0xffffffffu;
                if (vals_in) sval[i] = (i < tile_count) ? vals_in[tile_base + i] : 0u;
            }
        }
        if (!vals_in)
            for (int i = threadIdx.x; i < CHUNK; i += CTA) sval[i] = (uint32_t)(tile_base + i);
        cp_async_wait<0>();
        __syncthreads();
    }
    const int seg = w * (32 * ROUNDS);
    // rank of every key among the same-digit keys of its warp's segment (segment order = key order)
#pragma unroll
    for (int r = 0; r < ROUNDS; ++r) {
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
        const int d = threadIdx.x + q * CTA;
        total[q] = 0;
        if (d < D) {
            uint32_t run = 0;
#pragma unroll
            for (int ww = 0; ww < NW; ++ww) { const uint32_t t = wh[ww][d]; wh[ww][d] = run; run += t; }
            total[q] = run;
            atomicExch(&status[(size_t)tile * D + d], run | (tile == 0 ? OS_FLAG_INC : OS_FLAG_AGG));
        }
    }
    // tile-local exclusive scan over the digits (thread t holds digits t and t+CTA: two sweeps when DPT == 2); the
    // global digit histogram rides along in the same scans -> exclusive digit bases
    {
        uint32_t ia = total[0], ih = hcount[0], lo_total, lo_htotal;
        os_block_scan2<CTA>(ia, ih, lo_total, lo_htotal, wsum);
        if (threadIdx.x < D) lstart[threadIdx.x] = ia - total[0];
        hbase[0] = ih - hcount[0];
        if (DPT == 2) {
            uint32_t hi_total, hi_htotal;
            ia = total[DPT - 1]; ih = hcount[DPT - 1];
            os_block_scan2<CTA>(ia, ih, hi_total, hi_htotal, wsum);
            if (threadIdx.x + CTA < D) lstart[threadIdx.x + CTA] = lo_total + ia - total[DPT - 1];
            hbase[DPT - 1] = lo_htotal + ih - hcount[DPT - 1];
        }
    }
    // decoupled look-back: count of digit d in all earlier tiles.  The first batch reads the eight tiles behind us
    // (independent loads); when none of them has its inclusive count yet -- all tiles started together, the
    // coherent sort's few mover tiles or a small scene -- the walk goes on 32 tiles at a time, and a thread's two
    // digits walk side by side, so the chain is a couple of L2 round trips instead of one per eight tiles per digit
    {
        uint32_t excl[DPT];
        int64_t tb[DPT];
        bool done[DPT];
#pragma unroll
        for (int q = 0; q < DPT; ++q) { excl[q] = 0u; tb[q] = (int64_t)tile - 1; done[q] = (threadIdx.x + q * CTA >= D) || tile == 0; }
        os_lookback<D, CTA, DPT, 8>(status, excl, tb, done);
        for (;;) {
            bool all = true;
#pragma unroll
            for (int q = 0; q < DPT; ++q) all = all && done[q];
            if (all) break;
            os_lookback<D, CTA, DPT, 32>(status, excl, tb, done);
        }
#pragma unroll
        for (int q = 0; q < DPT; ++q) {
            const int d = threadIdx.x + q * CTA;
            if (d < D) {
                if (tile > 0) atomicExch(&status[(size_t)tile * D + d], (excl[q] + total[q]) | OS_FLAG_INC);
                gdelta[d] = hbase[q] + excl[q] - lstart[d];
            }
        }
    }
    __syncthreads();
    // re-stage the tile in sorted order (through registers: the staging area is rewritten in place)
    {
        uint32_t key[ROUNDS], val[ROUNDS];
#pragma unroll
        for (int r = 0; r < ROUNDS; ++r) {
            const int li = seg + r * 32 + (int)lane;
            key[r] = skey[li];
            val[r] = sval[li];
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < ROUNDS; ++r) {
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
    for (int r = 0; r < ROUNDS; ++r) {
        const int i = r * CTA + threadIdx.x;
        if (i < tile_count) {
            const uint32_t k = skey[i];
            const uint32_t pos = (uint32_t)i + gdelta[(k >> shift) & (D - 1)];
            keys_out[pos] = k;
            vals_out[pos] = sval[i];
        }
    }
    }   // next tile
}


// ---- coherent sort ----------------------------------------------------------------------
// g2p writes the particles in sorted order, so in the next substep the OLD key of physical slot r is still
// sitting in sorted_keys[r], and all but the few particles that changed cell ("movers", keys[r] != okeys[r]) are
// already in order among themselves.  The stable sort by (key, slot) is then the merge of
//     stayers -- sorted as they lie -- and movers -- compacted in slot order, then radix-sorted by key (stable),
// and the merge needs no comparisons between the two lists beyond searches in the old sorted keys:
//   mover (k, j): every stayer in a slot < x sorts before it, x = min(max(j, lower_bound(okeys, k)), upper_bound(okeys, k))
//                 (stayers keep their old key, and okeys is non-decreasing), so with t = its rank among the sorted
//                 movers and S(x) = stayers in slots [0, x):      position = t + S(x);
//   stayer i:     the movers in front of it are those with x <= i (x is non-decreasing in t):
//                                                                 position = S(i) + #{t : x_t <= i}.
// S comes from a chained scan of the mover flags (k_cs_flag_scan), #{t : x_t <= i} from a second chained scan over
// per-word insert counts (k_cs_place_stayers).  The result is the permutation a CPU stable sort produces, for ANY
// non-decreasing okeys; nearly-unchanged keys only make it cheap (every pass touches N keys once instead of
// three scatter passes).  Correctness does not depend on the mover fraction, only the speed does.
#define CS_CHUNK 4096   // slots per CTA of the scans = 128 mask words, 16 per warp
#define CS_THREADS 256
#define CS_CTRL_TICKET_A 5
#define CS_CTRL_TICKET_B 6
#define CS_CTRL_MOVERS 7

// Warp-parallel decoupled look-back: sum of the aggregates of chunks [0, chunk).  Chunks are handed out by ticket,
// so every predecessor has at least started.
__device__ __forceinline__ uint32_t cs_lookback(const uint32_t* status, uint32_t chunk, unsigned lane) {
    // Window of 128 predecessors per step (lane l reads chunk-1-l, -33-l, -65-l, -97-l): with every chunk resident at
    // once the walk and the front of inclusive prefixes approach each other 2 x 128 chunks per L2 round trip.
    uint32_t excl = 0;
    int64_t t = (int64_t)chunk - 1;
    while (t >= 0) {
        uint32_t v[4];
        bool ready;
        do {
            ready = true;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const int64_t mine = t - 32 * g - (int64_t)lane;
                v[g] = (mine >= 0) ? *((volatile const uint32_t*)&status[mine]) : OS_FLAG_INC;   // before chunk 0: prefix 0
                ready = ready && ((v[g] & (OS_FLAG_AGG | OS_FLAG_INC)) != 0u);
            }
        } while (!__all_sync(0xffffffffu, ready));
        bool found = false;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if (found) break;
            const unsigned inc = __ballot_sync(0xffffffffu, (v[g] & OS_FLAG_INC) != 0u);
            const int first = __ffs(inc) - 1;   // nearest predecessor of this group holding an inclusive prefix
            const uint32_t c = (inc == 0u || (int)lane <= first) ? (v[g] & OS_MASK) : 0u;
            excl += __reduce_add_sync(0xffffffffu, c);
            found = inc != 0u;
        }
        if (found) break;
        t -= 128;
    }
    return excl;
}

// CTA-wide exclusive offsets of the per-word counts `pc` (held by lanes 0..15 of each warp), chained across chunks.
// Returns, for lanes 0..15, the global exclusive prefix of that lane's word; *total_out = inclusive prefix of the chunk.
__device__ __forceinline__ uint32_t cs_chunk_scan(uint32_t pc, uint32_t chunk, uint32_t* status, uint32_t* wsum, uint32_t* excl_s,
                                                  uint32_t* total_out) {
    const int w = threadIdx.x >> 5;
    const unsigned lane = threadIdx.x & 31;
    uint32_t inc = pc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
        if ((int)lane >= o) inc += v;
    }
    if (lane == 31) wsum[w] = inc;
    __syncthreads();
    uint32_t wex = 0, total = 0;
#pragma unroll
    for (int ww = 0; ww < CS_THREADS / 32; ++ww) { const uint32_t t = wsum[ww]; if (ww < w) wex += t; total += t; }
    if (w == 0) {
        if (lane == 0) atomicExch(&status[chunk], total | (chunk == 0 ? OS_FLAG_INC : OS_FLAG_AGG));
        const uint32_t ex = chunk ? cs_lookback(status, chunk, lane) : 0u;
        if (lane == 0) {
            if (chunk) atomicExch(&status[chunk], (ex + total) | OS_FLAG_INC);
            *excl_s = ex;
        }
    }
    __syncthreads();
    *total_out = *excl_s + total;
    return *excl_s + wex + inc - pc;
}

__global__ void __launch_bounds__(CS_THREADS) k_cs_flag_scan(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ okeys,
                                                             int64_t n, uint32_t* __restrict__ ctrl, uint32_t* __restrict__ status,
                                                             uint32_t* __restrict__ mask, uint32_t* __restrict__ mprefix,
                                                             uint32_t* __restrict__ cnt, uint32_t* __restrict__ mk,
                                                             uint32_t* __restrict__ mv, int bits, int passes,
                                                             uint32_t* __restrict__ samples) {
    pdl_wait();   // programmatic dependent launch: nothing of the previous grid is read or overwritten before this
    __shared__ uint32_t chunk_s, excl_s;
    __shared__ uint32_t wsum[CS_THREADS / 32];
    __shared__ uint32_t hist_s[OS_MAX_PASSES * 512];
    // one resident wave of CTAs, each taking 4096-slot chunks by ticket until none are left: a chunk's predecessors
    // are then mostly finished when it looks back, and there is no second wave waiting for the first to drain
    for (;;) {
    __syncthreads();
    if (threadIdx.x == 0) chunk_s = atomicAdd(&ctrl[CS_CTRL_TICKET_A], 1u);
    __syncthreads();
    const uint32_t chunk = chunk_s;
    if ((int64_t)chunk * CS_CHUNK >= n) return;
    const int w = threadIdx.x >> 5;
    const unsigned lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const int64_t wbase = (int64_t)chunk * CS_CHUNK + (int64_t)w * 512;
    uint32_t k[16];
    uint32_t myword = 0;   // lane r < 16 keeps the mover mask of the warp's r-th word
    {   // 1024 evenly spaced old keys, the first level of k_cs_place_movers' searches: this chunk's share
        const uint32_t sstride = ((uint32_t)n + 1023u) / 1024u;
        const int64_t c0 = (int64_t)chunk * CS_CHUNK, c1 = min(c0 + CS_CHUNK, n);
        for (int64_t si = (c0 + sstride - 1) / sstride + threadIdx.x; si * sstride < c1; si += CS_THREADS) samples[si] = okeys[si * sstride];
    }
    uint32_t ok[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) {   // all 32 loads in flight before the first vote
        const int64_t slot = wbase + r * 32 + lane;
        k[r] = 0; ok[r] = 0;
        if (slot < n) { k[r] = __ldg(keys + slot); ok[r] = __ldg(okeys + slot); }
    }
#pragma unroll
    for (int r = 0; r < 16; ++r) {
        const unsigned m = __ballot_sync(0xffffffffu, k[r] != ok[r]);
        if ((int)lane == r) myword = m;
    }
    const uint32_t pc = __popc(myword);
    for (int i = threadIdx.x; i < passes * 512; i += CS_THREADS) hist_s[i] = 0u;
    uint32_t total;
    const uint32_t wordex = cs_chunk_scan(pc, chunk, status, wsum, &excl_s, &total);   // (its barriers also publish hist_s = 0)
    const uint32_t total_movers_cta = total - excl_s;
    if (threadIdx.x == 0 && (int64_t)(chunk + 1) * CS_CHUNK >= n) ctrl[CS_CTRL_MOVERS] = total;
    const int64_t word0 = wbase >> 5;
    if (lane < 16 && (word0 + lane) * 32 < n) {
        mask[word0 + lane] = myword;
        mprefix[word0 + lane] = wordex;
        cnt[word0 + lane] = 0u;
    }
#pragma unroll
    for (int r = 0; r < 16; ++r) {
        const unsigned m = __shfl_sync(0xffffffffu, myword, r);
        const uint32_t ex = __shfl_sync(0xffffffffu, wordex, r);
        if ((m >> lane) & 1u) {
            const uint32_t c = ex + __popc(m & lt);
            mk[c] = k[r];
            mv[c] = (uint32_t)(wbase + r * 32 + lane);
            // digit histograms of the movers for every radix pass, aggregated per CTA (a chunk's movers share
            // their high digits)
            for (int p = 0; p < passes; ++p) atomicAdd(&hist_s[p * 512 + ((k[r] >> (p * bits)) & ((1u << bits) - 1u))], 1u);
        }
    }
    __syncthreads();
    if (total_movers_cta) {
        for (int i = threadIdx.x; i < passes * 512; i += CS_THREADS) {
            const uint32_t v = hist_s[i];
            if (v) atomicAdd(&ctrl[OS_CTRL_HIST + i], v);
        }
    }
    }   // next chunk
}

__global__ void __launch_bounds__(256) k_cs_place_movers(const uint32_t* __restrict__ mk, const uint32_t* __restrict__ mv,
                                                         const uint32_t* __restrict__ ctrl, const uint32_t* __restrict__ okeys,
                                                         int64_t n, const uint32_t* __restrict__ mask,
                                                         const uint32_t* __restrict__ mprefix, uint32_t* __restrict__ cnt,
                                                         uint32_t* __restrict__ mx, uint32_t* __restrict__ out_keys,
                                                         uint32_t* __restrict__ out_perm, const uint32_t* __restrict__ samples,
                                                         uint32_t* __restrict__ rank) {
    pdl_wait();   // programmatic dependent launch: nothing of the previous grid is read or overwritten before this
    const uint32_t M = ctrl[CS_CTRL_MOVERS];
    const uint32_t N = (uint32_t)n;
    // first level of the searches: 1024 evenly spaced old keys in shared memory
    __shared__ uint32_t samp[1024];
    const uint32_t stride = (N + 1023u) / 1024u;
    for (uint32_t i = threadIdx.x; i < 1024u; i += blockDim.x) {
        const uint64_t idx = (uint64_t)i * stride;
        samp[i] = idx < N ? samples[i] : 0xffffffffu;
    }
    __syncthreads();
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < M; t += gridDim.x * blockDim.x) {
        const uint32_t kj = mk[t], j = mv[t];
        // A mover's new key differs from its old one.  Moved up (old < new): every slot up to j holds a smaller old
        // key, so x = lower_bound(okeys, kj) > j.  Moved down: x = upper_bound(okeys, kj) <= j.  One search either way.
        const bool up = okeys[j] < kj;
        uint32_t lo = 0, hi = 1024;
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            const uint32_t v = samp[mid];
            if (up ? (v < kj) : (v <= kj)) lo = mid + 1; else hi = mid;
        }
        // the answer lies in ((lo-1)*stride, lo*stride]
        const uint64_t whi = (uint64_t)lo * stride;
        hi = whi < N ? (uint32_t)whi : N;
        lo = lo ? (lo - 1u) * stride + 1u : 0u;
        if (up) lo = max(lo, j + 1u); else hi = min(hi, j);
        // 4-ary search: three independent probes per step halve the chain of dependent loads
        while (hi - lo > 2u) {
            const uint32_t q = (hi - lo) >> 2;
            const uint32_t m1 = lo + q, m2 = lo + 2u * q, m3 = lo + 3u * q;
            const uint32_t v1 = okeys[m1], v2 = okeys[m2], v3 = okeys[m3];
            const bool b1 = up ? (v1 < kj) : (v1 <= kj), b2 = up ? (v2 < kj) : (v2 <= kj), b3 = up ? (v3 < kj) : (v3 <= kj);
            // the predicate is monotone (true ... true false ... false): the answer is the first false position
            if (!b1) hi = m1;
            else if (!b2) { lo = m1 + 1u; hi = m2; }
            else if (!b3) { lo = m2 + 1u; hi = m3; }
            else lo = m3 + 1u;
        }
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            const uint32_t v = okeys[mid];
            if (up ? (v < kj) : (v <= kj)) lo = mid + 1; else hi = mid;
        }
        const uint32_t x = lo;
        const uint32_t movers_before = (x >= N) ? M : mprefix[x >> 5] + __popc(mask[x >> 5] & ((1u << (x & 31u)) - 1u));
        const uint32_t pos = t + (x - movers_before);
        out_keys[pos] = kj;
        out_perm[pos] = j;
        if (rank) rank[j] = pos;   // lagged batches: g2p writes slot j's particle to sorted position pos
        mx[t] = x;
        if (x < N) atomicAdd(&cnt[x >> 5], 1u);
    }
}

__global__ void __launch_bounds__(CS_THREADS) k_cs_place_stayers(const uint32_t* __restrict__ keys, int64_t n,
                                                                 uint32_t* __restrict__ ctrl, uint32_t* __restrict__ status,
                                                                 const uint32_t* __restrict__ mask,
                                                                 const uint32_t* __restrict__ mprefix,
                                                                 const uint32_t* __restrict__ cnt, const uint32_t* __restrict__ mx,
                                                                 uint32_t* __restrict__ out_keys, uint32_t* __restrict__ out_perm,
                                                                 uint32_t* __restrict__ rank) {
    pdl_wait();   // programmatic dependent launch: nothing of the previous grid is read or overwritten before this
    __shared__ uint32_t chunk_s, excl_s;
    __shared__ uint32_t wsum[CS_THREADS / 32];
    for (;;) {   // persistent over chunks, as k_cs_flag_scan
    __syncthreads();
    if (threadIdx.x == 0) chunk_s = atomicAdd(&ctrl[CS_CTRL_TICKET_B], 1u);
    __syncthreads();
    const uint32_t chunk = chunk_s;
    if ((int64_t)chunk * CS_CHUNK >= n) return;
    const int w = threadIdx.x >> 5;
    const unsigned lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const int64_t wbase = (int64_t)chunk * CS_CHUNK + (int64_t)w * 512;
    const int64_t word0 = wbase >> 5;
    // issue the key loads before the chained scan so that they overlap the look-back
    uint32_t k[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) {
        const int64_t slot = wbase + r * 32 + lane;
        k[r] = (slot < n) ? keys[slot] : 0u;
    }
    uint32_t pc = 0, myword = 0, mypre = 0;
    if (lane < 16 && (word0 + lane) * 32 < n) { pc = cnt[word0 + lane]; myword = mask[word0 + lane]; mypre = mprefix[word0 + lane]; }
    uint32_t total;
    const uint32_t cw = cs_chunk_scan(pc, chunk, status, wsum, &excl_s, &total);   // movers in front of the word's first slot
#pragma unroll
    for (int r = 0; r < 16; ++r) {
        const int64_t slot = wbase + r * 32 + lane;
        const unsigned m = __shfl_sync(0xffffffffu, myword, r);
        const uint32_t c0 = __shfl_sync(0xffffffffu, cw, r);
        const uint32_t nc = __shfl_sync(0xffffffffu, pc, r);
        const uint32_t mp = __shfl_sync(0xffffffffu, mypre, r);
        if (slot < n && !((m >> lane) & 1u)) {
            uint32_t c = c0;
            for (uint32_t e = 0; e < nc; ++e) {   // the movers inserted inside this word (x sorted: stop at the first one beyond)
                if (mx[c0 + e] <= (uint32_t)slot) ++c; else break;
            }
            const uint32_t pos = ((uint32_t)slot - (mp + __popc(m & lt))) + c;
            out_keys[pos] = k[r];
            out_perm[pos] = (uint32_t)slot;
            if (rank) rank[slot] = pos;
        }
    }
    }   // next chunk
}

// inverse of the permutation (full radix path of a lagged batch; the merge path writes the ranks as it places)
__global__ void __launch_bounds__(256) k_invert_perm(const uint32_t* __restrict__ perm, uint32_t* __restrict__ rank, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) rank[perm[i]] = (uint32_t)i;
}

template <int BITS>
static int sort_coherent(mpmk_ctx* ctx, int passes, cudaStream_t s, bool want_rank) {
    constexpr int D = 1 << BITS;
    const int ntiles = div_up(ctx->n, MPMK_SORT_CHUNK);   // == chunks of the scans (CS_CHUNK == MPMK_SORT_CHUNK)
    static_assert(CS_CHUNK == MPMK_SORT_CHUNK, "the scans and the radix tiles share the control block sizing");
    const int mtiles = want_rank ? div_up(ctx->n, 2048) : ntiles;   // tiles of the movers' radix passes (lean variant: 2048 keys)
    const size_t radix_status = (size_t)passes * mtiles * D;
    const size_t ctrl_words = OS_CTRL_STATUS + radix_status + 2 * (size_t)ntiles;
    MPMK_CHECK(ctx, cudaMemsetAsync(ctx->sort_ctrl, 0, ctrl_words * 4, s));
    uint32_t* ctrl = ctx->sort_ctrl;
    uint32_t* status_a = ctrl + OS_CTRL_STATUS + radix_status;
    uint32_t* status_b = status_a + ntiles;
    const uint32_t* okeys = ctx->sorted_keys;
    const uint32_t* keys = ctx->keys_cur;
    uint32_t* rank = want_rank ? ctx->rank : nullptr;
    const int first = passes & 1;   // the compacted movers start in this buffer so that the last radix pass lands in buffer 0
    mpmk_launch(k_cs_flag_scan, dim3(min(ntiles, ctx->sm_count * 5)), dim3(CS_THREADS), 0, s, keys, okeys, ctx->n, ctrl, status_a, ctx->cs_mask, ctx->cs_mprefix, ctx->cs_cnt,
                                                 ctx->cs_mk[first], ctx->cs_mv[first], BITS, passes, ctx->hist_tot);   // hist_tot: 1024 key samples
    MPMK_LAUNCH_CHECK(ctx);
    if (want_rank) mpmk_trace(ctx, s, 8);
    const uint32_t* m_dev = ctrl + CS_CTRL_MOVERS;
    constexpr int stage_bytes = 2 * MPMK_SORT_CHUNK * 4 + (512 / 32) * D * 4;   // staging + per-warp digit counters
    constexpr int stage_bytes_lean = 2 * 2048 * 4 + (256 / 32) * D * 4;
    int in = first;
    const int mgrid = ntiles < ctx->sm_count ? ntiles : ctx->sm_count;   // persistent: CTAs loop over the movers' tiles
    for (int p = 0; p < passes; ++p) {
        const int o = 1 - in;
        if (want_rank)   // beside the persistent p2g grid: 256 threads x 64 registers, 2048-key tiles (what p2g leaves free on an SM)
            mpmk_launch(k_sort_onesweep<BITS, 256, 8>, dim3(mgrid), dim3(256), stage_bytes_lean, s, ctx->cs_mk[in], ctx->cs_mv[in], ctx->cs_mk[o], ctx->cs_mv[o], 0, m_dev,
                                                                   p * BITS, ctx->sort_ctrl + OS_CTRL_HIST + p * 512,
                                                                   ctrl + OS_CTRL_STATUS + (size_t)p * mtiles * D, ctrl + p);
        else
            mpmk_launch(k_sort_onesweep<BITS, 512, 8>, dim3(mgrid), dim3(512), stage_bytes, s, ctx->cs_mk[in], ctx->cs_mv[in], ctx->cs_mk[o], ctx->cs_mv[o], 0, m_dev,
                                                                   p * BITS, ctx->sort_ctrl + OS_CTRL_HIST + p * 512,
                                                                   ctrl + OS_CTRL_STATUS + (size_t)p * mtiles * D, ctrl + p);
        MPMK_LAUNCH_CHECK(ctx);
        in = o;
    }
    if (want_rank) mpmk_trace(ctx, s, 9);
    // sorted movers are in buffer `in` (== 0); the other key buffer is free and receives x_t
    uint32_t* mx = ctx->cs_mk[1 - in];
    mpmk_launch(k_cs_place_movers, dim3(ctx->sm_count * 4), dim3(256), 0, s, ctx->cs_mk[in], ctx->cs_mv[in], ctrl, okeys, ctx->n, ctx->cs_mask, ctx->cs_mprefix,
                                                        ctx->cs_cnt, mx, ctx->keys_alt, ctx->perm, ctx->hist_tot, rank);
    MPMK_LAUNCH_CHECK(ctx);
    if (want_rank) mpmk_trace(ctx, s, 10);
    mpmk_launch(k_cs_place_stayers, dim3(min(ntiles, ctx->sm_count * 4)), dim3(CS_THREADS), 0, s, keys, ctx->n, ctrl, status_b, ctx->cs_mask, ctx->cs_mprefix, ctx->cs_cnt, mx,
                                                     ctx->keys_alt, ctx->perm, rank);
    MPMK_LAUNCH_CHECK(ctx);
    uint32_t* t = ctx->sorted_keys; ctx->sorted_keys = ctx->keys_alt; ctx->keys_alt = t;
    ctx->last_sort_coherent = true;
    return 0;
}

template <int BITS>
static int sort_all(mpmk_ctx* ctx, int passes, cudaStream_t s, bool want_rank) {
    constexpr int D = 1 << BITS;
    const int ntiles = div_up(ctx->n, MPMK_SORT_CHUNK);
    // below ~1.5 M particles the full sort's three or four kernels are quicker than the merge's six (measured on the
    // config-shaped scenes: 240 k particles 37 vs 47 us, 1 M 57 vs 65 us, 3.6 M 87 vs 68 us); MPMK_SORT_MERGE forces it
    if (ctx->coherent && !(ctx->prm.flags & MPMK_SORT_FULL) && (ctx->n >= MPMK_COHERENT_MIN_N || (ctx->prm.flags & MPMK_SORT_MERGE)))
        return sort_coherent<BITS>(ctx, passes, s, want_rank);
    ctx->last_sort_coherent = false;
    // small clouds re-sort in full every substep, and there a pass costs the latency of ONE tile (a warp's serial
    // instruction stream: ranking rounds, re-staging, write-out), not throughput: twice the threads, half the rounds
    // (quarter-size tiles were tried: the look-back chain over four times the tiles costs more than they save)
    const bool small = ctx->n < MPMK_COHERENT_MIN_N;
    const int otiles = ntiles;
    const size_t ctrl_words = OS_CTRL_STATUS + (size_t)passes * otiles * D;
    MPMK_CHECK(ctx, cudaMemsetAsync(ctx->sort_ctrl, 0, ctrl_words * 4, s));
    int hgrid = ntiles < ctx->sm_count * 4 ? ntiles : ctx->sm_count * 4;
    mpmk_launch(k_sort_digit_hist<BITS>, dim3(hgrid), dim3(256), 0, s, (const uint32_t*)ctx->keys_cur, ctx->n, nullptr, passes, ctx->sort_ctrl);
    MPMK_LAUNCH_CHECK(ctx);
    uint32_t* kb[2] = {ctx->sorted_keys, ctx->keys_alt};
    uint32_t* vb[2] = {ctx->perm, ctx->perm_alt};
    const uint32_t* kin = (const uint32_t*)ctx->keys_cur;
    const uint32_t* vin = nullptr;
    constexpr int stage_bytes = 2 * MPMK_SORT_CHUNK * 4 + (256 / 32) * D * 4;
    constexpr int stage_bytes_small = 2 * MPMK_SORT_CHUNK * 4 + (512 / 32) * D * 4;
    for (int p = 0; p < passes; ++p) {
        const int o = (passes - 1 - p) & 1;
        uint32_t* hist = ctx->sort_ctrl + OS_CTRL_HIST + p * 512;
        uint32_t* status = ctx->sort_ctrl + OS_CTRL_STATUS + (size_t)p * otiles * D;
        if (small)
            mpmk_launch(k_sort_onesweep<BITS, 512, 8>, dim3(otiles), dim3(512), stage_bytes_small, s, 
                kin, vin, kb[o], vb[o], ctx->n, nullptr, p * BITS, hist, status, ctx->sort_ctrl + p);
        else
            mpmk_launch(k_sort_onesweep<BITS, 256, 16>, dim3(otiles), dim3(256), stage_bytes, s, kin, vin, kb[o], vb[o], ctx->n, nullptr, p * BITS, hist, status,
                                                                            ctx->sort_ctrl + p);
        MPMK_LAUNCH_CHECK(ctx);
        kin = kb[o];
        vin = vb[o];
    }
    if (want_rank) {
        k_invert_perm<<<min(div_up(ctx->n, 256), ctx->sm_count * 8), 256, 0, s>>>(ctx->perm, ctx->rank, ctx->n);
        MPMK_LAUNCH_CHECK(ctx);
    }
    return 0;
}

// more than 48 KB of shared memory per CTA: opt in, per context (= per device), for every digit width
template <int BITS>
static int sort_init_bits(mpmk_ctx* ctx) {
    constexpr int D = 1 << BITS;
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_sort_onesweep<BITS, 256, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         2 * MPMK_SORT_CHUNK * 4 + (256 / 32) * D * 4));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_sort_onesweep<BITS, 512, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         2 * MPMK_SORT_CHUNK * 4 + (512 / 32) * D * 4));
    // lagged batches run the sort beside the persistent p2g grid: an SM only hosts kernels of one shared-memory
    // configuration at a time, so everything on the side branch asks for the same (maximal) carve-out as p2g
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_sort_onesweep<BITS, 256, 16>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_sort_onesweep<BITS, 512, 8>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_sort_onesweep<BITS, 256, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 2048 * 4 + (256 / 32) * D * 4));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_sort_onesweep<BITS, 256, 8>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_sort_digit_hist<BITS>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    return 0;
}

int mpmk_sort_init(mpmk_ctx* ctx) {
    int rc;
    if ((rc = sort_init_bits<5>(ctx)) || (rc = sort_init_bits<6>(ctx)) || (rc = sort_init_bits<7>(ctx)) ||
        (rc = sort_init_bits<8>(ctx)) || (rc = sort_init_bits<9>(ctx)))
        return rc;
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_cs_flag_scan, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_cs_place_movers, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_cs_place_stayers, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    MPMK_CHECK(ctx, cudaFuncSetAttribute(k_invert_perm, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    return 0;
}

static int launch_sort_impl(mpmk_ctx* ctx, cudaStream_t s, bool want_rank);
int mpmk_launch_sort(mpmk_ctx* ctx, cudaStream_t s) { return launch_sort_impl(ctx, s, false); }
// lagged batches: the same sort of the current keys (ctx->keys_cur), plus the inverse permutation for g2p's scatter
int mpmk_launch_sort_lagged(mpmk_ctx* ctx, cudaStream_t s) { return launch_sort_impl(ctx, s, true); }

static int launch_sort_impl(mpmk_ctx* ctx, cudaStream_t s, bool want_rank) {
    if (ctx->n == 0) { ctx->sorted_valid = true; return 0; }
    // (a lagged batch sorts beside the persistent p2g grid, which leaves 16 K registers per SM: the 256-thread mover
    // passes fit with digits of up to 8 bits -- 64 registers -- and not with 9 -- two digits per thread, 119 registers;
    // a 512^3 grid's 27-bit keys then take four 7-bit passes instead of three 9-bit ones)
    const int max_bits = want_rank ? 8 : MPMK_MAX_RADIX_BITS;
    int passes = (ctx->key_bits + max_bits - 1) / max_bits;
    int bits = (ctx->key_bits + passes - 1) / passes;
    if (bits < 5) bits = 5;
    int rc;
    switch (bits) {
        case 5: rc = sort_all<5>(ctx, passes, s, want_rank); break;
        case 6: rc = sort_all<6>(ctx, passes, s, want_rank); break;
        case 7: rc = sort_all<7>(ctx, passes, s, want_rank); break;
        case 8: rc = sort_all<8>(ctx, passes, s, want_rank); break;
        default: rc = sort_all<9>(ctx, passes, s, want_rank); break;
    }
    if (rc) return rc;
    ctx->sorted_valid = true;
    ctx->coherent = false;   // sorted_keys now holds the NEW keys; the next g2p re-establishes the relation
    ctx->prepared = false;
    return 0;
}

extern "C" int mpmk_sort_info(mpmk_ctx* ctx, int* coherent_out, int64_t* movers_out, void* stream) {
    if (!ctx || !ctx->bound) return MPMK_ERR_STATE;
    uint32_t m = 0;
    if (ctx->last_sort_coherent) {
        MPMK_CHECK(ctx, cudaMemcpyAsync(&m, ctx->sort_ctrl + CS_CTRL_MOVERS, 4, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
        MPMK_CHECK(ctx, cudaStreamSynchronize((cudaStream_t)stream));
    }
    if (coherent_out) *coherent_out = ctx->last_sort_coherent ? 1 : 0;
    if (movers_out) *movers_out = ctx->last_sort_coherent ? (int64_t)m : ctx->n;
    return 0;
}
