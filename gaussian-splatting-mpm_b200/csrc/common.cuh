// common.cuh -- context, launch helpers and small device utilities shared by the
// libmpmk kernels.  sm_100a only; see include/mpmk.h for the ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/mpmk.h"

#define MPMK_TILE 4                 // cells per tile edge; a tile's stencil covers 6^3 nodes
#define MPMK_TILE_NODES 216         // 6*6*6
#ifndef MPMK_ITEM_MAX
#define MPMK_ITEM_MAX 256           // particles per p2g/g2p work item (tile chunk)
#endif
#define MPMK_ITEM_MIN 64            // smallest work item (two batches: lego-like 0.083 -> 0.081 ms against one)
#ifndef MPMK_SORT_CTA
#define MPMK_SORT_CTA 256           // threads per sort CTA
#endif
#ifndef MPMK_SORT_ROUNDS
#define MPMK_SORT_ROUNDS 16         // keys per thread per sort CTA
#endif
#define MPMK_SORT_CHUNK (MPMK_SORT_CTA * MPMK_SORT_ROUNDS)
#define MPMK_MAX_RADIX_BITS 9
#define MPMK_COHERENT_MIN_N 1500000  // particles from which the merge sort (sort.cu) beats the full radix sort
#define MPMK_GRAPH_CACHE 8          // captured substep batches kept per context

struct DevCounters {
    uint32_t n_items;
    uint32_t n_active_blocks;
    uint32_t n_tiles;
    uint32_t n_clamped;
    uint32_t p2g_cursor;   // dynamic work fetch
    uint32_t g2p_cursor;
    uint32_t n_cross;      // lagged batches: particles whose cell left their work item's tile (p2g scatters them straight to the global grid, g2p gathers from it)
    uint32_t list_stamp;   // lagged batches: stamp value of the block list this work list goes with
};

// multi-GPU slab state (mg.cu)
struct MgState {
    void* comm;   // ncclComm_t
    int rank, world;
    bool enabled;
    int x_lo, x_hi, ghost;
    int y_lo, y_hi;          // row range the exchange / dense passes are restricted to
    int z_lo, z_hi;          // ... and the window along the fastest axis (peer-memory path; mpmk_mg_set_window)
    int lo0, lo1, hi0, hi1;  // shared plane ranges with rank-1 / rank+1 (empty when lo1 == lo0)
    void* recv_lo;
    void* recv_hi;
    // peer-memory halo path (mg.cu): this rank's exchange block (flags + double-buffered receive buffers, cudaMalloc'd
    // and exported over CUDA IPC) and the neighbours' blocks mapped into this process
    bool peer;
    unsigned char* my_block;
    unsigned char* peer_block[2];   // [0] rank-1's block, [1] rank+1's block
    size_t buf_bytes;               // one receive buffer
    size_t msg_bytes;               // one message slot (mpmk_mg_ll_send / _recv: the migration payload), 0 = none
};

struct mpmk_ctx {
    int device;
    int n_grid;        // nodes per edge
    int nt;            // tiles per edge = ceil(n_grid/4)
    int nb;            // node blocks per edge = ceil(n_grid/4)
    float dx, inv_dx;
    int64_t capacity;
    int64_t n;         // live particles
    int cur;           // which set holds the current state
    int key_bits;
    int sm_count;
    int g2p_ctas_per_sm, p2g_ctas_per_sm;  // occupancy-derived persistent grid sizes
    int p2g_lag_ctas_per_sm;
    // fused batches (g2p2g.cu): the next substep's scatter is already in grid_in when the substep starts, so the
    // clear must leave grid_in alone, and the grid pass zeroes grid_in behind itself for the scatter that follows it
    bool keep_grid_in, zero_grid_in;
    mpmk_bindings b;
    bool bound;
    mpmk_params prm;
    // scratch (library-owned)
    uint32_t* keys_alt;      // sort ping-pong
    uint32_t* sorted_keys;   // result of last sort
    uint32_t* perm;          // result of last sort: physical slot of the r-th particle in sorted order
    uint32_t* perm_alt;
    // coherent sort scratch (sort.cu): compacted movers (key, slot) ping-pong; per-32-slot-word mover mask, exclusive
    // mover prefix and count of movers inserted in front of the word's stayers
    uint32_t* cs_mk[2];
    uint32_t* cs_mv[2];
    uint32_t* cs_mask;
    uint32_t* cs_mprefix;
    uint32_t* cs_cnt;
    bool last_sort_coherent; // which path the last mpmk_sort took
    bool coherent;           // sorted_keys[r] is the OLD key of physical slot r of the current set (g2p wrote it in sorted order)
    uint32_t* sort_ctrl;     // one-sweep control block: tile counters, digit histograms, status words (sort.cu)
    uint32_t* hist_tot;      // 2048 words: the 1024 evenly spaced old keys the coherent sort searches first
    int64_t n_sort_ctas_cap;
    // work list, double-buffered: p2g / g2p consume set [wl]; prepare builds into [1-wl] and flips.  In a lagged batch
    // (api.cu) the sort + prepare of substep s run on a side branch while p2g / grid of substep s still consume set [wl].
    uint4* items_buf[2];     // {begin, count, tile, 0}
    DevCounters* counters_buf[2];
    int wl;
    uint4* items;            // == items_buf[wl]
    DevCounters* counters;   // == counters_buf[wl]
    uint2* tile_list;        // non-empty tiles of the current sort: {tile, first sorted index}
    uint32_t* tile_end;      // [nt^3] one past the last sorted index of each non-empty tile
    int64_t items_cap;
    uint32_t* block_stamp[2];// [nb^3] per block list (a lagged batch stamps list [cur_list] from p2g while the side branch builds the other)
    uint32_t* block_list[2]; // active node blocks: [cur_list] this substep, [1-cur_list] previous
    uint32_t* block_count[2];// device scalars
    int cur_list;
    uint32_t* dev_stamp;     // device scalar: substep stamp for block_stamp (device-side so graphs replay correctly)
    // lagged batches (api.cu): the sort of substep s leaves the critical path
    uint32_t* rank;          // inverse of perm: sorted position of each physical slot (g2p's destination slot)
    uint32_t* keys_b;        // second key buffer: g2p reads the keys of its source slots while it writes the next ones
    uint32_t* keys_cur;      // the buffer holding the current keys: b.keys at every API boundary, keys_b inside a batch
    bool lag_ready;          // set [1-wl] / list [1-cur_list] are prebuilt for the current physical order, grid_in is clean
    uint32_t* lag_done;      // device scalar: CTAs of k_lag_clear_out that have finished
    unsigned long long* trace;  // MPMK_TRACE: %globaltimer stamps of the last lagged substep (api.cu)
    cudaStream_t side_stream;
    cudaEvent_t ev_fork, ev_clear, ev_join, ev_grid, ev_end;
    bool sorted_valid;       // sorted_keys/perm describe the current set
    bool prepared;           // work list built for the current sort
    bool ftrial_valid;
    int64_t launches;
    // CUDA graph cache for mpmk_substeps
    struct HostState {
        int cur, cur_list, wl;
        bool coherent, lag_ready, last_sort_coherent;
        uint32_t* sorted_keys;
        uint32_t* keys_alt;
    };
    struct GraphEntry {
        cudaGraphExec_t exec;
        bool stale;               // arguments out of date: candidate for cudaGraphExecUpdate
        float dt;
        mpmk_substep_ops ops;
        int n_steps;
        uint32_t flags;
        int64_t n;
        mpmk_bindings b;
        HostState start, end;     // state the capture started from / leaves behind
        int64_t kernels;          // kernel launches inside the graph (mpmk_launch_count counts them per replay)
    } graphs[MPMK_GRAPH_CACHE];
    int n_graphs;
    cudaStream_t cap_stream;  // graphs are captured here (the legacy default stream cannot be captured)
    MgState mg;
    char err[512];
};

#define MPMK_CHECK(ctx, call)                                                                   \
    do {                                                                                        \
        cudaError_t e__ = (call);                                                               \
        if (e__ != cudaSuccess) {                                                               \
            snprintf((ctx)->err, sizeof((ctx)->err), "%s:%d %s: %s", __FILE__, __LINE__, #call, \
                     cudaGetErrorString(e__));                                                  \
            return (int)e__;                                                                    \
        }                                                                                       \
    } while (0)

#define MPMK_FAIL(ctx, code, ...)                            \
    do {                                                     \
        snprintf((ctx)->err, sizeof((ctx)->err), __VA_ARGS__); \
        return (code);                                       \
    } while (0)

#define MPMK_LAUNCH_CHECK(ctx)            \
    do {                                  \
        (ctx)->launches++;                \
        MPMK_CHECK(ctx, cudaGetLastError()); \
    } while (0)

static inline int div_up(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// ---- programmatic dependent launch (opt-in: MPMK_PDL=1) ------------------------------------------
// The substep is a chain of a dozen kernels, most of them a few microseconds long; each boundary costs the drain of
// one grid plus the launch of the next.  With MPMK_PDL=1 the chain's kernels are launched with the
// programmatic-serialization attribute: the next grid's CTAs become resident while the previous grid's last CTAs
// are still running and sit in griddepcontrol.wait -- the first statement of every such kernel -- until that grid
// has completed and its writes are visible.  Every kernel of the chain waits before it touches memory, so
// completion is transitive along the chain; kernels launched the ordinary way (memsets, the multi-GPU halo kernels,
// exports) serialise as before.  Measured: it shortens a loop of eager launches (lego-like sort +
// prepare 52 -> 41 us) but not the graph replay the solver uses (lego-like 0.083 -> 0.082 ms), and costs S-4M
// 3.6 % there (waiting CTAs of the next grid take residency from the big kernels' tails) -- hence off by default;
// without the attribute the waits are no-ops.
__device__ __forceinline__ void pdl_wait() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

bool mpmk_pdl_enabled();

template <typename... P, typename... A>
static inline cudaError_t mpmk_launch(void (*kern)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, A&&... a) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = mpmk_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<P>(a)...);
}

// ---- device helpers ---------------------------------------------------------------------
__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }

// 16-byte vector reduction to global memory (sm_90+): one L2 atomic transaction per node
__device__ __forceinline__ void red_add_v4(float4* addr, float4 v) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w)
                 : "memory");
}

// Base cell of a coordinate exactly as utils.py:94-95: (int)(x*inv_dx - 0.5), unfused, truncating;
// clamped into the grid for memory safety (the reference has no clamp, quirk Q9).
__device__ __forceinline__ int base_cell(float x, float inv_dx, int n_grid, bool& clamped) {
    float gp = __fmul_rn(x, inv_dx);
    int b = (int)__fsub_rn(gp, 0.5f);
    if (b < 0) { b = 0; clamped = true; }
    if (b > n_grid - 3) { b = n_grid - 3; clamped = true; }
    return b;
}

__device__ __forceinline__ uint32_t make_key(int bx, int by, int bz, int nt) {
    uint32_t tile = (uint32_t)(((bx >> 2) * nt + (by >> 2)) * nt + (bz >> 2));
    return (tile << 6) | (uint32_t)(((bx & 3) << 4) | ((by & 3) << 2) | (bz & 3));
}

// ---- asynchronous global->shared copies (LDGSTS) and the batch stream shared by p2g and g2p ----
__device__ __forceinline__ void cp_async16(void* smem, const void* gptr) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async4(void* smem, const void* gptr) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// ---- bulk asynchronous copies (the TMA engine's 1-D path: cp.async.bulk, SASS UBLKCP) and their mbarrier ----
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
    const unsigned a = (unsigned)__cvta_generic_to_shared(bar);
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(a), "r"(parity) : "memory");
}
// global -> shared, `bytes` a multiple of 16, both addresses 16-byte aligned; completion is signalled on `bar`
__device__ __forceinline__ void bulk_g2s(void* smem, const void* gptr, unsigned bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"((unsigned)__cvta_generic_to_shared(smem)),
                 "l"(gptr), "r"(bytes), "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
// shared -> global with an fp32 add at the destination (one bulk reduction instead of one RED per node)
__device__ __forceinline__ void bulk_red_add_f32(void* gptr, const void* smem, unsigned bytes) {
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(gptr), "r"((unsigned)__cvta_generic_to_shared(smem)),
                 "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// orders this thread's earlier generic-proxy accesses to shared memory before later async-proxy (bulk copy) accesses
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

struct BatchDesc {
    uint32_t base;  // first sorted index
    uint32_t tile;
    int count;      // 0 = end of stream
    bool first;     // first batch of its item
};

// Warp-uniform stream of batches over dynamically fetched work items (one atomic on the cursor per item).
// Requesting the next item ahead of time was tried twice and is not worth it: a whole item ahead is slower (a warp that
// reserves an item early lengthens the tail of the persistent grid; items are coarse, ~7 per warp at 4 M particles), and
// taking the ticket one batch ahead (the atomic's round trip, 8 % of g2p's stall samples, then overlaps a batch of work)
// measures the same as waiting for it: other warps fill the gap.
struct BatchStream {
    uint4 item;  // current item {begin, count, tile, -}
    uint32_t off;
    bool done;
    __device__ __forceinline__ void init() { item = make_uint4(0u, 0u, 0u, 0u); off = 0; done = false; }
    __device__ __forceinline__ BatchDesc next(int lane, const uint4* __restrict__ items, uint32_t* cursor, uint32_t n_items) {
        BatchDesc d;
        d.base = 0; d.tile = 0; d.count = 0; d.first = false;
        if (done) return d;
        if (off >= item.y) {
            uint32_t it = 0;
            if (lane == 0) it = atomicAdd(cursor, 1u);
            it = __shfl_sync(0xffffffffu, it, 0);
            if (it >= n_items) { done = true; return d; }
            item = items[it];
            off = 0;
        }
        d.base = item.x + off;
        d.tile = item.z;
        d.count = (int)min(32u, item.y - off);
        d.first = (off == 0);
        off += 32;
        return d;
    }
};

// internal launchers (one per translation unit)
int mpmk_launch_keys(mpmk_ctx* ctx, cudaStream_t s);
int mpmk_launch_sort(mpmk_ctx* ctx, cudaStream_t s);
int mpmk_launch_prepare(mpmk_ctx* ctx, cudaStream_t s);
int mpmk_launch_prepare_lagged(mpmk_ctx* ctx, cudaStream_t s);
int mpmk_launch_lag_clear_out(mpmk_ctx* ctx, cudaStream_t s);
int mpmk_launch_lag_clear_in(mpmk_ctx* ctx, cudaStream_t s);
int mpmk_worklist_init(mpmk_ctx* ctx);
int mpmk_launch_lag_enter(mpmk_ctx* ctx, cudaStream_t s);
int mpmk_launch_p2g_lagged(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, cudaStream_t s);
int mpmk_launch_grid_lagged(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, cudaStream_t s);
int mpmk_launch_g2p_lagged(mpmk_ctx* ctx, float dt, cudaStream_t s);
int mpmk_launch_sort_lagged(mpmk_ctx* ctx, cudaStream_t s);
void mpmk_lag_flip(mpmk_ctx* ctx);
void mpmk_trace(mpmk_ctx* ctx, cudaStream_t s, int slot);
int mpmk_item_size(const mpmk_ctx* ctx);
int mpmk_launch_p2g(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, cudaStream_t s);
int mpmk_launch_grid(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, cudaStream_t s);
int mpmk_launch_g2p(mpmk_ctx* ctx, float dt, cudaStream_t s);
int mpmk_sort_init(mpmk_ctx* ctx);
int mpmk_p2g_init(mpmk_ctx* ctx);
int mpmk_g2p_init(mpmk_ctx* ctx);
void mpmk_drop_graphs(mpmk_ctx* ctx);
int mpmk_mg_halo_sum(mpmk_ctx* ctx, cudaStream_t s);
extern "C" int mpmk_mg_finalize(mpmk_ctx* ctx);
