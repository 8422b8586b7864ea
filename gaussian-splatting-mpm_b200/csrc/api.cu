// api.cu -- C ABI (include/mpmk.h): context lifecycle, the substep driver (solver.py:27-52 order) and
// CUDA-graph replay of constant-dt substep batches.
#include "common.cuh"
#include <cstdlib>

bool mpmk_pdl_enabled() {
    static const bool on = getenv("MPMK_PDL") && atoi(getenv("MPMK_PDL")) != 0;   // opt-in: see common.cuh
    return on;
}

static int ceil_log2(uint64_t v) { int b = 0; while ((1ull << b) < v) ++b; return b; }

extern "C" int mpmk_version(void) { return MPMK_VERSION; }

extern "C" const char* mpmk_last_error(mpmk_ctx* ctx) { return ctx ? ctx->err : "null context"; }

#define ALLOC0(ctx, ptr, bytes)                                        \
    do {                                                               \
        MPMK_CHECK(ctx, cudaMalloc((void**)&(ptr), (bytes)));          \
        MPMK_CHECK(ctx, cudaMemset((ptr), 0, (bytes)));                \
    } while (0)

static int create_impl(mpmk_ctx* ctx, int device, int n_grid, double grid_extent, int64_t capacity) {
    MPMK_CHECK(ctx, cudaSetDevice(device));
    cudaDeviceProp prop;
    MPMK_CHECK(ctx, cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "libmpmk is built for sm_100a only (device is sm_%d%d)", prop.major, prop.minor);
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    ctx->n_grid = n_grid;
    ctx->nt = (n_grid + MPMK_TILE - 1) / MPMK_TILE;
    ctx->nb = (n_grid + 3) / 4;
    ctx->dx = (float)(grid_extent / (double)n_grid);      // model.py:15
    ctx->inv_dx = (float)((double)n_grid / grid_extent);  // model.py:16
    ctx->capacity = capacity;
    ctx->key_bits = 6 + ceil_log2((uint64_t)ctx->nt * ctx->nt * ctx->nt);
    if (ctx->key_bits > 32) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "grid too large for 32-bit keys");
    const size_t cap = (size_t)(capacity > 0 ? capacity : 1);
    ALLOC0(ctx, ctx->keys_alt, cap * 4);
    ALLOC0(ctx, ctx->sorted_keys, cap * 4);
    ALLOC0(ctx, ctx->perm, cap * 4);
    ALLOC0(ctx, ctx->perm_alt, cap * 4);
    ctx->n_sort_ctas_cap = div_up((int64_t)cap, MPMK_SORT_CHUNK);
    ALLOC0(ctx, ctx->sort_ctrl, ((size_t)8 + 4 * 512 + ((size_t)2 * 4 * (1 << MPMK_MAX_RADIX_BITS) + 2) * (ctx->n_sort_ctas_cap + 1)) * 4);   // (x 2: the lean mover passes of a lagged batch use 2048-key tiles)
    for (int i = 0; i < 2; ++i) { ALLOC0(ctx, ctx->cs_mk[i], cap * 4); ALLOC0(ctx, ctx->cs_mv[i], cap * 4); }
    ALLOC0(ctx, ctx->cs_mask, (cap / 32 + 2) * 4);
    ALLOC0(ctx, ctx->cs_mprefix, (cap / 32 + 2) * 4);
    ALLOC0(ctx, ctx->cs_cnt, (cap / 32 + 2) * 4);
    ALLOC0(ctx, ctx->hist_tot, 4 * 512 * 4);
    const size_t ntiles = (size_t)ctx->nt * ctx->nt * ctx->nt;
    ctx->items_cap = (int64_t)(cap / MPMK_ITEM_MIN + (cap < ntiles ? cap : ntiles) + 8);
    for (int i = 0; i < 2; ++i) {
        ALLOC0(ctx, ctx->items_buf[i], (size_t)ctx->items_cap * sizeof(uint4));
        ALLOC0(ctx, ctx->counters_buf[i], sizeof(DevCounters));
    }
    ctx->wl = 0; ctx->items = ctx->items_buf[0]; ctx->counters = ctx->counters_buf[0];
    ALLOC0(ctx, ctx->rank, cap * 4);
    ALLOC0(ctx, ctx->keys_b, cap * 4);
    ALLOC0(ctx, ctx->lag_done, 4);
    ALLOC0(ctx, ctx->trace, 16 * sizeof(unsigned long long));
    ALLOC0(ctx, ctx->tile_list, ((cap < ntiles ? cap : ntiles) + 8) * sizeof(uint2));
    ALLOC0(ctx, ctx->tile_end, ntiles * 4);
    const size_t nblocks = (size_t)ctx->nb * ctx->nb * ctx->nb;
    ALLOC0(ctx, ctx->block_stamp[0], nblocks * 4);
    ALLOC0(ctx, ctx->block_stamp[1], nblocks * 4);
    ALLOC0(ctx, ctx->block_list[0], nblocks * 4);
    ALLOC0(ctx, ctx->block_list[1], nblocks * 4);
    ALLOC0(ctx, ctx->block_count[0], 4);
    ALLOC0(ctx, ctx->block_count[1], 4);
    ALLOC0(ctx, ctx->dev_stamp, 4);
    MPMK_CHECK(ctx, cudaStreamCreateWithFlags(&ctx->cap_stream, cudaStreamNonBlocking));
    MPMK_CHECK(ctx, cudaStreamCreateWithFlags(&ctx->side_stream, cudaStreamNonBlocking));
    MPMK_CHECK(ctx, cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
    MPMK_CHECK(ctx, cudaEventCreateWithFlags(&ctx->ev_clear, cudaEventDisableTiming));
    MPMK_CHECK(ctx, cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
    MPMK_CHECK(ctx, cudaEventCreateWithFlags(&ctx->ev_grid, cudaEventDisableTiming));
    MPMK_CHECK(ctx, cudaEventCreateWithFlags(&ctx->ev_end, cudaEventDisableTiming));
    if (n_grid > 1023) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "n_grid > 1023 not supported (p2g packs a base cell into 30 bits)");
    { int rc = mpmk_sort_init(ctx); if (rc) return rc; }
    { int rc = mpmk_p2g_init(ctx); if (rc) return rc; }
    { int rc = mpmk_g2p_init(ctx); if (rc) return rc; }
    { int rc = mpmk_worklist_init(ctx); if (rc) return rc; }
    ctx->prm.alpha = 0.0f;
    return 0;
}

extern "C" int mpmk_create(mpmk_ctx** out, int device, int n_grid, double grid_extent, int64_t capacity) {
    if (!out || n_grid < 3 || grid_extent <= 0.0 || capacity < 0) return MPMK_ERR_INVALID;
    mpmk_ctx* ctx = new mpmk_ctx();
    memset(ctx, 0, sizeof(*ctx));
    *out = ctx;  // returned even on failure so the caller can read the message, then destroy
    return create_impl(ctx, device, n_grid, grid_extent, capacity);
}

extern "C" int mpmk_destroy(mpmk_ctx* ctx) {
    if (!ctx) return 0;
    mpmk_mg_finalize(ctx);
    for (int i = 0; i < ctx->n_graphs; ++i) cudaGraphExecDestroy(ctx->graphs[i].exec);
    cudaFree(ctx->keys_alt); cudaFree(ctx->sorted_keys); cudaFree(ctx->perm); cudaFree(ctx->perm_alt);
    for (int i = 0; i < 2; ++i) { cudaFree(ctx->cs_mk[i]); cudaFree(ctx->cs_mv[i]); }
    cudaFree(ctx->cs_mask); cudaFree(ctx->cs_mprefix); cudaFree(ctx->cs_cnt);
    cudaFree(ctx->sort_ctrl); cudaFree(ctx->hist_tot); cudaFree(ctx->tile_list); cudaFree(ctx->tile_end);
    for (int i = 0; i < 2; ++i) { cudaFree(ctx->items_buf[i]); cudaFree(ctx->counters_buf[i]); cudaFree(ctx->block_stamp[i]); }
    cudaFree(ctx->block_list[0]); cudaFree(ctx->block_list[1]); cudaFree(ctx->block_count[0]); cudaFree(ctx->block_count[1]);
    cudaFree(ctx->dev_stamp); cudaFree(ctx->rank); cudaFree(ctx->keys_b); cudaFree(ctx->lag_done); cudaFree(ctx->trace);
    if (ctx->cap_stream) cudaStreamDestroy(ctx->cap_stream);
    if (ctx->side_stream) cudaStreamDestroy(ctx->side_stream);
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->ev_clear) cudaEventDestroy(ctx->ev_clear);
    if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    if (ctx->ev_grid) cudaEventDestroy(ctx->ev_grid);
    if (ctx->ev_end) cudaEventDestroy(ctx->ev_end);
    delete ctx;
    return 0;
}

// The captured batches no longer describe the context (particle count, bindings, slab, scalar parameters changed).
// The executable graphs are kept, marked stale: the next capture of a batch of the same shape updates one of them in
// place (cudaGraphExecUpdate: same topology, new kernel arguments) instead of instantiating a new one, which is what a
// slab pays after every migration and a frame loop after every reload.
void mpmk_drop_graphs(mpmk_ctx* ctx) {
    for (int i = 0; i < ctx->n_graphs; ++i) ctx->graphs[i].stale = true;
}

extern "C" int mpmk_bind(mpmk_ctx* ctx, const mpmk_bindings* b) {
    if (!ctx || !b) return MPMK_ERR_INVALID;
    for (int s = 0; s < 2; ++s) {
        const mpmk_particle_set& p = b->set[s];
        if (!p.xm || !p.vv || !p.c0 || !p.c1 || !p.cs || !p.s1 || !p.f0 || !p.f1 || !p.f2 || !p.meta)
            MPMK_FAIL(ctx, MPMK_ERR_INVALID, "mpmk_bind: particle set %d has a null array", s);
    }
    if (!b->keys || !b->grid_in || !b->grid_out) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "mpmk_bind: keys/grid_in/grid_out must be bound");
    ctx->b = *b;
    ctx->keys_cur = (uint32_t*)b->keys;
    ctx->lag_ready = false;
    ctx->bound = true;
    ctx->coherent = false;
    ctx->sorted_valid = false;
    ctx->prepared = false;
    ctx->ftrial_valid = false;
    mpmk_drop_graphs(ctx);
    return 0;
}

extern "C" int mpmk_set_params(mpmk_ctx* ctx, const mpmk_params* p) {
    if (!ctx || !p) return MPMK_ERR_INVALID;
    // scalar parameters (gravity, alpha, xi, ...) are passed to the kernels by value, so a captured graph has them
    // baked in: anything but a pure flags change invalidates the captured batches (flags are part of the cache key)
    mpmk_params a = ctx->prm, b = *p;
    a.flags = 0u; b.flags = 0u;
    if (memcmp(&a, &b, sizeof(a)) != 0) mpmk_drop_graphs(ctx);
    ctx->prm = *p;
    return 0;
}

extern "C" int mpmk_set_count(mpmk_ctx* ctx, int64_t n, int current_set) {
    if (!ctx) return MPMK_ERR_INVALID;
    if (n < 0 || n > ctx->capacity || (current_set != 0 && current_set != 1))
        MPMK_FAIL(ctx, MPMK_ERR_INVALID, "mpmk_set_count: n=%lld capacity=%lld set=%d", (long long)n, (long long)ctx->capacity, current_set);
    ctx->n = n;
    ctx->cur = current_set;
    ctx->lag_ready = false;
    ctx->coherent = false;
    ctx->sorted_valid = false;
    ctx->prepared = false;
    ctx->ftrial_valid = false;
    return 0;
}

extern "C" int mpmk_current_set(mpmk_ctx* ctx) { return ctx ? ctx->cur : -1; }
extern "C" int64_t mpmk_count(mpmk_ctx* ctx) { return ctx ? ctx->n : -1; }
extern "C" int64_t mpmk_launch_count(mpmk_ctx* ctx) { return ctx ? ctx->launches : -1; }

#define REQUIRE_BOUND(ctx)                                                        \
    do {                                                                          \
        if (!(ctx)) return MPMK_ERR_INVALID;                                      \
        if (!(ctx)->bound) MPMK_FAIL(ctx, MPMK_ERR_STATE, "call mpmk_bind first"); \
    } while (0)

extern "C" int mpmk_compute_keys(mpmk_ctx* ctx, void* stream) {
    REQUIRE_BOUND(ctx);
    return mpmk_launch_keys(ctx, (cudaStream_t)stream);
}
extern "C" int mpmk_sort(mpmk_ctx* ctx, void* stream) {
    REQUIRE_BOUND(ctx);
    return mpmk_launch_sort(ctx, (cudaStream_t)stream);
}
extern "C" int mpmk_get_sort_result(mpmk_ctx* ctx, void* sorted_keys_dev, void* perm_dev, void* stream) {
    REQUIRE_BOUND(ctx);
    if (!ctx->sorted_valid) MPMK_FAIL(ctx, MPMK_ERR_STATE, "no valid sort result");
    if (sorted_keys_dev) MPMK_CHECK(ctx, cudaMemcpyAsync(sorted_keys_dev, ctx->sorted_keys, ctx->n * 4, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    if (perm_dev) MPMK_CHECK(ctx, cudaMemcpyAsync(perm_dev, ctx->perm, ctx->n * 4, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return 0;
}
extern "C" int mpmk_get_last_sort(mpmk_ctx* ctx, void* sorted_keys_dev, void* perm_dev, void* rank_dev, void* stream) {
    REQUIRE_BOUND(ctx);
    cudaStream_t s = (cudaStream_t)stream;
    if (sorted_keys_dev) MPMK_CHECK(ctx, cudaMemcpyAsync(sorted_keys_dev, ctx->sorted_keys, ctx->n * 4, cudaMemcpyDeviceToDevice, s));
    if (perm_dev) MPMK_CHECK(ctx, cudaMemcpyAsync(perm_dev, ctx->perm, ctx->n * 4, cudaMemcpyDeviceToDevice, s));
    if (rank_dev) MPMK_CHECK(ctx, cudaMemcpyAsync(rank_dev, ctx->rank, ctx->n * 4, cudaMemcpyDeviceToDevice, s));
    return 0;
}
extern "C" int mpmk_prepare_grid(mpmk_ctx* ctx, void* stream) {
    REQUIRE_BOUND(ctx);
    return mpmk_launch_prepare(ctx, (cudaStream_t)stream);
}
extern "C" int mpmk_p2g(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, void* stream) {
    REQUIRE_BOUND(ctx);
    return mpmk_launch_p2g(ctx, dt, ops, (cudaStream_t)stream);
}
extern "C" int mpmk_grid_update(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, void* stream) {
    REQUIRE_BOUND(ctx);
    return mpmk_launch_grid(ctx, dt, ops, (cudaStream_t)stream);
}
extern "C" int mpmk_g2p(mpmk_ctx* ctx, float dt, void* stream) {
    REQUIRE_BOUND(ctx);
    return mpmk_launch_g2p(ctx, dt, (cudaStream_t)stream);
}

static int substep_impl(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, cudaStream_t s) {
    int rc;
    if ((rc = mpmk_launch_sort(ctx, s))) return rc;
    if ((rc = mpmk_launch_prepare(ctx, s))) return rc;
    if ((rc = mpmk_launch_p2g(ctx, dt, ops, s))) return rc;
    if ((rc = mpmk_mg_halo_sum(ctx, s))) return rc;  // no-op unless a slab is set (mg.cu)
    if ((rc = mpmk_launch_grid(ctx, dt, ops, s))) return rc;
    return mpmk_launch_g2p(ctx, dt, s);
}

extern "C" int mpmk_substep(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, void* stream) {
    REQUIRE_BOUND(ctx);
    return substep_impl(ctx, dt, ops, (cudaStream_t)stream);
}

// ---- lagged substeps ----------------------------------------------------------------------------
// Inside a batch the sort leaves the critical path.  The physical order of the particles is the sorted order of the
// previous substep (g2p writes it), and one substep moves ~3.5 % of the particles to another cell and ~1 % to
// another tile, so p2g and g2p of substep s run on the work list built from the sort of substep s-1: the true cell of
// a particle comes from its current key; one that left its item's tile scatters straight to the global grid and
// gathers from it.  Meanwhile, on a forked branch of the graph, the sort of substep s (still the
// bit-exact stable sort of the current keys), the work list and the block list for substep s+1 are built; g2p joins
// that branch and writes every particle to the slot its rank says, which re-establishes physical order == sorted
// order of substep s.  The clear pass disappears as well: the grid pass zeroes grid_in behind itself and the side
// branch resets the velocities of the substep before.  MPMK_LAG=0 (or multi-GPU slabs) keeps every substep on the
// sequential sort -> prepare -> p2g -> grid -> g2p chain; the first substep after a (re)load always is.
static bool lag_enabled() {
    static const bool on = !(getenv("MPMK_LAG") && atoi(getenv("MPMK_LAG")) == 0);
    return on;
}

// MPMK_TRACE=1: one-thread kernels between the stages write %globaltimer into a small buffer, which gives the
// timeline of the two branches inside a replayed graph (read back with mpmk_debug_trace; tuning aid)
__global__ void k_trace_stamp(unsigned long long* buf, int slot) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    buf[slot] = t;
}
static bool trace_enabled() {
    static const bool on = getenv("MPMK_TRACE") && atoi(getenv("MPMK_TRACE")) != 0;
    return on;
}
void mpmk_trace(mpmk_ctx* ctx, cudaStream_t s, int slot) { if (trace_enabled()) k_trace_stamp<<<1, 1, 0, s>>>(ctx->trace, slot); }
#define TRACE(ctx, stream, slot) mpmk_trace(ctx, stream, slot)

extern "C" int mpmk_debug_trace(mpmk_ctx* ctx, unsigned long long* out16) {
    if (!ctx || !out16) return MPMK_ERR_INVALID;
    MPMK_CHECK(ctx, cudaMemcpy(out16, ctx->trace, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    return 0;
}

static int lagged_substep(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, cudaStream_t s) {
    int rc;
    static const bool serial = getenv("MPMK_LAG_SERIAL") && atoi(getenv("MPMK_LAG_SERIAL")) != 0;   // diagnosis: no fork
    cudaStream_t side = serial ? s : ctx->side_stream;
    if (!ctx->lag_ready && (rc = mpmk_launch_lag_enter(ctx, s))) return rc;
    mpmk_lag_flip(ctx);   // the prebuilt work list / block list are this substep's
    TRACE(ctx, s, 0);
    MPMK_CHECK(ctx, cudaEventRecord(ctx->ev_fork, s));
    MPMK_CHECK(ctx, cudaStreamWaitEvent(side, ctx->ev_fork, 0));
    // side branch: velocities of the substep before; stable sort of the current keys (+ ranks); lists for the next substep
    if ((rc = mpmk_launch_lag_clear_out(ctx, side))) return rc;
    TRACE(ctx, side, 5);
    MPMK_CHECK(ctx, cudaEventRecord(ctx->ev_clear, side));
    if ((rc = mpmk_launch_sort_lagged(ctx, side))) return rc;
    TRACE(ctx, side, 6);
    if ((rc = mpmk_launch_prepare_lagged(ctx, side))) return rc;
    TRACE(ctx, side, 7);
    MPMK_CHECK(ctx, cudaEventRecord(ctx->ev_join, side));
    // critical path
    if ((rc = mpmk_launch_p2g_lagged(ctx, dt, ops, s))) return rc;
    if ((rc = mpmk_mg_halo_sum(ctx, s))) return rc;  // no-op unless a slab is set (mg.cu)
    TRACE(ctx, s, 1);
    MPMK_CHECK(ctx, cudaStreamWaitEvent(s, ctx->ev_clear, 0));
    if ((rc = mpmk_launch_grid_lagged(ctx, dt, ops, s))) return rc;
    TRACE(ctx, s, 2);
    // grid_in is reset for the next scatter beside g2p (side stream again: its work for this substep is done by then)
    MPMK_CHECK(ctx, cudaEventRecord(ctx->ev_grid, s));
    MPMK_CHECK(ctx, cudaStreamWaitEvent(side, ctx->ev_grid, 0));
    if ((rc = mpmk_launch_lag_clear_in(ctx, side))) return rc;
    MPMK_CHECK(ctx, cudaEventRecord(ctx->ev_end, side));
    MPMK_CHECK(ctx, cudaStreamWaitEvent(s, ctx->ev_join, 0));
    if ((rc = mpmk_launch_g2p_lagged(ctx, dt, s))) return rc;
    MPMK_CHECK(ctx, cudaStreamWaitEvent(s, ctx->ev_end, 0));
    TRACE(ctx, s, 3);
    ctx->lag_ready = true;
    return 0;
}

// n >= 1 substeps of constant dt and ops.  F_trial (and the yield stress before the return map) is only consumed by
// export and by mpmk_constitutive_ex: it is stored by the last g2p of the batch.
static int substeps_impl(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, int n_steps, cudaStream_t s, bool allow_lag) {
    const uint32_t user_flags = ctx->prm.flags;
    int rc = 0;
    for (int i = 0; i < n_steps && rc == 0; ++i) {
        const bool last = (i == n_steps - 1);
        ctx->prm.flags = last ? user_flags : (user_flags & ~MPMK_WRITE_FTRIAL);
        // (slabs: the peer-memory halo is plain kernels and is captured like everything else; the NCCL halo path is issued
        // eagerly and stays sequential)
        if (allow_lag && ctx->coherent && ctx->n > 0 && (!ctx->mg.enabled || ctx->mg.peer) && !(user_flags & MPMK_NO_LAG))
            rc = lagged_substep(ctx, dt, ops, s);
        else rc = substep_impl(ctx, dt, ops, s);
    }
    ctx->prm.flags = user_flags;
    if (rc == 0 && ctx->keys_cur != (uint32_t*)ctx->b.keys) {   // the caller's key array holds the current keys at every API boundary
        MPMK_CHECK(ctx, cudaMemcpyAsync(ctx->b.keys, ctx->keys_cur, (size_t)ctx->n * 4, cudaMemcpyDeviceToDevice, s));
        ctx->keys_cur = (uint32_t*)ctx->b.keys;
    }
    return rc;
}

static bool same_bindings(const mpmk_bindings& a, const mpmk_bindings& b) { return memcmp(&a, &b, sizeof(a)) == 0; }

// the host-side state a batch starts from / leaves behind (graphs are keyed on the former and restore the latter)
static void host_state_get(const mpmk_ctx* ctx, mpmk_ctx::HostState* h) {
    h->cur = ctx->cur; h->cur_list = ctx->cur_list; h->wl = ctx->wl; h->coherent = ctx->coherent; h->lag_ready = ctx->lag_ready;
    h->last_sort_coherent = ctx->last_sort_coherent; h->sorted_keys = ctx->sorted_keys; h->keys_alt = ctx->keys_alt;
}
static void host_state_set(mpmk_ctx* ctx, const mpmk_ctx::HostState& h) {
    ctx->cur = h.cur; ctx->cur_list = h.cur_list; ctx->wl = h.wl; ctx->coherent = h.coherent; ctx->lag_ready = h.lag_ready;
    ctx->last_sort_coherent = h.last_sort_coherent; ctx->sorted_keys = h.sorted_keys; ctx->keys_alt = h.keys_alt;
    ctx->items = ctx->items_buf[h.wl]; ctx->counters = ctx->counters_buf[h.wl];
    ctx->keys_cur = (uint32_t*)ctx->b.keys;
}
static bool host_state_same(const mpmk_ctx::HostState& a, const mpmk_ctx::HostState& b) {
    return a.cur == b.cur && a.cur_list == b.cur_list && a.wl == b.wl && a.coherent == b.coherent && a.lag_ready == b.lag_ready &&
           a.sorted_keys == b.sorted_keys && a.keys_alt == b.keys_alt;
}

// find or capture the CUDA graph of n_steps substeps from the current context state, and replay it
static int graph_substeps(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, int n_steps, cudaStream_t s) {
    const uint32_t user_flags = ctx->prm.flags;
    mpmk_ctx::HostState now;
    host_state_get(ctx, &now);
    // look for a captured graph with identical launch parameters
    for (int i = 0; i < ctx->n_graphs; ++i) {
        mpmk_ctx::GraphEntry& g = ctx->graphs[i];
        if (!g.stale && g.n_steps == n_steps && g.dt == dt && g.flags == user_flags && g.n == ctx->n && host_state_same(g.start, now) &&
            memcmp(&g.ops, ops, sizeof(*ops)) == 0 && same_bindings(g.b, ctx->b)) {
            MPMK_CHECK(ctx, cudaGraphLaunch(g.exec, s));
            ctx->launches += g.kernels;
            host_state_set(ctx, g.end);
            ctx->ftrial_valid = (user_flags & MPMK_WRITE_FTRIAL) != 0;
            ctx->sorted_valid = false;
            ctx->prepared = false;
            return 0;
        }
    }
    // capture
    mpmk_ctx::GraphEntry e;
    memset(&e, 0, sizeof(e));
    e.n_steps = n_steps; e.dt = dt; e.flags = user_flags; e.n = ctx->n;
    e.ops = *ops; e.b = ctx->b; e.start = now;
    const int64_t launches_before = ctx->launches;
    cudaStream_t cs = ctx->cap_stream;
    MPMK_CHECK(ctx, cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
    const int rc = substeps_impl(ctx, dt, ops, n_steps, cs, lag_enabled());
    cudaGraph_t graph = nullptr;
    cudaError_t ce = cudaStreamEndCapture(cs, &graph);
    cudaGraphExec_t exec = nullptr;
    int reused = -1;
    if (rc == 0 && ce == cudaSuccess) {
        // a stale executable of the same shape (substeps, lagged or not from the start) takes the new arguments in place
        for (int i = 0; i < ctx->n_graphs && reused < 0; ++i) {
            mpmk_ctx::GraphEntry& g = ctx->graphs[i];
            if (!g.stale || g.n_steps != n_steps || g.start.lag_ready != now.lag_ready || g.start.coherent != now.coherent) continue;
            cudaGraphExecUpdateResultInfo info;
            if (cudaGraphExecUpdate(g.exec, graph, &info) == cudaSuccess) { exec = g.exec; reused = i; }
            else cudaGetLastError();   // topology differs (other op count, flags, merge vs full sort): instantiate below
        }
        if (reused < 0) ce = cudaGraphInstantiate(&exec, graph, 0);
    }
    if (graph) cudaGraphDestroy(graph);
    if (rc || ce != cudaSuccess) {
        // nothing ran: put the context back where the capture started (ping-pong state, key buffer, validity flags)
        host_state_set(ctx, now);
        ctx->sorted_valid = false; ctx->prepared = false; ctx->prm.flags = user_flags;
        ctx->launches = launches_before;
        if (rc) return rc;
        MPMK_CHECK(ctx, ce);
    }
    e.exec = exec;
    e.stale = false;
    e.kernels = ctx->launches - launches_before;   // kernels one replay executes
    host_state_get(ctx, &e.end);
    if (reused >= 0) {
        ctx->graphs[reused] = e;
    } else {
        if (ctx->n_graphs == MPMK_GRAPH_CACHE) {  // evict a stale entry if there is one, else the oldest
            int victim = 0;
            for (int i = 0; i < ctx->n_graphs; ++i) if (ctx->graphs[i].stale) { victim = i; break; }
            cudaGraphExecDestroy(ctx->graphs[victim].exec);
            for (int i = victim + 1; i < MPMK_GRAPH_CACHE; ++i) ctx->graphs[i - 1] = ctx->graphs[i];
            ctx->n_graphs = MPMK_GRAPH_CACHE - 1;
        }
        ctx->graphs[ctx->n_graphs++] = e;
    }
    MPMK_CHECK(ctx, cudaGraphLaunch(exec, s));
    ctx->ftrial_valid = (user_flags & MPMK_WRITE_FTRIAL) != 0;
    return 0;
}

extern "C" int mpmk_read_counters(mpmk_ctx* ctx, mpmk_counters* out, void* stream) {
    REQUIRE_BOUND(ctx);
    if (!out) return MPMK_ERR_INVALID;
    DevCounters h;
    uint32_t nb = 0;
    cudaStream_t s = (cudaStream_t)stream;
    MPMK_CHECK(ctx, cudaMemcpyAsync(&h, ctx->counters, sizeof(h), cudaMemcpyDeviceToHost, s));
    MPMK_CHECK(ctx, cudaMemcpyAsync(&nb, ctx->block_count[ctx->cur_list], 4, cudaMemcpyDeviceToHost, s));
    MPMK_CHECK(ctx, cudaStreamSynchronize(s));
    out->n_items = h.n_items;
    out->n_active_blocks = nb;
    out->n_tiles = h.n_tiles;
    out->n_clamped = h.n_clamped;
    out->n_crossers = h.n_cross;
    return 0;
}

extern "C" int mpmk_substeps(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, int n_steps, void* stream) {
    REQUIRE_BOUND(ctx);
    if (n_steps <= 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    mpmk_substep_ops zero_ops;
    memset(&zero_ops, 0, sizeof(zero_ops));
    if (!ops) ops = &zero_ops;
    if (!ctx->mg.enabled) return graph_substeps(ctx, dt, ops, n_steps, s);
    // slab mode over NCCL: the substeps carry ncclSend/ncclRecv; issued eagerly on the caller's stream (no host
    // synchronisation in between), not captured
    if (!ctx->mg.peer) return substeps_impl(ctx, dt, ops, n_steps, s, false);
    // slab mode over peer memory: plain kernels, so they can be replayed -- in graphs of TWO substeps, which leave
    // the ping-pong state (particle set, block list, sorted-key buffer) where they found it: one capture serves
    // until the next migration changes the particle count or the row range (mg.cu drops the graphs then)
    const uint32_t user_flags = ctx->prm.flags;
    const uint32_t quiet = user_flags & ~MPMK_WRITE_FTRIAL;   // F_trial: last substep of the batch only
    int left = n_steps, rc = 0;
    if (!ctx->coherent) {   // first substep after a migration: full radix sort, not worth a graph of its own
        ctx->prm.flags = left == 1 ? user_flags : quiet;
        rc = substeps_impl(ctx, dt, ops, 1, s, false);
        left -= 1;
    }
    // (the classic -> lagged transition outside the graphs: every two-substep graph then starts from the same state)
    if (rc == 0 && left >= 2 && lag_enabled() && !(user_flags & MPMK_NO_LAG) && ctx->coherent && ctx->n > 0 && !ctx->lag_ready)
        rc = mpmk_launch_lag_enter(ctx, s);
    while (rc == 0 && left >= 2) {
        ctx->prm.flags = left == 2 ? user_flags : quiet;
        rc = graph_substeps(ctx, dt, ops, 2, s);
        left -= 2;
    }
    if (rc == 0 && left == 1) {
        ctx->prm.flags = user_flags;
        rc = substeps_impl(ctx, dt, ops, 1, s, false);
    }
    ctx->prm.flags = user_flags;
    return rc;
}
