// mg.cu -- multi-GPU slab decomposition of the substep (no counterpart in the reference, which is
// single-GPU: SURVEY.md 2.3 / 8e).  One process per GPU; rank r owns the base-cell planes
// [x_lo, x_hi) of the grid's slowest axis, and between migrations its particles may sit up to
// `ghost` cells outside that range.  A particle with base plane b scatters to planes b..b+2, so the
// only nodes two ranks both contribute to are the SHARED planes
//     S(r, r+1) = [x_hi(r) - ghost, x_hi(r) + ghost + 2).
// After p2g both neighbours exchange their partial sums of the shared planes (planes are contiguous
// n^2 float4 slices of grid_in: no packing) with ncclSend/ncclRecv in one group on the caller's
// stream, and each adds what it received: a + b on one side, b + a on the other, bit-identical.  Both
// then run the (deterministic, per-node) grid update on the shared planes themselves, so no
// velocities travel back.  The shared planes are cleared and updated densely; everything else stays
// on the sparse block lists.
//
// NCCL is reached through dlopen of the library torch already loaded (path passed by the host layer):
// torch.distributed does the rendezvous plumbing (broadcast of the ncclUniqueId), the data path is
// this file's ncclSend/ncclRecv over NVLink.
#include <dlfcn.h>
#include <nccl.h>

#include "common.cuh"

struct MgApi {
    void* handle;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*);
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*GroupStart)();
    ncclResult_t (*GroupEnd)();
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    const char* (*GetErrorString)(ncclResult_t);
};
static MgApi g_nccl = {nullptr};

static int load_nccl(const char* path, char* err, size_t errlen) {
    if (g_nccl.handle) return 0;
    void* h = dlopen(path && path[0] ? path : "libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { snprintf(err, errlen, "dlopen(%s): %s", path ? path : "libnccl.so.2", dlerror()); return MPMK_ERR_STATE; }
#define SYM(field, name)                                                                      \
    do {                                                                                      \
        *(void**)(&g_nccl.field) = dlsym(h, name);                                            \
        if (!g_nccl.field) { snprintf(err, errlen, "dlsym(%s) failed", name); return MPMK_ERR_STATE; } \
    } while (0)
    SYM(GetUniqueId, "ncclGetUniqueId");
    SYM(CommInitRank, "ncclCommInitRank");
    SYM(CommDestroy, "ncclCommDestroy");
    SYM(GroupStart, "ncclGroupStart");
    SYM(GroupEnd, "ncclGroupEnd");
    SYM(Send, "ncclSend");
    SYM(Recv, "ncclRecv");
    SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
    g_nccl.handle = h;
    return 0;
}

#define NCCL_CHECK(ctx, call)                                                                          \
    do {                                                                                               \
        ncclResult_t r__ = (call);                                                                     \
        if (r__ != ncclSuccess) {                                                                      \
            snprintf((ctx)->err, sizeof((ctx)->err), "%s:%d %s: %s", __FILE__, __LINE__, #call,        \
                     g_nccl.GetErrorString(r__));                                                      \
            return MPMK_ERR_STATE;                                                                     \
        }                                                                                              \
    } while (0)

extern "C" int mpmk_mg_unique_id(const char* nccl_lib_path, void* id_out, int id_bytes) {
    char err[256];
    if (!id_out || id_bytes < (int)sizeof(ncclUniqueId)) return MPMK_ERR_INVALID;
    if (load_nccl(nccl_lib_path, err, sizeof(err))) { fprintf(stderr, "libmpmk: %s\n", err); return MPMK_ERR_STATE; }
    ncclUniqueId id;
    if (g_nccl.GetUniqueId(&id) != ncclSuccess) return MPMK_ERR_STATE;
    memcpy(id_out, &id, sizeof(id));
    return 0;
}

extern "C" int mpmk_mg_init(mpmk_ctx* ctx, const char* nccl_lib_path, const void* id, int id_bytes, int rank, int world) {
    if (!ctx || !id || id_bytes < (int)sizeof(ncclUniqueId) || world < 1 || rank < 0 || rank >= world) return MPMK_ERR_INVALID;
    if (load_nccl(nccl_lib_path, ctx->err, sizeof(ctx->err))) return MPMK_ERR_STATE;
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof(uid));
    MPMK_CHECK(ctx, cudaSetDevice(ctx->device));
    ncclComm_t comm;
    NCCL_CHECK(ctx, g_nccl.CommInitRank(&comm, world, uid, rank));
    ctx->mg.comm = (void*)comm;
    ctx->mg.rank = rank;
    ctx->mg.world = world;
    ctx->mg.enabled = false;  // until a slab is set
    return 0;
}

extern "C" int mpmk_mg_finalize(mpmk_ctx* ctx) {
    if (!ctx) return 0;
    if (ctx->mg.comm && g_nccl.handle) g_nccl.CommDestroy((ncclComm_t)ctx->mg.comm);
    ctx->mg.comm = nullptr;
    ctx->mg.enabled = false;
    return 0;
}

extern "C" int mpmk_mg_set_slab(mpmk_ctx* ctx, int x_lo, int x_hi, int ghost, int y_lo, int y_hi, void* recv_lo_dev,
                                void* recv_hi_dev) {
    if (!ctx) return MPMK_ERR_INVALID;
    MgState& m = ctx->mg;
    if (m.world < 1) MPMK_FAIL(ctx, MPMK_ERR_STATE, "mpmk_mg_set_slab: call mpmk_mg_init first");
    const int n = ctx->n_grid;
    if (ghost < 0 || x_lo < 0 || x_hi > n || x_hi - x_lo < 2 * ghost + 2)
        MPMK_FAIL(ctx, MPMK_ERR_INVALID, "slab [%d,%d) too thin for ghost=%d (needs >= 2*ghost+2 planes)", x_lo, x_hi, ghost);
    if (y_lo < 0 || y_hi > n || y_hi <= y_lo) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "bad y range [%d,%d)", y_lo, y_hi);
    m.x_lo = x_lo; m.x_hi = x_hi; m.ghost = ghost; m.y_lo = y_lo; m.y_hi = y_hi;
    // shared plane ranges, clamped to the grid; empty at the outer ranks
    m.lo0 = m.lo1 = m.hi0 = m.hi1 = 0;
    if (m.rank > 0) { m.lo0 = max(0, x_lo - ghost); m.lo1 = min(n, x_lo + ghost + 2); }
    if (m.rank < m.world - 1) { m.hi0 = max(0, x_hi - ghost); m.hi1 = min(n, x_hi + ghost + 2); }
    if ((m.lo1 > m.lo0 && !recv_lo_dev) || (m.hi1 > m.hi0 && !recv_hi_dev))
        MPMK_FAIL(ctx, MPMK_ERR_INVALID, "mpmk_mg_set_slab: receive buffers missing");
    m.recv_lo = recv_lo_dev; m.recv_hi = recv_hi_dev;
    m.enabled = m.world > 1;
    return 0;
}

// ---- kernels ----------------------------------------------------------------------------
// region = planes [x0, x0+nx) x rows [y0, y0+ny) x all z; `buf` is the dense [nx][ny][n] exchange buffer
__global__ void k_halo_add(float4* __restrict__ gin, const float4* __restrict__ buf, int x0, int nx, int y0, int ny, int n) {
    const int64_t per_plane = (int64_t)ny * n, total = per_plane * nx;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t px = i / per_plane, rem = i - px * per_plane;
        float4* g = gin + (((int64_t)(x0 + px) * n + y0) * n + rem);
        float4 a = *g;
        const float4 b = buf[i];
        a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        *g = a;
    }
}

__global__ void k_region_clear(float4* __restrict__ gin, float4* __restrict__ gout, int x0, int nx, int y0, int ny, int n) {
    const int64_t per_plane = (int64_t)ny * n, total = per_plane * nx;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t px = i / per_plane, rem = i - px * per_plane;
        const int64_t node = ((int64_t)(x0 + px) * n + y0) * n + rem;
        gin[node] = z;
        gout[node] = z;
    }
}

// model.py:124-127 on the shared region (dense; the block lists do not know what neighbours add there)
int mpmk_mg_clear_shared(mpmk_ctx* ctx, cudaStream_t s) {
    const MgState& m = ctx->mg;
    const int r0[2] = {m.lo0, m.hi0}, r1[2] = {m.lo1, m.hi1};
    for (int k = 0; k < 2; ++k) {
        if (r1[k] <= r0[k]) continue;
        k_region_clear<<<ctx->sm_count * 4, 256, 0, s>>>((float4*)ctx->b.grid_in, (float4*)ctx->b.grid_out, r0[k], r1[k] - r0[k],
                                                         m.y_lo, m.y_hi - m.y_lo, ctx->n_grid);
        MPMK_LAUNCH_CHECK(ctx);
    }
    return 0;
}

// symmetric halo sum of the shared region with both neighbours: one send + one receive per plane
// (rows [y_lo, y_hi) of a plane are one contiguous run of (y_hi - y_lo) * n float4), all in one group
int mpmk_mg_halo_sum(mpmk_ctx* ctx, cudaStream_t s) {
    const MgState& m = ctx->mg;
    if (!m.enabled) return 0;
    ncclComm_t comm = (ncclComm_t)m.comm;
    const int n = ctx->n_grid, ny = m.y_hi - m.y_lo;
    const size_t run_f = (size_t)ny * n * 4;  // floats per plane run
    float* gin = (float*)ctx->b.grid_in;
    const int r0[2] = {m.lo0, m.hi0}, r1[2] = {m.lo1, m.hi1};
    const int peer[2] = {m.rank - 1, m.rank + 1};
    float* rbuf[2] = {(float*)m.recv_lo, (float*)m.recv_hi};
    NCCL_CHECK(ctx, g_nccl.GroupStart());
    for (int k = 0; k < 2; ++k)
        for (int x = r0[k]; x < r1[k]; ++x) {
            NCCL_CHECK(ctx, g_nccl.Send(gin + (((size_t)x * n + m.y_lo) * n) * 4, run_f, ncclFloat32, peer[k], comm, s));
            NCCL_CHECK(ctx, g_nccl.Recv(rbuf[k] + (size_t)(x - r0[k]) * run_f, run_f, ncclFloat32, peer[k], comm, s));
        }
    NCCL_CHECK(ctx, g_nccl.GroupEnd());
    for (int k = 0; k < 2; ++k) {
        if (r1[k] <= r0[k]) continue;
        k_halo_add<<<ctx->sm_count * 4, 256, 0, s>>>((float4*)gin, (const float4*)rbuf[k], r0[k], r1[k] - r0[k], m.y_lo, ny, n);
        MPMK_LAUNCH_CHECK(ctx);
    }
    return 0;
}

extern "C" int mpmk_mg_halo_exchange(mpmk_ctx* ctx, void* stream) {
    if (!ctx || !ctx->bound) return MPMK_ERR_STATE;
    return mpmk_mg_halo_sum(ctx, (cudaStream_t)stream);
}
