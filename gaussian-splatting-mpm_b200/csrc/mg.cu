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
// velocities travel back.  The shared planes are updated densely (and reset behind the update, grid.cu
// k_grid_planes); everything else stays on the sparse block lists.
//
// NCCL is reached through dlopen of the library torch already loaded (path passed by the host layer):
// torch.distributed does the rendezvous plumbing (broadcast of the ncclUniqueId), the data path is
// this file's ncclSend/ncclRecv over NVLink.
#include <dlfcn.h>
#include <stdlib.h>
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
    for (int k = 0; k < 2; ++k) if (ctx->mg.peer_block[k]) { cudaIpcCloseMemHandle(ctx->mg.peer_block[k]); ctx->mg.peer_block[k] = nullptr; }
    if (ctx->mg.my_block) { cudaFree(ctx->mg.my_block); ctx->mg.my_block = nullptr; }
    ctx->mg.peer = false;
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
    m.z_lo = 0; m.z_hi = n;   // until mpmk_mg_set_window narrows it
    // shared plane ranges, clamped to the grid; empty at the outer ranks
    m.lo0 = m.lo1 = m.hi0 = m.hi1 = 0;
    if (m.rank > 0) { m.lo0 = max(0, x_lo - ghost); m.lo1 = min(n, x_lo + ghost + 2); }
    if (m.rank < m.world - 1) { m.hi0 = max(0, x_hi - ghost); m.hi1 = min(n, x_hi + ghost + 2); }
    if ((m.lo1 > m.lo0 && !recv_lo_dev) || (m.hi1 > m.hi0 && !recv_hi_dev))
        MPMK_FAIL(ctx, MPMK_ERR_INVALID, "mpmk_mg_set_slab: receive buffers missing");
    m.recv_lo = recv_lo_dev; m.recv_hi = recv_hi_dev;
    m.enabled = m.world > 1;
    ctx->lag_ready = false;
    mpmk_drop_graphs(ctx);   // captured substeps hold the old plane / row ranges
    return 0;
}

// [z_lo, z_hi): nodes along the fastest axis that any rank's stencils can reach before the next migration (the same
// on every rank, like the row range).  The peer-memory halo and the dense update of the shared planes are restricted
// to rows x this window; the NCCL halo path keeps whole rows (contiguous runs) and ignores it.
extern "C" int mpmk_mg_set_window(mpmk_ctx* ctx, int z_lo, int z_hi) {
    if (!ctx) return MPMK_ERR_INVALID;
    MgState& m = ctx->mg;
    if (z_lo < 0 || z_hi > ctx->n_grid || z_hi <= z_lo) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "bad z window [%d,%d)", z_lo, z_hi);
    if (z_lo != m.z_lo || z_hi != m.z_hi) mpmk_drop_graphs(ctx);
    m.z_lo = z_lo; m.z_hi = z_hi;
    return 0;
}

// after a migration: the live count changed, everything else (current set, F_trial validity) stands --
// the host moved the F_trial records together with the particle records
extern "C" int mpmk_mg_set_count(mpmk_ctx* ctx, int64_t n) {
    if (!ctx) return MPMK_ERR_INVALID;
    if (n < 0 || n > ctx->capacity) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "mpmk_mg_set_count: n=%lld capacity=%lld", (long long)n, (long long)ctx->capacity);
    ctx->n = n;
    ctx->coherent = false;
    ctx->lag_ready = false;
    ctx->sorted_valid = false;
    ctx->prepared = false;
    mpmk_drop_graphs(ctx);
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

// ---- peer-memory halo path ---------------------------------------------------------------------------------
// Instead of ncclSend/ncclRecv, every rank PUSHES its partial sums of the shared planes straight into a receive
// buffer in the neighbour's memory (stores over NVLink through a CUDA-IPC mapping), then raises a flag there with a
// system-scope release; the neighbour's add kernel spins on its own flag (acquire) and adds the buffer.  No
// collective, no rendezvous kernel, nothing on the host: a substep costs one small copy kernel per neighbour and
// one add kernel.  Exchange block layout (bytes): [0,256) control words, then buf[side][parity], side 0 = data coming
// from rank-1, side 1 = from rank+1, parity = substep epoch & 1.  Two buffers per side are enough: a sender can only
// be two pushes ahead of the receiver's add if it has already consumed the receiver's push of the substep in
// between, which the receiver issues after that add (stream order).
#define MGP_FLAG_LO 0    // epoch of the last complete push from rank-1 (written by rank-1)
#define MGP_FLAG_HI 1    // ... from rank+1
#define MGP_EPOCH 2      // substeps exchanged so far (local)
#define MGP_DONE_LO 3    // block-done counters of the two push kernels and the add kernel (local)
#define MGP_DONE_HI 4
#define MGP_DONE_ADD 5
#define MGP_CTRL_BYTES 256

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

#define MGP_ERR 6        // epoch at which a wait for a neighbour's push timed out (0 = never); read by the host at the next migration
#define MGP_SPIN_NS 4000000000ull   // a neighbour that has not pushed after 4 s is dead or has skipped a substep

__device__ __forceinline__ unsigned long long mg_now_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// bounded wait for a neighbour's flag to reach `want`: a rank that died or threw between substeps must not hang the
// others inside a kernel (the host could not even reach mpmk_mg_peer_close).  On a timeout the error word is set, the
// add proceeds on whatever the buffer holds, and every rank raises at its next migration (slabs.py).
__device__ __forceinline__ void mg_wait_flag(const uint32_t* flag, uint32_t want, uint32_t* err) {
    if (ld_acquire_sys(flag) >= want) return;
    const unsigned long long t0 = mg_now_ns();
    while (ld_acquire_sys(flag) < want) {
        if (mg_now_ns() - t0 > MGP_SPIN_NS) { *err = want; return; }
    }
}

// region of one side: planes [x0, x0+nx) x rows [y0, y0+ny) x nodes [z0, z0+nz); the exchange buffers are the dense
// [nx][ny][nz] images of it
struct MgRegion { int x0, nx; };

// copy my partial sums of both shared regions into the neighbours' buffers of this epoch's parity (stores over NVLink),
// then -- last block out -- publish epoch+1 in both neighbours' flags
__global__ void __launch_bounds__(256) k_halo_push(const float4* __restrict__ gin, float4* __restrict__ peer_lo, float4* __restrict__ peer_hi,
                                                   size_t buf_f4, uint32_t* flag_lo, uint32_t* flag_hi, uint32_t* my_ctrl, MgRegion lo,
                                                   MgRegion hi, int y0, int ny, int z0, int nz, int n) {
    const uint32_t e = my_ctrl[MGP_EPOCH];
    const size_t par = (size_t)(e & 1u) * buf_f4;
    const int64_t per_plane = (int64_t)ny * nz;
    const int64_t tlo = per_plane * lo.nx, total = tlo + per_plane * hi.nx;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const bool up = i >= tlo;
        const int64_t j = up ? i - tlo : i;
        const int64_t px = j / per_plane, rem = j - px * per_plane;
        const int64_t py = rem / nz, pz = rem - py * nz;
        const float4 v = gin[(((int64_t)((up ? hi.x0 : lo.x0) + px) * n + y0 + py) * n) + z0 + pz];
        (up ? peer_hi : peer_lo)[par + j] = v;
    }
    // One system-scope fence per push, not one per thread.  The block's stores reach thread 0 through the barrier, its
    // device-scope fence orders them before the block counter, and the last block out -- which has observed every other
    // block's counter update -- issues the one fence.acq_rel.sys that, being cumulative, orders ALL of the push's stores
    // before the two flag stores.  (__threadfence_system() is fence.sc.sys: MEMBAR.SC.SYS + ERRBAR + CCTL.IVALL in every
    // warp, and a system-scope fence queues behind whatever the copy engines have in flight towards the host: with a
    // frame's download running beside the substeps, 75 k of them per push cost the 8-GPU frame loop 11 ms per frame.)
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
        if (atomicAdd(&my_ctrl[MGP_DONE_LO], 1u) == gridDim.x - 1) {
            my_ctrl[MGP_DONE_LO] = 0u;
            asm volatile("fence.acq_rel.sys;" ::: "memory");
            if (lo.nx > 0) st_relaxed_sys(flag_lo, e + 1u);
            if (hi.nx > 0) st_relaxed_sys(flag_hi, e + 1u);
        }
    }
}

// wait for both neighbours' pushes of this epoch, add them (a + b here, b + a there: bit-identical sums), advance the epoch
__global__ void __launch_bounds__(256) k_halo_wait_add(float4* __restrict__ gin, const float4* __restrict__ buf_lo,
                                                       const float4* __restrict__ buf_hi, size_t buf_f4, uint32_t* my_ctrl,
                                                       MgRegion lo, MgRegion hi, int y0, int ny, int z0, int nz, int n) {
    const uint32_t e = my_ctrl[MGP_EPOCH];
    if (threadIdx.x == 0) {
        if (lo.nx > 0) mg_wait_flag(&my_ctrl[MGP_FLAG_LO], e + 1u, &my_ctrl[MGP_ERR]);
        if (hi.nx > 0) mg_wait_flag(&my_ctrl[MGP_FLAG_HI], e + 1u, &my_ctrl[MGP_ERR]);
    }
    __syncthreads();
    const int64_t per_plane = (int64_t)ny * nz;
    const int64_t tlo = per_plane * lo.nx, total = tlo + per_plane * hi.nx;
    const float4* blo = buf_lo + (size_t)(e & 1u) * buf_f4;
    const float4* bhi = buf_hi + (size_t)(e & 1u) * buf_f4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const bool up = i >= tlo;
        const int64_t j = up ? i - tlo : i;
        const int64_t px = j / per_plane, rem = j - px * per_plane;
        const int64_t py = rem / nz, pz = rem - py * nz;
        float4* g = gin + ((((int64_t)((up ? hi.x0 : lo.x0) + px) * n + y0 + py) * n) + z0 + pz);
        const float4 b = __ldcg((up ? bhi : blo) + j);   // written by the peer: never through this SM's L1
        float4 a = *g;
        a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        *g = a;
    }
    __syncthreads();
    if (threadIdx.x == 0 && atomicAdd(&my_ctrl[MGP_DONE_ADD], 1u) == gridDim.x - 1) {
        my_ctrl[MGP_DONE_ADD] = 0u;
        my_ctrl[MGP_EPOCH] = e + 1u;   // read by the next substep's kernels (later launches on the stream)
    }
}

// ---- the same exchange without any fence: every 8 bytes carry their own flag ("LL" protocol) ---------------------
// A node travels as two 16-byte stores {p_x, f, p_y, f} {p_z, f, m, f} with f = epoch + 1 (never 0, never a value an
// earlier use of the buffer left behind: epochs only grow).  NVLink delivers aligned 8-byte words whole, so a receiver
// that reads a word with the right flag has its payload: no flag word per push, no block counter, and above all no
// system-scope fence on either side.  A fence.sys waits for everything the GPU has in flight towards the system, which
// includes a copy engine's bulk download to the host: with 8 GPUs streaming results out over a shared PCIe root, the one
// fence per push that the flag protocol needs cost 0.15 ms per substep.  Twice the bytes over NVLink (2 x 1.2 MB per
// neighbour and substep) is nothing beside that.
__device__ __forceinline__ void st_volatile_v4(uint4* p, uint4 v) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_volatile_v4(const uint4* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(256) k_halo_push_ll(const float4* __restrict__ gin, uint4* __restrict__ peer_lo, uint4* __restrict__ peer_hi,
                                                      size_t buf_u4, const uint32_t* my_ctrl, MgRegion lo, MgRegion hi, int y0, int ny,
                                                      int z0, int nz, int n) {
    const uint32_t e = my_ctrl[MGP_EPOCH];
    const uint32_t f = e + 1u;
    const size_t par = (size_t)(e & 1u) * buf_u4;
    const int64_t per_plane = (int64_t)ny * nz;
    const int64_t tlo = per_plane * lo.nx, total = tlo + per_plane * hi.nx;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const bool up = i >= tlo;
        const int64_t j = up ? i - tlo : i;
        const int64_t px = j / per_plane, rem = j - px * per_plane;
        const int64_t py = rem / nz, pz = rem - py * nz;
        const float4 v = gin[(((int64_t)((up ? hi.x0 : lo.x0) + px) * n + y0 + py) * n) + z0 + pz];
        uint4* dst = (up ? peer_hi : peer_lo) + par + 2 * j;
        st_volatile_v4(dst, make_uint4(__float_as_uint(v.x), f, __float_as_uint(v.y), f));
        st_volatile_v4(dst + 1, make_uint4(__float_as_uint(v.z), f, __float_as_uint(v.w), f));
    }
}

// add both neighbours' pushes of this epoch as their words arrive (a + b here, b + a there: bit-identical sums), advance
// the epoch.  A word that has not arrived after MGP_SPIN_NS sets the error word (see mg_wait_flag) and counts as zero.
__global__ void __launch_bounds__(256) k_halo_wait_add_ll(float4* __restrict__ gin, const uint4* __restrict__ buf_lo,
                                                          const uint4* __restrict__ buf_hi, size_t buf_u4, uint32_t* my_ctrl, MgRegion lo,
                                                          MgRegion hi, int y0, int ny, int z0, int nz, int n) {
    const uint32_t e = my_ctrl[MGP_EPOCH];
    const uint32_t f = e + 1u;
    const int64_t per_plane = (int64_t)ny * nz;
    const int64_t tlo = per_plane * lo.nx, total = tlo + per_plane * hi.nx;
    const uint4* blo = buf_lo + (size_t)(e & 1u) * buf_u4;
    const uint4* bhi = buf_hi + (size_t)(e & 1u) * buf_u4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const bool up = i >= tlo;
        const int64_t j = up ? i - tlo : i;
        const int64_t px = j / per_plane, rem = j - px * per_plane;
        const int64_t py = rem / nz, pz = rem - py * nz;
        float4* g = gin + ((((int64_t)((up ? hi.x0 : lo.x0) + px) * n + y0 + py) * n) + z0 + pz);
        const uint4* src = (up ? bhi : blo) + 2 * j;
        uint4 w0 = ld_volatile_v4(src), w1 = ld_volatile_v4(src + 1);
        if (w0.y != f || w0.w != f || w1.y != f || w1.w != f) {
            const unsigned long long t0 = mg_now_ns();
            for (;;) {
                w0 = ld_volatile_v4(src); w1 = ld_volatile_v4(src + 1);
                if (w0.y == f && w0.w == f && w1.y == f && w1.w == f) break;
                if (mg_now_ns() - t0 > MGP_SPIN_NS) { my_ctrl[MGP_ERR] = f; w0.x = w0.z = w1.x = w1.z = 0u; break; }
            }
        }
        float4 a = *g;
        a.x += __uint_as_float(w0.x); a.y += __uint_as_float(w0.z); a.z += __uint_as_float(w1.x); a.w += __uint_as_float(w1.z);
        *g = a;
    }
    __syncthreads();
    if (threadIdx.x == 0 && atomicAdd(&my_ctrl[MGP_DONE_ADD], 1u) == gridDim.x - 1) {
        my_ctrl[MGP_DONE_ADD] = 0u;
        my_ctrl[MGP_EPOCH] = e + 1u;   // read by the next substep's kernels (later launches on the stream)
    }
}

static int halo_sum_peer(mpmk_ctx* ctx, cudaStream_t s) {
    const MgState& m = ctx->mg;
    const int n = ctx->n_grid, ny = m.y_hi - m.y_lo, nz = m.z_hi - m.z_lo;
    const size_t buf_f4 = m.buf_bytes / 16;
    if ((size_t)(2 * m.ghost + 2) * ny * nz > buf_f4) MPMK_FAIL(ctx, MPMK_ERR_STATE, "peer exchange buffers too small for the shared region");
    uint32_t* my_ctrl = (uint32_t*)m.my_block;
    float4* gin = (float4*)ctx->b.grid_in;
    const MgRegion lo = {m.lo0, m.lo1 - m.lo0}, hi = {m.hi0, m.hi1 - m.hi0};
    // every buffer is a slot of 2 x buf_bytes: the dense float4 image under the flag protocol, 32-byte flagged nodes under LL;
    // side s, parity p at slot 2 s + p
    const size_t slot = 2 * m.buf_bytes;
    static const bool ll = !(getenv("MPMK_MG_LL") && atoi(getenv("MPMK_MG_LL")) == 0);
    if (ll) {
        // my data for rank-1 lands in ITS side-1 slots (I am its rank+1); for rank+1: side 0
        uint4* pl = lo.nx > 0 ? (uint4*)(m.peer_block[0] + MGP_CTRL_BYTES + 2 * slot) : nullptr;
        uint4* ph = hi.nx > 0 ? (uint4*)(m.peer_block[1] + MGP_CTRL_BYTES) : nullptr;
        k_halo_push_ll<<<ctx->sm_count * 2, 256, 0, s>>>(gin, pl, ph, slot / 16, my_ctrl, lo, hi, m.y_lo, ny, m.z_lo, nz, n);
        MPMK_LAUNCH_CHECK(ctx);
        const uint4* ml = (const uint4*)(m.my_block + MGP_CTRL_BYTES);
        const uint4* mh = (const uint4*)(m.my_block + MGP_CTRL_BYTES + 2 * slot);
        k_halo_wait_add_ll<<<ctx->sm_count * 2, 256, 0, s>>>(gin, ml, mh, slot / 16, my_ctrl, lo, hi, m.y_lo, ny, m.z_lo, nz, n);
        MPMK_LAUNCH_CHECK(ctx);
        return 0;
    }
    // my data for rank-1 lands in ITS side-1 buffers (I am its rank+1) and raises its FLAG_HI; for rank+1: side 0, FLAG_LO
    float4* pb_lo = lo.nx > 0 ? (float4*)(m.peer_block[0] + MGP_CTRL_BYTES + 2 * slot) : nullptr;
    float4* pb_hi = hi.nx > 0 ? (float4*)(m.peer_block[1] + MGP_CTRL_BYTES) : nullptr;
    uint32_t* fl_lo = lo.nx > 0 ? (uint32_t*)m.peer_block[0] + MGP_FLAG_HI : nullptr;
    uint32_t* fl_hi = hi.nx > 0 ? (uint32_t*)m.peer_block[1] + MGP_FLAG_LO : nullptr;
    k_halo_push<<<ctx->sm_count * 2, 256, 0, s>>>(gin, pb_lo, pb_hi, slot / 16, fl_lo, fl_hi, my_ctrl, lo, hi, m.y_lo, ny, m.z_lo, nz, n);
    MPMK_LAUNCH_CHECK(ctx);
    const float4* my_lo = (const float4*)(m.my_block + MGP_CTRL_BYTES);
    const float4* my_hi = (const float4*)(m.my_block + MGP_CTRL_BYTES + 2 * slot);
    k_halo_wait_add<<<ctx->sm_count * 2, 256, 0, s>>>(gin, my_lo, my_hi, slot / 16, my_ctrl, lo, hi, m.y_lo, ny, m.z_lo, nz, n);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}

// epoch at which this rank gave up waiting for a neighbour's halo push (0: never).  Synchronises `stream`.
extern "C" int mpmk_mg_peer_error(mpmk_ctx* ctx, uint32_t* epoch_out, void* stream) {
    if (!ctx || !epoch_out) return MPMK_ERR_INVALID;
    *epoch_out = 0;
    if (!ctx->mg.peer || !ctx->mg.my_block) return 0;
    MPMK_CHECK(ctx, cudaMemcpyAsync(epoch_out, (uint32_t*)ctx->mg.my_block + MGP_ERR, 4, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    MPMK_CHECK(ctx, cudaStreamSynchronize((cudaStream_t)stream));
    return 0;
}

// allocate this rank's exchange block and export it; handle_out receives the 64-byte cudaIpcMemHandle_t.
// Layout: control words, 2 sides x 2 parities of halo slots (a slot = 2 x buf_bytes, halo_sum_peer), then 2 sides x 2
// parities of message slots of msg_bytes (mpmk_mg_ll_send / _recv).
extern "C" int mpmk_mg_peer_alloc_ex(mpmk_ctx* ctx, int64_t buf_bytes, int64_t msg_bytes, void* handle_out, int handle_bytes) {
    if (!ctx || !handle_out || handle_bytes < (int)sizeof(cudaIpcMemHandle_t) || buf_bytes <= 0 || (buf_bytes & 15) || msg_bytes < 0 ||
        (msg_bytes & 15))
        return MPMK_ERR_INVALID;
    MgState& m = ctx->mg;
    if (m.my_block) MPMK_FAIL(ctx, MPMK_ERR_STATE, "mpmk_mg_peer_alloc: already allocated");
    MPMK_CHECK(ctx, cudaSetDevice(ctx->device));
    const size_t total = MGP_CTRL_BYTES + 4 * 2 * (size_t)buf_bytes + 4 * (size_t)msg_bytes;
    m.msg_bytes = (size_t)msg_bytes;
    MPMK_CHECK(ctx, cudaMalloc((void**)&m.my_block, total));
    MPMK_CHECK(ctx, cudaMemset(m.my_block, 0, total));
    MPMK_CHECK(ctx, cudaDeviceSynchronize());
    cudaIpcMemHandle_t h;
    MPMK_CHECK(ctx, cudaIpcGetMemHandle(&h, m.my_block));
    memcpy(handle_out, &h, sizeof(h));
    m.buf_bytes = (size_t)buf_bytes;
    return 0;
}

extern "C" int mpmk_mg_peer_alloc(mpmk_ctx* ctx, int64_t buf_bytes, void* handle_out, int handle_bytes) {
    return mpmk_mg_peer_alloc_ex(ctx, buf_bytes, 0, handle_out, handle_bytes);
}

// ---- flagged messages between neighbours (the migration payload) ---------------------------------------------------
// The LL encoding of the halo on an arbitrary word array: words (2 i, 2 i + 1) travel as one 16-byte store
// {w0, tag, w1, tag} into the neighbour's message slot [side][tag & 1]; the receiver spins per 16 bytes until both tags
// match.  No NCCL, no fence, no rendezvous: under a congested host link (8 GPUs downloading a frame each) ncclSend /
// ncclRecv of the migration records took ~6 ms per migration -- their chunk hand-shakes are system-scope fences -- against
// 0.15 ms for the same records without contention.  `tag` must differ from the tag of the previous message through the
// same slot and must not be 0; both sides pass the same value (the migration count).
__global__ void __launch_bounds__(256) k_ll_send(uint4* __restrict__ dst, const uint32_t* __restrict__ src, int64_t n_words, uint32_t tag) {
    const int64_t pairs = (n_words + 1) / 2;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < pairs; i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t w0 = src[2 * i], w1 = (2 * i + 1 < n_words) ? src[2 * i + 1] : 0u;
        st_volatile_v4(dst + i, make_uint4(w0, tag, w1, tag));
    }
}

__global__ void __launch_bounds__(256) k_ll_recv(uint32_t* __restrict__ dst, const uint4* __restrict__ src, int64_t n_words, uint32_t tag,
                                                 uint32_t* err) {
    const int64_t pairs = (n_words + 1) / 2;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < pairs; i += (int64_t)gridDim.x * blockDim.x) {
        uint4 w = ld_volatile_v4(src + i);
        if (w.y != tag || w.w != tag) {
            const unsigned long long t0 = mg_now_ns();
            for (;;) {
                w = ld_volatile_v4(src + i);
                if (w.y == tag && w.w == tag) break;
                if (mg_now_ns() - t0 > MGP_SPIN_NS) { *err = 0x80000000u | tag; w.x = w.z = 0u; break; }
            }
        }
        dst[2 * i] = w.x;
        if (2 * i + 1 < n_words) dst[2 * i + 1] = w.z;
    }
}

static unsigned char* mg_msg_slot(const MgState& m, unsigned char* block, int side, uint32_t tag) {
    return block + MGP_CTRL_BYTES + 4 * 2 * m.buf_bytes + (size_t)(2 * side + (int)(tag & 1u)) * m.msg_bytes;
}

// side 0: to rank-1 (lands in ITS side-1 slots), side 1: to rank+1 (its side-0 slots)
extern "C" int mpmk_mg_ll_send(mpmk_ctx* ctx, int side, const void* src_dev, int64_t n_words, uint32_t tag, void* stream) {
    if (!ctx || side < 0 || side > 1 || n_words < 0 || tag == 0u) return MPMK_ERR_INVALID;
    const MgState& m = ctx->mg;
    if (!m.peer || !m.peer_block[side]) MPMK_FAIL(ctx, MPMK_ERR_STATE, "mpmk_mg_ll_send: no mapped neighbour on that side");
    if ((size_t)((n_words + 1) / 2) * 16 > m.msg_bytes) MPMK_FAIL(ctx, MPMK_ERR_STATE, "mpmk_mg_ll_send: message larger than the slot");
    if (n_words == 0) return 0;
    if (!src_dev) return MPMK_ERR_INVALID;
    const int64_t pairs = (n_words + 1) / 2;
    const int grid = (int)(pairs / 256 + 1 < (int64_t)ctx->sm_count * 4 ? pairs / 256 + 1 : (int64_t)ctx->sm_count * 4);
    k_ll_send<<<grid, 256, 0, (cudaStream_t)stream>>>((uint4*)mg_msg_slot(m, m.peer_block[side], 1 - side, tag), (const uint32_t*)src_dev, n_words, tag);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}

// side 0: from rank-1, side 1: from rank+1
extern "C" int mpmk_mg_ll_recv(mpmk_ctx* ctx, int side, void* dst_dev, int64_t n_words, uint32_t tag, void* stream) {
    if (!ctx || side < 0 || side > 1 || n_words < 0 || tag == 0u) return MPMK_ERR_INVALID;
    const MgState& m = ctx->mg;
    if (!m.peer || !m.my_block) MPMK_FAIL(ctx, MPMK_ERR_STATE, "mpmk_mg_ll_recv: peer-memory path is not open");
    if ((size_t)((n_words + 1) / 2) * 16 > m.msg_bytes) MPMK_FAIL(ctx, MPMK_ERR_STATE, "mpmk_mg_ll_recv: message larger than the slot");
    if (n_words == 0) return 0;
    if (!dst_dev) return MPMK_ERR_INVALID;
    const int64_t pairs = (n_words + 1) / 2;
    const int grid = (int)(pairs / 256 + 1 < (int64_t)ctx->sm_count * 4 ? pairs / 256 + 1 : (int64_t)ctx->sm_count * 4);
    k_ll_recv<<<grid, 256, 0, (cudaStream_t)stream>>>((uint32_t*)dst_dev, (const uint4*)mg_msg_slot(m, m.my_block, side, tag), n_words, tag,
                                                      (uint32_t*)m.my_block + MGP_ERR);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}

// map the neighbours' exchange blocks (handles from their mpmk_mg_peer_alloc; NULL at the outer ranks) and switch
// the halo sum of mpmk_substep(s) to the peer-memory path
extern "C" int mpmk_mg_peer_open(mpmk_ctx* ctx, const void* handle_lo, const void* handle_hi) {
    if (!ctx) return MPMK_ERR_INVALID;
    MgState& m = ctx->mg;
    if (!m.my_block) MPMK_FAIL(ctx, MPMK_ERR_STATE, "mpmk_mg_peer_open: call mpmk_mg_peer_alloc first");
    MPMK_CHECK(ctx, cudaSetDevice(ctx->device));
    const void* hs[2] = {handle_lo, handle_hi};
    for (int k = 0; k < 2; ++k) {
        if (!hs[k]) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, hs[k], sizeof(h));
        void* p = nullptr;
        MPMK_CHECK(ctx, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        m.peer_block[k] = (unsigned char*)p;
    }
    m.peer = true;
    return 0;
}

extern "C" int mpmk_mg_peer_enabled(mpmk_ctx* ctx) { return ctx && ctx->mg.peer ? 1 : 0; }

// back to the NCCL halo path (a neighbour could not map this rank's block, say): unmap, free, forget
extern "C" int mpmk_mg_peer_close(mpmk_ctx* ctx) {
    if (!ctx) return MPMK_ERR_INVALID;
    MgState& m = ctx->mg;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    for (int k = 0; k < 2; ++k) if (m.peer_block[k]) { cudaIpcCloseMemHandle(m.peer_block[k]); m.peer_block[k] = nullptr; }
    if (m.my_block) { cudaFree(m.my_block); m.my_block = nullptr; }
    m.peer = false;
    mpmk_drop_graphs(ctx);
    cudaGetLastError();
    return 0;
}

// symmetric halo sum of the shared region with both neighbours: one send + one receive per plane
// (rows [y_lo, y_hi) of a plane are one contiguous run of (y_hi - y_lo) * n float4), all in one group
int mpmk_mg_halo_sum(mpmk_ctx* ctx, cudaStream_t s) {
    const MgState& m = ctx->mg;
    if (!m.enabled) return 0;
    static const int skip = getenv("MPMK_MG_SKIP") ? atoi(getenv("MPMK_MG_SKIP")) : 0;   // timing experiments only: wrong physics
    if (skip >= 2) return 0;
    if (m.peer) return halo_sum_peer(ctx, s);
    ncclComm_t comm = (ncclComm_t)m.comm;
    const int n = ctx->n_grid, ny = m.y_hi - m.y_lo;
    const size_t run_f = (size_t)ny * n * 4;  // floats per plane run
    float* gin = (float*)ctx->b.grid_in;
    const int r0[2] = {m.lo0, m.hi0}, r1[2] = {m.lo1, m.hi1};
    const int peer[2] = {m.rank - 1, m.rank + 1};
    float* rbuf[2] = {(float*)m.recv_lo, (float*)m.recv_hi};
    if (skip < 1) {
    NCCL_CHECK(ctx, g_nccl.GroupStart());
    for (int k = 0; k < 2; ++k)
        for (int x = r0[k]; x < r1[k]; ++x) {
            NCCL_CHECK(ctx, g_nccl.Send(gin + (((size_t)x * n + m.y_lo) * n) * 4, run_f, ncclFloat32, peer[k], comm, s));
            NCCL_CHECK(ctx, g_nccl.Recv(rbuf[k] + (size_t)(x - r0[k]) * run_f, run_f, ncclFloat32, peer[k], comm, s));
        }
    NCCL_CHECK(ctx, g_nccl.GroupEnd());
    }
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

// ---- migration helpers (the protocol itself is mpm_solver/slabs.py) ------------------------------------------
// stats layout (int32[12]): {leavers down, leavers up, beyond ghost, -min row, max row, max(|v_x|,|v_y|,|v_z|) (float bits),
//                            -min z, max z, halo timeout epoch (peer path), 0, 0, 0}
struct MgRecordPtrs {
    const float4* src[16];
    float4* dst[16];
    int n_arrays;
};

__global__ void k_mg_stats_init(int* stats, int n_grid, const uint32_t* ctrl) {
    const int t = threadIdx.x;
    if (t < 12) stats[t] = (t == 3 || t == 6) ? -n_grid : ((t == 4 || t == 7) ? -1 : (t == 8 && ctrl ? (int)ctrl[MGP_ERR] : 0));
}

__global__ void __launch_bounds__(256) k_mg_scan(const uint32_t* __restrict__ keys, const float4* __restrict__ vv, int64_t n, int nt,
                                                 int x_lo, int x_hi, int ghost, int* __restrict__ stats,
                                                 int* __restrict__ idx_lo, int* __restrict__ idx_hi) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned lane = threadIdx.x & 31;
    int by = -1, nby = -(1 << 30), bz = -1, nbz = -(1 << 30), far = 0;
    float av = 0.f;
    bool lo = false, hi = false;
    if (p < n) {
        const uint32_t k = keys[p];
        const uint32_t tile = k >> 6;
        const int bx = (int)(tile / (uint32_t)(nt * nt)) * 4 + (int)((k >> 4) & 3u);
        by = (int)((tile / (uint32_t)nt) % (uint32_t)nt) * 4 + (int)((k >> 2) & 3u);
        nby = -by;
        bz = (int)(tile % (uint32_t)nt) * 4 + (int)(k & 3u);
        nbz = -bz;
        lo = bx < x_lo; hi = bx >= x_hi;
        far = (bx < x_lo - ghost || bx >= x_hi + ghost) ? 1 : 0;
        const float4 v = vv[p];
        av = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fabsf(v.z));   // drift along the slab axis and inside the exchange window
    }
    // leavers: one atomic per warp and direction, order inside the lists is irrelevant
    const unsigned mlo = __ballot_sync(0xffffffffu, lo), mhi = __ballot_sync(0xffffffffu, hi);
    if (mlo) {
        int base = 0;
        if (lane == 0) base = atomicAdd(&stats[0], __popc(mlo));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (lo) idx_lo[base + __popc(mlo & ((1u << lane) - 1u))] = (int)p;
    }
    if (mhi) {
        int base = 0;
        if (lane == 0) base = atomicAdd(&stats[1], __popc(mhi));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (hi) idx_hi[base + __popc(mhi & ((1u << lane) - 1u))] = (int)p;
    }
    far = __reduce_add_sync(0xffffffffu, far);
    nby = __reduce_max_sync(0xffffffffu, nby);
    by = __reduce_max_sync(0xffffffffu, by);
    nbz = __reduce_max_sync(0xffffffffu, nbz);
    bz = __reduce_max_sync(0xffffffffu, bz);
    int avb = __reduce_max_sync(0xffffffffu, __float_as_int(av));   // non-negative floats order like their bits
    if (lane == 0) {
        if (far) atomicAdd(&stats[2], far);
        atomicMax(&stats[3], nby);
        atomicMax(&stats[4], by);
        atomicMax(&stats[5], avb);
        atomicMax(&stats[6], nbz);
        atomicMax(&stats[7], bz);
    }
}

// rows[count][n_arrays*4 + 1]: the 16-byte records of every array, then the key
__global__ void __launch_bounds__(256) k_mg_pack(MgRecordPtrs P, const uint32_t* __restrict__ keys, const int* __restrict__ idx,
                                                 int64_t count, int* __restrict__ rows) {
    const int W = P.n_arrays * 4 + 1;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t r = t / (P.n_arrays + 1);
    const int a = (int)(t % (P.n_arrays + 1));
    if (r >= count) return;
    const int slot = idx[r];
    if (a < P.n_arrays) {
        const float4 v = P.src[a][slot];
        int* o = rows + r * W + 4 * a;
        o[0] = __float_as_int(v.x); o[1] = __float_as_int(v.y); o[2] = __float_as_int(v.z); o[3] = __float_as_int(v.w);
    } else {
        rows[r * W + W - 1] = (int)keys[slot];
    }
}

__global__ void __launch_bounds__(256) k_mg_unpack(MgRecordPtrs P, uint32_t* __restrict__ keys, const int* __restrict__ idx,
                                                   int64_t idx_offset, int64_t count, const int* __restrict__ rows) {
    // destination slot of row r: idx[r] when an index list is given, else idx_offset + r (append)
    const int W = P.n_arrays * 4 + 1;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t r = t / (P.n_arrays + 1);
    const int a = (int)(t % (P.n_arrays + 1));
    if (r >= count) return;
    const int64_t slot = idx ? (int64_t)idx[r] : idx_offset + r;
    if (a < P.n_arrays) {
        const int* i = rows + r * W + 4 * a;
        P.dst[a][slot] = make_float4(__int_as_float(i[0]), __int_as_float(i[1]), __int_as_float(i[2]), __int_as_float(i[3]));
    } else {
        keys[slot] = (uint32_t)rows[r * W + W - 1];
    }
}

// closing the holes a shrinking slab leaves below n_new: pass 1 splits the unfilled holes into those below n_new
// (list) and those in the tail [n_new, n) (flags); pass 2 moves each live tail record into one of the listed holes
__global__ void k_mg_split_holes(const int* __restrict__ holes, int64_t count, int64_t n_new, int* __restrict__ low_holes,
                                 int* __restrict__ counters, unsigned char* __restrict__ tail_flag) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    const int h = holes[t];
    if (h < n_new) low_holes[atomicAdd(&counters[0], 1)] = h;
    else tail_flag[h - n_new] = 1;
}

__global__ void k_mg_move_tail(MgRecordPtrs P, uint32_t* __restrict__ keys, int64_t n_new, int64_t tail_len,
                               const int* __restrict__ low_holes, int* __restrict__ counters,
                               const unsigned char* __restrict__ tail_flag) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= tail_len || tail_flag[t]) return;
    const int dst = low_holes[atomicAdd(&counters[1], 1)];
    const int64_t src = n_new + t;
    for (int a = 0; a < P.n_arrays; ++a) P.dst[a][dst] = P.src[a][src];
    keys[dst] = keys[src];
}

static MgRecordPtrs record_ptrs(mpmk_ctx* ctx) {
    // order = mpm_solver/slabs.py RECORD_ARRAYS
    const mpmk_particle_set& s = ctx->b.set[ctx->cur];
    void* arr[15] = {s.xm, s.vv, s.c0, s.c1, s.cs, s.s1, s.f0, s.f1, s.f2, s.meta, s.cv0, s.cv1, ctx->b.ft0, ctx->b.ft1, ctx->b.ft2};
    MgRecordPtrs P;
    P.n_arrays = 15;
    for (int i = 0; i < 15; ++i) { P.src[i] = (const float4*)arr[i]; P.dst[i] = (float4*)arr[i]; }
    P.src[15] = nullptr; P.dst[15] = nullptr;
    return P;
}

#define REQUIRE_SLAB(ctx)                                                                          \
    do {                                                                                           \
        if (!(ctx) || !(ctx)->bound) return MPMK_ERR_STATE;                                        \
        if ((ctx)->mg.world < 1) MPMK_FAIL(ctx, MPMK_ERR_STATE, "call mpmk_mg_init / mpmk_mg_set_slab first"); \
    } while (0)

extern "C" int mpmk_mg_scan(mpmk_ctx* ctx, void* stats_dev, void* idx_lo_dev, void* idx_hi_dev, void* stream) {
    REQUIRE_SLAB(ctx);
    cudaStream_t s = (cudaStream_t)stream;
    const MgState& m = ctx->mg;
    k_mg_stats_init<<<1, 32, 0, s>>>((int*)stats_dev, ctx->n_grid, (m.peer && m.my_block) ? (const uint32_t*)m.my_block : nullptr);
    MPMK_LAUNCH_CHECK(ctx);
    if (ctx->n > 0) {
        k_mg_scan<<<div_up(ctx->n, 256), 256, 0, s>>>((const uint32_t*)ctx->b.keys, (const float4*)ctx->b.set[ctx->cur].vv, ctx->n,
                                                      ctx->nt, m.x_lo, m.x_hi, m.ghost, (int*)stats_dev, (int*)idx_lo_dev,
                                                      (int*)idx_hi_dev);
        MPMK_LAUNCH_CHECK(ctx);
    }
    return 0;
}

extern "C" int mpmk_mg_pack(mpmk_ctx* ctx, const void* idx_dev, int64_t count, void* rows_dev, void* stream) {
    REQUIRE_SLAB(ctx);
    if (count <= 0) return 0;
    const MgRecordPtrs P = record_ptrs(ctx);
    k_mg_pack<<<div_up(count * (P.n_arrays + 1), 256), 256, 0, (cudaStream_t)stream>>>(P, (const uint32_t*)ctx->b.keys,
                                                                                         (const int*)idx_dev, count, (int*)rows_dev);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}

extern "C" int mpmk_mg_unpack(mpmk_ctx* ctx, const void* idx_dev, int64_t first_slot, int64_t count, const void* rows_dev,
                              void* stream) {
    REQUIRE_SLAB(ctx);
    if (count <= 0) return 0;
    if (!idx_dev && (first_slot < 0 || first_slot + count > ctx->capacity)) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "mpmk_mg_unpack: beyond capacity");
    const MgRecordPtrs P = record_ptrs(ctx);
    k_mg_unpack<<<div_up(count * (P.n_arrays + 1), 256), 256, 0, (cudaStream_t)stream>>>(P, (uint32_t*)ctx->b.keys, (const int*)idx_dev,
                                                                                           first_slot, count, (const int*)rows_dev);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}

extern "C" int mpmk_mg_close_holes(mpmk_ctx* ctx, const void* holes_dev, int64_t n_holes, int64_t n_new, void* scratch_dev,
                                   int64_t scratch_bytes, void* stream) {
    REQUIRE_SLAB(ctx);
    const int64_t tail_len = ctx->n - n_new;
    if (n_holes <= 0 || tail_len <= 0) return 0;
    // scratch: int counters[2] | int low_holes[n_holes] | uchar tail_flag[tail_len]
    const int64_t need = 8 + 4 * n_holes + tail_len;
    if (scratch_bytes < need) MPMK_FAIL(ctx, MPMK_ERR_INVALID, "mpmk_mg_close_holes: scratch %lld < %lld bytes", (long long)scratch_bytes, (long long)need);
    cudaStream_t s = (cudaStream_t)stream;
    int* counters = (int*)scratch_dev;
    int* low = counters + 2;
    unsigned char* flag = (unsigned char*)(low + n_holes);
    MPMK_CHECK(ctx, cudaMemsetAsync(scratch_dev, 0, (size_t)need, s));
    k_mg_split_holes<<<div_up(n_holes, 256), 256, 0, s>>>((const int*)holes_dev, n_holes, n_new, low, counters, flag);
    MPMK_LAUNCH_CHECK(ctx);
    const MgRecordPtrs P = record_ptrs(ctx);
    k_mg_move_tail<<<div_up(tail_len, 256), 256, 0, s>>>(P, (uint32_t*)ctx->b.keys, n_new, tail_len, low, counters, flag);
    MPMK_LAUNCH_CHECK(ctx);
    return 0;
}

// ---- small read-back past the copy engines (mpmk.h: mpmk_read_small) ----------------------------------------
__global__ void k_store_host(uint32_t* __restrict__ dst, const uint32_t* __restrict__ src, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[i];
    __threadfence_system();
}

extern "C" int mpmk_read_small(mpmk_ctx* ctx, void* dst_pinned, const void* src_dev, int n_words, void* stream) {
    if (!ctx || !dst_pinned || !src_dev || n_words < 0) return MPMK_ERR_INVALID;
    if (n_words == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    k_store_host<<<div_up(n_words, 256), 256, 0, s>>>((uint32_t*)dst_pinned, (const uint32_t*)src_dev, n_words);
    MPMK_LAUNCH_CHECK(ctx);
    MPMK_CHECK(ctx, cudaStreamSynchronize(s));
    return 0;
}

