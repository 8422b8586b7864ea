/*
 * mpmk.h -- C ABI of libmpmk.so: the B200 (sm_100a) MPM substep kernels.
 *
 * This is the drop-in boundary for the hot path `MPM_Simulator.p2g2p()` of
 * ranrandy/gaussian-splatting-mpm (mpm_solver/solver.py:27-52).  The reference
 * has no FFI of its own: its kernels are Taichi `@ti.kernel`s JIT-compiled by
 * taichi==1.5.0.  Each entry point below therefore names the reference kernel
 * (file:line, relative to the reference root) whose work it replaces.
 *
 * Conventions
 *   - plain pointers and sizes only; every `dev` pointer is a CUDA device
 *     pointer owned by the caller (torch tensors in the Python host layer);
 *     the library borrows them and owns only its scratch (sort buffers,
 *     histograms, work lists).
 *   - every launch goes to the caller's stream (`stream` = cudaStream_t cast to
 *     void*); no entry point synchronises unless its comment says so.
 *   - return value 0 = success, otherwise a cudaError_t-compatible code (or
 *     MPMK_ERR_*); the message is available from mpmk_last_error().
 *   - a context is bound to one device and is not thread-safe.
 */
#ifndef MPMK_H
#define MPMK_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPMK_VERSION 1

#if defined(__GNUC__)
#define MPMK_API __attribute__((visibility("default")))
#else
#define MPMK_API
#endif

#define MPMK_ERR_INVALID 10001 /* bad argument                      */
#define MPMK_ERR_STATE 10002   /* call order / missing bind         */

/* material ids: mpm_solver/utils.py:5-10 (0..3); 4 and 5 are additions of
 * this build, parity unpinned                               */
#define MPMK_MAT_JELLY 0
#define MPMK_MAT_METAL 1
#define MPMK_MAT_SAND 2
#define MPMK_MAT_FOAM 3
#define MPMK_MAT_FLUID 4
#define MPMK_MAT_PLASTICINE 5

/* behaviour flags (mpmk_params.flags) */
#define MPMK_STRICT_Q1 1u  /* utils.py:37: jelly stress is never evaluated -> 0       */
#define MPMK_STRICT_Q3 2u  /* constitutive_models.py:256: foam F through Hadamard     */
#define MPMK_WRITE_FTRIAL 4u /* g2p also stores F_trial (needed by export for plastic)  */
#define MPMK_TRACK_COV 8u  /* g2p applies update_cov every substep (utils.py:187-215) */
#define MPMK_SORT_MERGE 64u /* take the merge path whenever it is valid, also below the size where it pays (tests) */
#define MPMK_SORT_FULL 16u /* always run the full radix sort (no merge of the few particles that changed cell) */
#define MPMK_NO_LAG 128u   /* mpmk_substeps: keep the sort on the critical path in every substep (see api.cu, lagged substeps) */

typedef struct mpmk_ctx mpmk_ctx;

/* One set of particle arrays in the packed SoA the kernels use; each array
 * holds `capacity` 16-byte records, record r belongs to the particle in
 * physical slot r.  Two sets are bound (ping-pong): g2p reads the source set
 * through the sort permutation and writes the destination set in sorted
 * order, which re-establishes physical order == sorted order every substep.  */
typedef struct {
    void* xm;   /* float4 {x, y, z, mass}              model.py:85,87           */
    void* vv;   /* float4 {vx, vy, vz, -}              model.py:88              */
    void* c0;   /* float4 {C00, C01, C02, C10}         model.py:91              */
    void* c1;   /* float4 {C11, C12, C20, C21}                                  */
    void* cs;   /* float4 {C22, s00, s01, s02}         stress: model.py:90      */
    void* s1;   /* float4 {s11, s12, s22, vol}         symmetric Kirchhoff stress */
    void* f0;   /* float4 {F00, F01, F02, F10}         model.py:89              */
    void* f1;   /* float4 {F11, F12, F20, F21}                                  */
    void* f2;   /* float4 {F22, mu, lam, yield_stress} model.py:38-39,54        */
    void* meta; /* int4   {orig_id, material, vol(bits), 0}                     */
    void* cv0;  /* float4 {cov0..cov3}  only read/written with MPMK_TRACK_COV   */
    void* cv1;  /* float4 {cov4, cov5, -, -}                                    */
} mpmk_particle_set;

typedef struct {
    mpmk_particle_set set[2];
    void* ft0;      /* float4 x3: F_trial of the last g2p, sorted order (model.py:98); ft2 = {Ft22, yield_stress before the return map, 0, 0} */
    void* ft1;
    void* ft2;
    void* keys;     /* uint32[capacity]: cell key of the particle in each physical slot */
    void* grid_in;  /* float4[n^3] {px, py, pz, mass}  grid_v_in + grid_mass, model.py:118-120 */
    void* grid_out; /* float4[n^3] {vx, vy, vz, 0}     grid_v_out, model.py:121        */
} mpmk_bindings;

typedef struct {
    float gravity[3];        /* model.py:32                                   */
    float alpha;             /* model.py:49-51                                */
    float plastic_viscosity; /* model.py:58                                   */
    float xi;                /* model.py:57                                   */
    float softening;         /* model.py:59                                   */
    int32_t hardening;       /* model.py:56                                   */
    uint32_t flags;          /* MPMK_STRICT_* | MPMK_WRITE_FTRIAL | MPMK_TRACK_COV */
} mpmk_params;

/* grid operations applied after normalisation, in list order (solver.py:41-46) */
#define MPMK_OP_FIXED_CUBE 0 /* boundary_conditions.py:23-27: a=center, b=size          */
#define MPMK_OP_COLLIDER 1   /* collider.py:13-44: a=point, b=unit normal, c=friction   */
typedef struct {
    int32_t type;
    float a[3];
    float b[3];
    float c;
} mpmk_grid_op;

/* particle pre-process: boundary_conditions.py:41-45 ImpulseBC.apply */
typedef struct {
    float center[3];
    float size[3];
    float force[3];
    float substep_dt; /* the BC's own substep_dt, not the call's dt */
} mpmk_impulse;

#define MPMK_MAX_GRID_OPS 16
#define MPMK_MAX_IMPULSES 8

typedef struct {
    int32_t n_ops;
    mpmk_grid_op ops[MPMK_MAX_GRID_OPS];
    int32_t n_impulses;
    mpmk_impulse impulses[MPMK_MAX_IMPULSES];
} mpmk_substep_ops;

/* device-side counters of the last substep (mpmk_read_counters copies them) */
typedef struct {
    uint32_t n_items;         /* p2g work items (tile chunks)                  */
    uint32_t n_active_blocks; /* 4x4x4-node blocks touched by p2g              */
    uint32_t n_tiles;         /* non-empty 4x4x4-cell tiles                    */
    uint32_t n_clamped;       /* particles whose base cell was clamped (Q9)    */
    uint32_t n_crossers;      /* lagged substeps: particles whose cell had left their work item's tile */
} mpmk_counters;

/* ---- lifecycle ---------------------------------------------------------- */
MPMK_API int mpmk_version(void);
/* dx = grid_extent/n_grid and inv_dx = n_grid/grid_extent are computed in double
 * and rounded to f32 as model.py:15-16 does.  Allocates the scratch.          */
MPMK_API int mpmk_create(mpmk_ctx** out, int device, int n_grid, double grid_extent, int64_t capacity);
MPMK_API int mpmk_destroy(mpmk_ctx* ctx);
MPMK_API const char* mpmk_last_error(mpmk_ctx* ctx);
MPMK_API int mpmk_bind(mpmk_ctx* ctx, const mpmk_bindings* b);
MPMK_API int mpmk_set_params(mpmk_ctx* ctx, const mpmk_params* p);
/* number of live particles and which set currently holds them (0 or 1) */
MPMK_API int mpmk_set_count(mpmk_ctx* ctx, int64_t n_particles, int current_set);
MPMK_API int mpmk_current_set(mpmk_ctx* ctx);
MPMK_API int64_t mpmk_count(mpmk_ctx* ctx);

/* ---- init-time kernels --------------------------------------------------- */
/* utils.py:350-362 compute_mu_lam_from_E_nu */
MPMK_API int mpmk_compute_mu_lam(mpmk_ctx* ctx, int64_t n, const void* logE_dev, const void* y_dev, void* mu_dev,
                        void* lam_dev, void* stream);
/* internel_filling/filling.py:12-24; cell_count_dev = int32[n_grid^3] scratch, zeroed by the caller */
MPMK_API int mpmk_particle_volume(mpmk_ctx* ctx, int64_t n, const void* pos_dev /* float[n][3] */, float grid_dx,
                         void* cell_count_dev, void* vol_dev, void* stream);
/* mpmk_particle_volume without a caller-provided scratch and without any allocation.  No counterpart in the reference:
 * added for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  The per-cell counts live in a file-scope device array inside the
 * library (csrc/export.cu: 128^3 int32 = 8 MiB, held by every GPU that loads the library until the process exits), so
 * n_grid must be <= 128: larger grids are rejected with MPMK_ERR_INVALID and a message naming the limit.  The context's
 * device must be current.  The call clears the cells it uses on `stream` and launches the same two kernels as
 * mpmk_particle_volume: same result, bit for bit.  It calls no cudaMalloc / cudaFree / allocating library routine; the
 * first call on a device resolves the array's address (cudaGetSymbolAddress).  Every call on a device shares the array:
 * calls that overlap on the GPU (different streams without an ordering between them) race.                       */
MPMK_API int mpmk_particle_volume_noalloc(mpmk_ctx* ctx, int64_t n, const void* pos_dev /* float[n][3] */, float grid_dx,
                                          void* vol_dev, void* stream);
/* cell keys of the current set from positions (utils.py:94-95 base cell) */
MPMK_API int mpmk_compute_keys(mpmk_ctx* ctx, void* stream);
/* utils.py:14-54 compute_stress_from_F_trial on the current set, in place:
 * F <- return_map(F), stress <- tau(F).  Used after init / from_torch edits. */
MPMK_API int mpmk_constitutive(mpmk_ctx* ctx, float dt, void* stream);
/* The library evaluates compute_stress_from_F_trial for substep s+1 in the epilogue of substep s's g2p, with the
 * per-particle parameters of that moment; the reference evaluates it at the START of substep s+1 (solver.py:36), i.e.
 * with whatever the caller wrote to mu / lam / yield_stress / material / the scalar parameters in between.  After such a
 * write the host layer calls this to redo the update: use_ftrial != 0 takes F_trial from the stored ft0..ft2 (valid after
 * a g2p with MPMK_WRITE_FTRIAL) instead of the already return-mapped F; use_ys_before != 0 takes the yield stress as it
 * was BEFORE that g2p's return map (ft2.y), so hardening / softening is applied once, not twice.                    */
MPMK_API int mpmk_constitutive_ex(mpmk_ctx* ctx, float dt, int use_ftrial, int use_ys_before, void* stream);

/* ---- the substep, stage by stage (solver.py:27-52) ------------------------ */
/* stable sort of (key, slot) by key; results stay in library scratch.  Full LSD radix sort, or -- when the
 * previous g2p left the particles in sorted order -- a radix sort of only the particles whose cell changed,
 * merged with the rest; both produce the permutation of a CPU stable sort, bit for bit                   */
MPMK_API int mpmk_sort(mpmk_ctx* ctx, void* stream);
/* copy the last sort's sorted keys / permutation (uint32[n]) to caller buffers */
MPMK_API int mpmk_get_sort_result(mpmk_ctx* ctx, void* sorted_keys_dev, void* perm_dev, void* stream);
/* which path the last mpmk_sort took (1 = merge of the particles that changed cell, 0 = full radix sort) and how
 * many (key, slot) pairs went through the radix passes; synchronises `stream`                              */
MPMK_API int mpmk_sort_info(mpmk_ctx* ctx, int* coherent_out, int64_t* movers_out, void* stream);
/* model.py:124-127 reset_grid_state, restricted to the blocks the previous
 * substep touched; then tile work list + active-block list of this substep   */
MPMK_API int mpmk_prepare_grid(mpmk_ctx* ctx, void* stream);
/* boundary_conditions.py:41-45 (impulses) + utils.py:90-134 p2g */
MPMK_API int mpmk_p2g(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, void* stream);
/* utils.py:178-183 + boundary_conditions.py:23-27 + collider.py:13-44, fused */
MPMK_API int mpmk_grid_update(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, void* stream);
/* utils.py:219-282 g2p + :14-54 compute_stress_from_F_trial (for the next
 * substep) + new cell keys; writes the other set in sorted order and flips   */
MPMK_API int mpmk_g2p(mpmk_ctx* ctx, float dt, void* stream);
/* all of the above in order: one reference p2g2p() call */
MPMK_API int mpmk_substep(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, void* stream);
/* n substeps with constant dt and ops, replayed from a CUDA graph */
MPMK_API int mpmk_substeps(mpmk_ctx* ctx, float dt, const mpmk_substep_ops* ops, int n, void* stream);

/* ---- export (solver.py:135-137 postprocess + main.py:311-312) ------------- */
/* utils.py:402-433 compute_cov_from_F and :377-398 compute_R_from_F fused with
 * the scatter back to original particle order.  Any output may be NULL.
 * init_cov_dev: float[n][6] in ORIGINAL order.                                */
MPMK_API int mpmk_export(mpmk_ctx* ctx, const void* init_cov_dev, void* xyz_out_dev /* float[n][3] */,
                void* cov_out_dev /* float[n][6] */, void* R_out_dev /* float[n][9] */, void* stream);

/* export straight into the rasteriser's frame: postprocess (utils.py:402-433) + grid2world
 * (utils/transform_utils.py:18-21, main.py:310-313) + render_frame's undo chain (main.py:139-146) in one pass:
 *   position = (x * alpha + beta) M,   covariance = M^T (gamma * F Sigma0 F^T) M   (upper triangle, 6 floats)
 * M (row-major) is the product of the pre-processing rotations in inverse order (transform_utils.py:106-123).    */
typedef struct {
    float alpha;
    float gamma;
    float beta[3];
    float M[9];
} mpmk_world_xform;
MPMK_API int mpmk_export_world(mpmk_ctx* ctx, const void* init_cov_dev, const mpmk_world_xform* xf,
                               void* xyz_out_dev /* float[n][3] */, void* cov_out_dev /* float[n][6] or NULL */, void* stream);

/* ---- multi-GPU slab decomposition (csrc/mg.cu) ------------------------------
 * The reference is single-GPU (one ti.init(arch=ti.cuda), dense global fields:
 * main.py:28, model.py:118-121); these entry points have no counterpart there.
 * One process per GPU.  Rank r owns the base-cell planes [x_lo, x_hi) of the
 * grid's slowest axis; between migrations its particles may sit up to `ghost`
 * cells outside.  After p2g the partial sums on the planes shared with each
 * neighbour are exchanged (ncclSend/ncclRecv on the caller's stream) and added
 * symmetrically, after which both neighbours run the grid update on them.
 * Particle migration is host-orchestrated (mpm_solver/slabs.py) between batches.  */
/* rank 0: create the 128-byte NCCL unique id that the host layer broadcasts     */
MPMK_API int mpmk_mg_unique_id(const char* nccl_lib_path, void* id_out, int id_bytes);
/* all ranks: join the communicator (nccl_lib_path: the libnccl.so.2 torch loaded) */
MPMK_API int mpmk_mg_init(mpmk_ctx* ctx, const char* nccl_lib_path, const void* id, int id_bytes, int rank, int world);
MPMK_API int mpmk_mg_finalize(mpmk_ctx* ctx);
/* [y_lo, y_hi): rows of each shared plane that can hold particles' stencils until the
 * next migration (the exchange, dense clear and dense update are restricted to them);
 * recv_*_dev: float4[(2*ghost+2) * (y_hi-y_lo) * n_grid] scratch each (NULL at the
 * outer ranks); from here on mpmk_substep/mpmk_substeps include the halo sum      */
MPMK_API int mpmk_mg_set_slab(mpmk_ctx* ctx, int x_lo, int x_hi, int ghost, int y_lo, int y_hi, void* recv_lo_dev,
                              void* recv_hi_dev);
/* [z_lo, z_hi): the nodes along the fastest grid axis any rank's stencils can reach before the next migration (same
 * on every rank).  Narrows the peer-memory halo and the dense update of the shared planes to rows x this window (the
 * NCCL halo path keeps whole rows).  mpmk_mg_set_slab resets it to [0, n_grid).                       */
MPMK_API int mpmk_mg_set_window(mpmk_ctx* ctx, int z_lo, int z_hi);
/* peer-memory halo: substep epoch at which this rank gave up waiting (4 s) for a neighbour's push; 0 = never.
 * Synchronises `stream`.  mpmk_mg_scan also reports it in stats[8], which is how slabs.py raises on every rank. */
MPMK_API int mpmk_mg_peer_error(mpmk_ctx* ctx, uint32_t* epoch_out, void* stream);
/* Small device -> host read-back that does not use a copy engine: a kernel stores `n_words` 32-bit words from
 * `src_dev` into `dst_pinned`, page-locked host memory the device can address (cudaHostAlloc / torch pin_memory()
 * under unified addressing), and `stream` is synchronised.  A cudaMemcpy of a few bytes queues on the device-to-host
 * copy engine behind whatever bulk download the caller has in flight on another stream (measured at 8 GPUs: the 72-byte
 * migration table waited ~12 ms per frame behind the previous frame's 147 MB download); a store from an SM does not.
 * No reference counterpart (slab plumbing, mpm_solver/slabs.py).                                              */
MPMK_API int mpmk_read_small(mpmk_ctx* ctx, void* dst_pinned, const void* src_dev, int n_words, void* stream);
/* new live particle count after a host-side migration (mpm_solver/slabs.py); unlike mpmk_set_count
 * it keeps the current set and the validity of the stored F_trial records                       */
MPMK_API int mpmk_mg_set_count(mpmk_ctx* ctx, int64_t n_particles);
/* Migration helpers; the protocol is mpm_solver/slabs.py.  A record = the 16-byte entries of the 12 set arrays
 * and ft0..ft2 (15 x 4 words) + the key = 61 x int32.
 * scan: one pass over the live keys; stats_dev int32[12] = {leavers to rank-1, leavers to rank+1, particles more
 * than `ghost` planes outside the slab, -(min base row), max base row, max |v| component (float bits), -(min base z),
 * max base z, halo timeout epoch (peer path, else 0), 0, 0, 0};
 * idx_lo_dev / idx_hi_dev int32[capacity] receive the slots of the leavers (unordered).                      */
MPMK_API int mpmk_mg_scan(mpmk_ctx* ctx, void* stats_dev, void* idx_lo_dev, void* idx_hi_dev, void* stream);
/* rows_dev int32[count][61] <- records at slots idx_dev[0..count)                                          */
MPMK_API int mpmk_mg_pack(mpmk_ctx* ctx, const void* idx_dev, int64_t count, void* rows_dev, void* stream);
/* records at slots idx_dev[r] (or first_slot + r when idx_dev is NULL) <- rows_dev[r]                        */
MPMK_API int mpmk_mg_unpack(mpmk_ctx* ctx, const void* idx_dev, int64_t first_slot, int64_t count, const void* rows_dev,
                            void* stream);
/* the slab shrinks to n_new < n: move the live records of [n_new, n) into the n_holes unfilled leaver slots
 * (holes_dev) that lie below n_new; scratch_dev: 8 + 4*n_holes + (n - n_new) bytes                           */
MPMK_API int mpmk_mg_close_holes(mpmk_ctx* ctx, const void* holes_dev, int64_t n_holes, int64_t n_new, void* scratch_dev,
                                 int64_t scratch_bytes, void* stream);
/* Peer-memory halo path (same node, NVLink): every rank allocates an exchange block (control words + 4 receive
 * buffers of buf_bytes: two neighbours x double buffering) and exports it as a 64-byte CUDA IPC handle; the host layer
 * passes the handles to the neighbours, which map them.  From mpmk_mg_peer_open on, the halo sum of mpmk_substep(s) is
 * one copy kernel per neighbour that stores this rank's partial sums into the neighbour's memory and raises a flag
 * there (st.release.sys), and one kernel that waits for the neighbours' flags and adds what they pushed -- no NCCL
 * call on the substep path.  buf_bytes >= (2*ghost+2) * n_grid^2 * 16 covers any row range.                     */
MPMK_API int mpmk_mg_peer_alloc(mpmk_ctx* ctx, int64_t buf_bytes, void* handle_out, int handle_bytes);
/* the same with 2 x 2 message slots of msg_bytes each behind the halo slots (mpmk_mg_ll_send / _recv)           */
MPMK_API int mpmk_mg_peer_alloc_ex(mpmk_ctx* ctx, int64_t buf_bytes, int64_t msg_bytes, void* handle_out, int handle_bytes);
/* Flagged message to / from a neighbour through the mapped exchange blocks (the migration payload of
 * mpm_solver/slabs.py; replaces ncclSend / ncclRecv there).  n_words 32-bit words travel as 16-byte stores
 * {w0, tag, w1, tag} over NVLink into slot [side][tag & 1]; the receiver spins until the tags match: no collective,
 * no fence, nothing on the host.  side 0 = rank-1, 1 = rank+1.  tag != 0, different from the previous message's tag
 * through the same slot, same on both sides.  A receive that times out (4 s) sets the error word mpmk_mg_peer_error
 * and mpmk_mg_scan report.  No reference counterpart.                                                          */
MPMK_API int mpmk_mg_ll_send(mpmk_ctx* ctx, int side, const void* src_dev, int64_t n_words, uint32_t tag, void* stream);
MPMK_API int mpmk_mg_ll_recv(mpmk_ctx* ctx, int side, void* dst_dev, int64_t n_words, uint32_t tag, void* stream);
MPMK_API int mpmk_mg_peer_open(mpmk_ctx* ctx, const void* handle_lo, const void* handle_hi);
MPMK_API int mpmk_mg_peer_enabled(mpmk_ctx* ctx);
/* unmap / free the exchange blocks and return to the NCCL halo path */
MPMK_API int mpmk_mg_peer_close(mpmk_ctx* ctx);
/* the halo sum alone (stagewise use: after mpmk_p2g, before mpmk_grid_update)     */
MPMK_API int mpmk_mg_halo_exchange(mpmk_ctx* ctx, void* stream);

/* The most recent sort, wherever it ran (mpmk_sort, or the side branch of a lagged substep inside mpmk_substeps):
 * sorted keys, the permutation (physical slot, AT THE TIME OF THAT SORT, of the r-th particle in sorted order) and --
 * lagged substeps only -- its inverse, the slot g2p moved each particle to.  Any pointer may be NULL.  For tests. */
MPMK_API int mpmk_get_last_sort(mpmk_ctx* ctx, void* sorted_keys_dev, void* perm_dev, void* rank_dev, void* stream);

/* ---- differentiable ("fitting") path: solver.py:54-108, extra.py:199-236 ---------------------------
 * Forward kernels with a substep history and hand-written adjoints (csrc/opt.cu) in place of Taichi's autodiff.
 * All arrays are plain row-major device arrays in ORIGINAL particle order, one substep slice per call:
 * x, v float[n][3]; F, C, stress float[n][9]; cov, init_cov float[n][6]; mu, lam, logE, y, mass, vol float[n];
 * grid_v_in, grid_v_out float[n_grid^3][3]; grid_mass float[n_grid^3] (dense, zeroed by the caller per substep as
 * MPM_state_opt.reset_grid_state does).  g_* are the adjoints (".grad" fields); the *_grad entry points ACCUMULATE
 * into their outputs like Taichi's reverse-mode kernels.  n_grid, dx, inv_dx come from the context.              */
/* utils.py:56-76 compute_stress_from_F_opt and its adjoint */
MPMK_API int mpmk_opt_stress(mpmk_ctx* ctx, int64_t n, const float* F, const float* mu, const float* lam, float* stress, void* stream);
MPMK_API int mpmk_opt_stress_grad(mpmk_ctx* ctx, int64_t n, const float* F, const float* mu, const float* lam, const float* g_stress,
                                  float* g_F, float* g_mu, float* g_lam, void* stream);
/* utils.py:137-174 p2g_opt and its adjoint (grid_mass, mass and vol carry no gradient) */
MPMK_API int mpmk_opt_p2g(mpmk_ctx* ctx, int64_t n, const float* x, const float* v, const float* C, const float* stress, const float* mass,
                          const float* vol, float* grid_v_in, float* grid_mass, float dt, void* stream);
MPMK_API int mpmk_opt_p2g_grad(mpmk_ctx* ctx, int64_t n, const float* x, const float* v, const float* C, const float* stress,
                               const float* mass, const float* vol, const float* g_grid_v_in, float* g_x, float* g_v, float* g_C,
                               float* g_stress, float dt, void* stream);
/* utils.py:178-183 + grid_postprocess[0].apply (boundary_conditions.py:23-27; box_* NULL = none) and their adjoint */
MPMK_API int mpmk_opt_grid(mpmk_ctx* ctx, const float* grid_v_in, const float* grid_mass, float* grid_v_out, float dt, const float* gravity3,
                           const float* box_center3, const float* box_size3, void* stream);
/* block_*: NULL = reference-literal (Taichi leaves the adjoint of a node the box overwrote in place, so its gradient
 * still flows back); the box = textbook adjoint (nothing flows through a zeroed node) */
MPMK_API int mpmk_opt_grid_grad(mpmk_ctx* ctx, const float* grid_mass, const float* g_grid_v_out, float* g_grid_v_in,
                                const float* block_center3, const float* block_size3, void* stream);
/* utils.py:285-347 g2p_opt (writes slice s+1 from slice s) and its adjoint */
MPMK_API int mpmk_opt_g2p(mpmk_ctx* ctx, int64_t n, const float* x, const float* F, const float* grid_v_out, float* x1, float* v1, float* C1,
                          float* F1, float dt, void* stream);
MPMK_API int mpmk_opt_g2p_grad(mpmk_ctx* ctx, int64_t n, const float* x, const float* F, const float* grid_v_out, const float* g_x1,
                               const float* g_v1, const float* g_C1, const float* g_F1, float* g_grid_v_out, float* g_x, float* g_F, float dt,
                               void* stream);
/* utils.py:436-467 compute_cov_from_F_opt and its adjoint */
MPMK_API int mpmk_opt_cov(mpmk_ctx* ctx, int64_t n, const float* F, const float* init_cov, float* cov, void* stream);
MPMK_API int mpmk_opt_cov_grad(mpmk_ctx* ctx, int64_t n, const float* F, const float* init_cov, const float* g_cov, float* g_F, void* stream);
/* adjoint of utils.py:350-362 compute_mu_lam_from_E_nu, and solver.py:92-108 learn (clipped gradient step) */
MPMK_API int mpmk_opt_mu_lam_grad(mpmk_ctx* ctx, int64_t n, const float* logE, const float* y, const float* g_mu, const float* g_lam,
                                  float* g_logE, float* g_y, void* stream);
MPMK_API int mpmk_opt_learn(mpmk_ctx* ctx, int64_t n, float* logE, float* y, const float* g_logE, const float* g_y, void* stream);

/* ---- introspection -------------------------------------------------------- */
/* With MPMK_TRACE=1 in the environment: %globaltimer (ns) at the stage boundaries of the last lagged substep of the last
 * batch -- [0] start, [1] p2g done, [2] grid done, [3] g2p done; side branch [5] clear done, [6] sort done, [7] work
 * lists done.  Synchronises the device.  Tuning aid. */
MPMK_API int mpmk_debug_trace(mpmk_ctx* ctx, unsigned long long* out16);
/* synchronises `stream` */
MPMK_API int mpmk_read_counters(mpmk_ctx* ctx, mpmk_counters* out, void* stream);
/* how many kernels the library has launched since create (for bench's gpu_launches) */
MPMK_API int64_t mpmk_launch_count(mpmk_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* MPMK_H */
