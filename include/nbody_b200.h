/* include/nbody_b200.h — C-ABI of the B200-native physics step.
 *
 * Drop-in boundary for the hot path of tr0yspradling/raylib-nbody: the three private
 * statics its flecs systems call (src/systems/Physics.hpp:32-64). The reference has no
 * FFI of its own; the seam is the gather/scatter code inside those statics
 * (Physics.hpp:197-205, :120-121, :260), and these entry points are what a maintainer
 * binds in their place (INTEGRATION.md shows the patch to register_systems).
 *
 * Conventions
 *   - plain C, POD only, no exceptions across the boundary; every call returns an
 *     nb_status (0 = ok) and nb_last_error() gives the message for the calling thread.
 *   - host arrays are borrowed for the duration of a call; device state is owned by the
 *     opaque nb_ctx. Body identity = array index in upload order.
 *   - 2-D entry points use the reference's layouts exactly: interleaved {x,y} doubles
 *     (DVec2, src/core/Math.hpp:6-9), f32 mass (Components.hpp:22-24), u8 pinned
 *     (Components.hpp:36-38). The *3 variants carry {x,y,z}; a 2-D upload is the 3-D
 *     state with z == 0 (every z term is then exactly zero).
 *   - there is NO CPU fallback: without a CUDA device nb_create fails with
 *     NB_ERR_NO_DEVICE and nothing else can be called.
 */
#ifndef NBODY_B200_H
#define NBODY_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NB_ABI_VERSION 1

typedef enum nb_status {
    NB_OK = 0,
    NB_ERR_INVALID = 1,    /* bad argument / call order */
    NB_ERR_NO_DEVICE = 2,  /* no usable CUDA device (never falls back to the CPU) */
    NB_ERR_CUDA = 3,       /* CUDA runtime failure, message in nb_last_error() */
    NB_ERR_NCCL = 4,       /* NCCL failure or NCCL not loadable */
    NB_ERR_ALLOC = 5
} nb_status;

typedef struct nb_ctx nb_ctx;

/* Physics fields of the reference's Config singleton (src/core/Config.hpp:13-29,
 * defaults src/core/Constants.hpp:86-100). bh_threshold/bh_theta are set with nb_set_solver:
 * by default this library always evaluates the direct sum (Physics.hpp:231-258). */
typedef struct nb_config {
    double g;                   /* Config::g                      6.6743e-11 */
    float softening;            /* Config::softening (epsilon, m) 4.0        */
    float max_speed;            /* Config::max_speed, 0 = off     0.0        */
    int32_t integrator;         /* 0 semi-implicit Euler, 1 velocity Verlet (default 1) */
    float max_substep;          /* Config::max_substep            1e9        */
    int32_t max_substeps_per_frame; /* Config::max_substeps_per_frame 200    */
} nb_config;

/* Arithmetic of the force evaluation (nb_set_precision).
 *   FP32       the hot path: FP32 pair terms, FP64 accumulation across 64-source chunks and
 *              FP64 integrator state; per-body acceleration within 1e-5 of the reference.
 *   FP64_EXACT the reference's own FP64 operation order (Physics.hpp:237-254), bit-identical
 *              accelerations and trajectories; about 1/15 of the FP32 throughput.
 *   AUTO       FP64_EXACT for worlds of at most NB_EXACT_AUTO_MAX bodies (where the force
 *              pass is launch-bound anyway and configurations such as the three-body seed are
 *              dynamically unstable), FP32 above. Default. */
#define NB_PRECISION_AUTO 0
#define NB_PRECISION_FP32 1
#define NB_PRECISION_FP64_EXACT 2
#define NB_EXACT_AUTO_MAX 512

/* Which branch of Physics::compute_gravity runs (nb_set_solver).
 *   DIRECT     always the all-pairs sum (Physics.hpp:231-258) — the reference's behaviour when
 *              n <= bh_threshold, and this library's hot path at every size. Default.
 *   REFERENCE  the reference's own switch (Physics.hpp:212): a world with more than bh_threshold
 *              bodies passing the gather filter is evaluated with the reference's float
 *              Barnes-Hut quadtree (src/physics/SpatialPartition.hpp:36-220, opening angle
 *              bh_theta), reproduced bit for bit on the device together with the FP64 integrator
 *              arithmetic, so whole trajectories equal the reference's at its DEFAULT settings
 *              (bh_threshold = 100, bh_theta = 0.5: Constants.hpp:90-91). Smaller worlds take the
 *              direct sum as under DIRECT. 2-D uploads, single-GPU contexts. Two bodies whose
 *              float coordinates coincide make the reference subdivide without end; here the
 *              force evaluation returns NB_ERR_INVALID instead. */
#define NB_SOLVER_DIRECT 0
#define NB_SOLVER_REFERENCE 1

#define NB_INTEGRATOR_EULER 0
#define NB_INTEGRATOR_VERLET 1

/* Physics::Diagnostics (Physics.hpp:22-30) widened to three components; for 2-D state
 * the [2] entries are 0. */
typedef struct nb_diag {
    double kinetic;
    double potential;
    double energy;
    double momentum[3];
    double com[3];
    double total_mass;
    int32_t ok; /* Diagnostics::ok: 0 when any sum is non-finite (the shim then sets cfg->paused) */
    int32_t _pad;
} nb_diag;

/* ---- lifetime -------------------------------------------------------------------- */

/* One context on CUDA device `device`. */
int nb_create(int device, nb_ctx** out);

/* A single-process context over n_dev GPUs of one box (1 <= n_dev <= 8; every pair must have peer access:
 * NVLink/NVSwitch). The reference is one process with one thread (src/main.cpp:100-143), so this — not
 * nb_create_sharded — is what its flecs systems can hold: the targets are partitioned over the devices like
 * nb_partition(n, i, n_dev), every call on the returned handle drives all devices, uploads and downloads use
 * the caller's full arrays, and the per-step exchange of packed positions is done by the packing kernels'
 * own stores into the peers' memory (no NCCL). n_dev == 1 is nb_create. Same restrictions as a sharded
 * context: FP32 path only (no FP64_EXACT, no NB_SOLVER_REFERENCE, no collision pass). */
int nb_create_multi(const int* devices, int n_dev, nb_ctx** out);
/* Devices behind a handle (1 for nb_create / nb_create_sharded contexts). */
int nb_num_devices(nb_ctx* ctx);

/* Target-partitioned context: rank `rank` of `world` uploads and downloads rows [lo,hi) of the caller's
 * arrays (nb_local_range), computes one P-th of the targets (nb_partition_info below says which) and
 * exchanges packed positions with its peers through NCCL once per force evaluation. `nccl_unique_id` is the 128-byte id from nb_comm_unique_id() on rank 0,
 * distributed by the caller (torch.distributed / MPI / a file). Collective over all ranks. */
int nb_comm_unique_id(void* out_id_128);
int nb_create_sharded(int device, int rank, int world, const void* nccl_unique_id, nb_ctx** out);

void nb_destroy(nb_ctx* ctx);
const char* nb_last_error(void);
int nb_abi_version(void);

/* ---- configuration ---------------------------------------------------------------- */

int nb_default_config(nb_config* out);            /* Constants.hpp:86-100 + Config.hpp:25 */
int nb_set_config(nb_ctx* ctx, const nb_config* cfg);
int nb_get_config(nb_ctx* ctx, nb_config* out);
int nb_set_precision(nb_ctx* ctx, int precision); /* NB_PRECISION_*; single-GPU contexts only for FP64_EXACT */
/* Config::bh_threshold / Config::bh_theta (Config.hpp:17-18) and whether to honour them. */
int nb_set_solver(nb_ctx* ctx, int solver, int32_t bh_threshold, float bh_theta);
int nb_get_solver(nb_ctx* ctx, int* solver, int32_t* bh_threshold, float* bh_theta);
/* Tree of the last Barnes-Hut pass: values[8] = passes so far, cells, deepest level, bodies in the
 * tree, and — counted only after nb_solver_profile(ctx, 1) — cells visited, terms added and accept
 * tests re-evaluated in FP64 by the walks of the last pass; values[7] = 0. Synchronises. */
int nb_solver_info(nb_ctx* ctx, int64_t* values);
int nb_solver_profile(nb_ctx* ctx, int enable);

/* ---- state ------------------------------------------------------------------------ */

/* Replaces all bodies (the gather of Physics.hpp:197-205). `pinned` may be NULL.
 * Accelerations start at zero like freshly created entities (Physics.hpp:97-98); the next
 * step evaluates a(x) first, as the gravity system does. Sharded: every rank passes the
 * same full arrays. */
int nb_upload(nb_ctx* ctx, int64_t n, const double* pos_xy, const double* vel_xy, const float* mass,
              const uint8_t* pinned);
int nb_upload3(nb_ctx* ctx, int64_t n, const double* pos_xyz, const double* vel_xyz, const float* mass,
               const uint8_t* pinned);

/* Same, additionally taking the Acceleration components the caller holds (may be NULL). When
 * given, the caller asserts they equal a(x) for these positions — true for components this
 * library scattered back after the previous frame, the situation the reference's comment at
 * Physics.hpp:324 describes — so a Verlet frame needs one force evaluation instead of two.
 * dim is 2 (reference layout) or 3. */
int nb_upload_state(nb_ctx* ctx, int64_t n, int dim, const double* pos, const double* vel, const double* acc,
                    const float* mass, const uint8_t* pinned);

/* Scatter back (Physics.hpp:260 and the in-place writes of :276-323). Any pointer may be
 * NULL. Arrays are sized for all n bodies; a sharded context fills rows [lo,hi) only. */
int nb_download(nb_ctx* ctx, double* pos_xy, double* vel_xy, double* acc_xy);
int nb_download3(nb_ctx* ctx, double* pos_xyz, double* vel_xyz, double* acc_xyz);
/* Same with the PrevAcceleration component (Components.hpp:19-21; written by the Verlet drift,
 * Physics.hpp:303) as a fourth array; dim is 2 or 3. */
int nb_download_state(nb_ctx* ctx, int dim, double* pos, double* vel, double* acc, double* acc_prev);

/* The order in which the FP32 path streams the bodies (a Hilbert-curve order of the positions at the last
 * sort; order[k] = the caller's row of the body at stream position k; a sharded context fills and reads entries
 * [lo,hi) only — stream positions, which in a multi-GPU context also say which GPU computes the body) and
 * the number of generations packed since that sort. The order fixes the summation order of every force sum,
 * so a checkpoint that must resume bit-identically carries both: nb_upload_state, then nb_set_stream_order.
 * The library re-sorts on its own every 64 generations (NB_RESORT_EVERY). */
int nb_stream_order(nb_ctx* ctx, int32_t* order, int64_t* generations_since_sort);
int nb_set_stream_order(nb_ctx* ctx, const int32_t* order, int64_t generations_since_sort);

/* Who owns which body in a multi-GPU context (nb_create_multi, nb_create_sharded) is the library's business: the
 * interface keeps speaking the caller's rows (uploads and downloads of rows [lo,hi), nb_local_range), while
 * every GPU computes the targets of one stretch of a Hilbert curve through the whole world (compact groups, few
 * two-float visits) and the rows travel between GPUs over NVLink at upload, download, and whenever the stream is
 * re-sorted. values[4] = {ownership along the curve in use (0 for one GPU or NB_CURVE_PARTITION=0), times the
 * ownership was computed, same-size uploads that re-used it since, Hilbert sorts of the stream}. */
int nb_partition_info(nb_ctx* ctx, int64_t* values);

int64_t nb_num_bodies(nb_ctx* ctx);
int nb_local_range(nb_ctx* ctx, int64_t* lo, int64_t* hi);
/* The row ranges of a multi-GPU context (the caller's rows a rank uploads / downloads, and the stretch of the
 * stream it computes): contiguous, aligned to the 512-body source tile, equal tile counts per rank (so the
 * all-gathers are uniform). Host-only, no context. */
int nb_partition(int64_t n, int rank, int world, int64_t* lo, int64_t* hi);

/* ---- the hot path ----------------------------------------------------------------- */

/* Physics::compute_gravity (Physics.hpp:183-261): Acceleration <- a(x) by direct sum.
 * Bodies with non-finite x/v or m <= 0 neither attract nor receive a new acceleration;
 * pinned bodies attract and get a = 0. */
int nb_compute_gravity(nb_ctx* ctx);

/* Physics::integrate (Physics.hpp:263-327) with the reference's substep plan. Requires
 * Acceleration == a(x) on entry, i.e. nb_compute_gravity ran since the last upload, exactly
 * as sys2 runs before sys3. */
int nb_integrate(nb_ctx* ctx, float dt);

/* n_frames x (gravity system + integrate system), Physics.hpp:45-57, with dt_eff =
 * baseDt * max(0, time_scale) computed by the caller (Physics.hpp:54-55). This is the fused
 * path: one force kernel per substep with the kick/drift in its epilogue; the gravity
 * system's evaluation is skipped when Acceleration already equals a(x). Asynchronous. */
int nb_step(nb_ctx* ctx, float dt_eff, int64_t n_frames);

/* Same, additionally accumulating the potential in the force pass so nb_step_diagnostics
 * returns the energy/momentum of the state at the end of the call without another O(N^2)
 * pass. For Euler it describes the state the last force evaluation saw (frame start), as
 * the reference's diagnostics do (Physics.hpp:34-42). */
int nb_step_with_diagnostics(nb_ctx* ctx, float dt_eff, int64_t n_frames);
int nb_step_diagnostics(nb_ctx* ctx, nb_diag* out);

/* Physics::compute_diagnostics (Physics.hpp:117-177) over ALL bodies, unfiltered.
 * Sharded: collective, every rank receives the global values. Synchronises. */
int nb_compute_diagnostics(nb_ctx* ctx, nb_diag* out);

/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). The same sums as
 * nb_compute_diagnostics on a context in NB_PRECISION_FP64_EXACT (the FP64 kernel of that path, same operation
 * order), without a context and without any allocation: inputs and output are DEVICE pointers the caller owns
 * (d_pos, d_vel: n x dim doubles in the reference's component layout, dim 2 or 3; d_mass: n floats; d_out: one
 * nb_diag), the two kernels are enqueued on `cuda_stream` (a cudaStream_t, NULL = the default stream) of device
 * `device`, and the call returns without synchronising: read *d_out after the stream has reached that point.
 * Scratch (per-CTA partial sums) lives in a static device array of 96 KiB that every GPU loading the library
 * holds until exit, sized for n <= NB_NOALLOC_MAX_BODIES; a larger n returns NB_ERR_INVALID (nothing enqueued).
 * Calls no cudaMalloc / cudaFree / allocating routine; the first call on a device looks the array's device
 * address up (cudaGetSymbolAddress), so make one call outside a capture first. All calls on a device share the
 * array: calls that overlap on the GPU (different streams) race on it — keep them on one stream or order them. */
#define NB_NOALLOC_MAX_BODIES 262144
int nb_diagnostics_noalloc(int device, int64_t n, int dim, const double* d_pos, const double* d_vel,
                           const float* d_mass, double g, float softening, nb_diag* d_out, void* cuda_stream);

/* Physics::zero_net_momentum (Physics.hpp:66-85). */
int nb_zero_net_momentum(nb_ctx* ctx);

/* Detection phase of Collision::resolve (src/systems/Collision.hpp:63-84) on the bodies the
 * context holds: the number of pairs i < j, both passing its input filter (finite position, finite
 * mass > 0), with dist2 <= (r_i + r_j)^2 evaluated in the reference's FP64 operation order, and
 * (first_pair, may be NULL) the first such pair in the reference's visiting order, (-1,-1) if none.
 * `radius` is the caller's Collision::radius_of per body (Radius component, or mass + density,
 * Collision.hpp:28-35). Zero overlaps means resolve() would change nothing this frame, so the
 * binding only runs the reference's merge code on frames where the count is non-zero.
 * Single-GPU contexts. Synchronises. */
int nb_count_overlaps(nb_ctx* ctx, const double* radius, int64_t* n_overlaps, int64_t* first_pair);

/* Collision::resolve in its default merge mode (src/systems/Collision.hpp:49-132) on the bodies the
 * context holds (2-D uploads), with the reference's outcome bit for bit: which bodies are destructed,
 * and the survivors' new Mass (float), Velocity, Position (momentum- and mass-weighted, :107-112),
 * Pinned (:120) and radius (:37-46). Every pair test runs on the device; the merges, which the
 * reference applies one after another with the survivor's CHANGED position, mass and radius feeding
 * the following tests (and with a row that keeps running after its own body died, :76), are replayed
 * on the host in the reference's order from the device's detections.
 *   radius[i]   in: Collision::radius_of per body on entry (:28-35)
 *   density[i]  in: the Density component, or 5510 (Constants.hpp:34)
 *   alive[i]    out: 0 where the reference would destruct the body (:124)
 *   changed[i]  out: 1 where a body survived a merge; its pos_xy/vel_xy/mass/pinned/radius_out entries
 *               are then written (others are left untouched) and the caller zeroes its Acceleration
 *               and PrevAcceleration (:118-119)
 *   merge_log   out (may be NULL): the merges in the order the reference performs them, as pairs
 *               {survivor, absorbed}; at most merge_log_cap pairs are written, n_merges is the full
 *               count. An ECS caller needs the order: flecs applies the deferred destruct() calls in
 *               the order they were issued, and the order decides the row order of the bodies left
 * The device state is not modified: the caller applies the result to its bodies and uploads the new
 * world. When no pair overlaps (the common frame) the call costs one detection pass and returns
 * n_merges = 0. Single-GPU contexts. Synchronises. */
int nb_resolve_collisions(nb_ctx* ctx, const double* radius, const double* density, int64_t* n_merges,
                          uint8_t* alive, uint8_t* changed, double* pos_xy, double* vel_xy, float* mass,
                          uint8_t* pinned, double* radius_out, int64_t* merge_log, int64_t merge_log_cap);

/* ---- scenario / checkpoint files (host-only; the reference keeps scenarios in memory, Scenario.hpp:14-128) ---- */

/* The Config subset a Scenario carries to replay itself (Scenario.hpp:26-38), physics fields only. */
typedef struct nb_scenario_config {
    nb_config physics;      /* g, softening, max_speed, integrator, max_substep, max_substeps_per_frame */
    int32_t bh_threshold;   /* Config.hpp:17 */
    float bh_theta;         /* :18 */
    int32_t use_fixed_dt;   /* :20 */
    float fixed_dt;         /* :21 */
    float time_scale;       /* :22 */
} nb_scenario_config;
#define NB_SCENARIO_HAS_ACC 1
#define NB_SCENARIO_HAS_ACC_PREV 2
#define NB_SCENARIO_HAS_ORDER 4
int nb_scenario_default_config(nb_scenario_config* out);
/* Bodies (BodySnapshot: pos, vel, mass, pinned; dim 2 = the reference's DVec2 layout) + the config subset, and
 * optionally what a bit-exact resume needs: Acceleration, PrevAcceleration, the stream order and its age
 * (nb_stream_order). One little-endian binary file, shared by scenarios.py and PhysicsB200.hpp. */
int nb_scenario_write(const char* path, int64_t n, int dim, const double* pos, const double* vel, const double* acc,
                      const double* acc_prev, const float* mass, const uint8_t* pinned, const nb_scenario_config* cfg,
                      const int32_t* order, int64_t order_age, int64_t frame);
int nb_scenario_read_header(const char* path, int64_t* n, int* dim, int* flags, nb_scenario_config* cfg,
                            int64_t* order_age, int64_t* frame);
int nb_scenario_read(const char* path, double* pos, double* vel, double* acc, double* acc_prev, float* mass,
                     uint8_t* pinned, int32_t* order);

/* Test hook: the device restatement of glibc's cbrt (update_radius_from_mass, Collision.hpp:37-46) on n values. */
int nb_debug_cbrt(nb_ctx* ctx, const double* in, double* out, int64_t n);

/* Developer counters of the last nb_resolve_collisions that merged: values[4] = {merges, events popped and
 * re-tested, bodies visited by the scans after the merges, initially overlapping pairs}. */
int nb_merge_info(nb_ctx* ctx, int64_t* values);

/* Substep plan of Physics.hpp:268-271 (float arithmetic). Host-only helper. */
int nb_substep_plan(float dt, float max_substep, int32_t max_substeps_per_frame, float* dt_sub);

/* Physics::reset_scenario (Physics.hpp:87-115): uploads the three-body seed. */
int nb_reset_scenario(nb_ctx* ctx);

/* ---- synchronisation and measurement ---------------------------------------------- */

int nb_synchronize(nb_ctx* ctx);
/* CUDA events on the library's own stream around whatever is enqueued between them. */
int nb_timer_begin(nb_ctx* ctx);
int nb_timer_end(nb_ctx* ctx, float* elapsed_ms); /* synchronises */
/* SM count, compute capability and max SM clock (kHz) of the context's device. */
int nb_device_info(nb_ctx* ctx, int* sm_count, int* cc_major, int* cc_minor, int* sm_clock_khz);
/* Evicts L2 by overwriting a 256 MiB scratch buffer on the library's stream (benchmark hygiene). */
int nb_flush_l2(nb_ctx* ctx);
/* Page-locked host memory for callers that want full-speed uploads/downloads. */
int nb_alloc_pinned(size_t bytes, void** out);
int nb_free_pinned(void* p);
/* Kernels launched / force evaluations / body-body interactions since creation. */
int nb_counters(nb_ctx* ctx, int64_t* kernel_launches, int64_t* force_evals, double* interactions);
/* Device-time of the force kernels alone since the last call (events around every force
 * launch when profiling is on). */
int nb_profile_forces(nb_ctx* ctx, int enable);
int nb_profile_read(nb_ctx* ctx, int64_t* n_launches, double* total_ms);
/* Sharded contexts, Velocity Verlet: CUDA events around the last own-sources launch and around the exchange
 * (all-gather of the packed stream) it overlapped. ms4 = {own-sources launch duration, exchange start and
 * exchange end relative to the launch's start (negative = before it), exchange duration}. Call after a step;
 * enable = 1 arms the events for the following steps. Synchronises. */
int nb_overlap_profile(nb_ctx* ctx, int enable, float* ms4);
/* Developer counters of the force kernel's two-float ("near") path: how many (warp, source group) visits
 * the launches since the last call made and how many of them took the near path. Reads and clears the
 * counters; enable = 0 also releases them (the kernel then skips the counting). Synchronises. */
int nb_near_stats(nb_ctx* ctx, int enable, int64_t* visits, int64_t* near_visits);

#ifdef __cplusplus
}
#endif
#endif /* NBODY_B200_H */
