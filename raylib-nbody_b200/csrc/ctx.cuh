// ctx.cuh — the context behind the opaque nb_ctx handle: device buffers, streams, the
// force-launch plan, and (for target-partitioned contexts) the NCCL communicator.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/nbody_b200.h"
#include "common.cuh"
#include "state.cuh"

namespace nb {

// NCCL is resolved at run time (dlopen) so single-GPU users never need it on the loader path.
struct NcclApi;

// How one force launch is cut into CTAs.
struct LaunchPlan {
    int variant = 0;   // index into the kernel table (R, packed)
    int r = 4;         // targets per thread
    int nseg = 1;      // source segments
    long long n_itiles = 0;
    int ctas_per_sm = 1;
};

struct Variant {
    int nt, r, packed;
};

}  // namespace nb

struct nb_ctx;
// A single-process multi-GPU context (nb_create_multi): one member context per device, member 0 is the
// handle the caller holds. Members exchange the packed stream through peer memory (the packing kernels
// store to every member's buffers) and order themselves with CUDA events; there is no NCCL on this path.
struct nb_group {
    std::vector<nb_ctx*> m;
};

struct nb_ctx {
    nb_group* group = nullptr;           // non-null on every member of a multi-GPU context
    cudaEvent_t ev_pub = nullptr;        // "my rows of the newest generation are in everyone's buffers, and I am
                                         //  done reading the previous one" (recorded after each packing launch)
    cudaEvent_t ev_bar = nullptr;        // team barriers around the rare multi-phase operations
    int device = 0;
    int rank = 0, world = 1;
    int sm_count = 148;
    cudaStream_t stream = nullptr;       // all kernels
    cudaStream_t comm_stream = nullptr;  // NCCL exchange, high priority
    cudaEvent_t ev_timer0 = nullptr, ev_timer1 = nullptr;
    cudaEvent_t ev_packed = nullptr, ev_gathered = nullptr;

    nb_config cfg;
    int dim = 2;  // 2: reference layout uploads, 3: xyz uploads
    int precision = NB_PRECISION_AUTO;
    bool step_diag_exact = false;  // nb_step_diagnostics recomputes (exact path has no fused potential)

    // ---- bodies ----
    long long n = 0;          // all bodies
    long long cap = 0;        // allocated rows
    long long n_sized = -1;   // body count the buffers were last initialised for
    long long lo = 0, hi = 0; // rows owned by this rank
    long long n_tiles = 0;    // j-stream tiles (padded to world * tiles_per_rank)
    long long tiles_per_rank = 0;
    double* state = nullptr;  // 15 * cap doubles: x[3], v[3], a[3], a0[3], and the shadow positions xs[3]
    int x_sel = 0;            // which of the two position slots (0: rows 0-2, 1: rows 12-14) is Position
    float* mass = nullptr;
    uint8_t* pinned = nullptr;
    uint8_t* live2[2] = {nullptr, nullptr};  // gather-filter flags per generation parity (body rows)
    double* tx[2] = {nullptr, nullptr}; // FP64 positions in stream order per generation, 3 * cap each (targets' coordinates)
    double* dtx = nullptr;              // ... of the stand-alone diagnostics stream
    int* order = nullptr;               // stream row -> body row (Hilbert-curve order of this rank's rows)
    float* pk[2] = {nullptr, nullptr};  // pair-SoA j-stream generations: hi parts
    float* pk_lo[2] = {nullptr, nullptr};   // ... residuals (near path)
    nb::GroupHdr* hdr[2] = {nullptr, nullptr};  // ... group origins / half-extents
    float* dpk = nullptr;               // stand-alone diagnostics stream (hi, lo, headers)
    float* dpk_lo = nullptr;
    nb::GroupHdr* dhdr = nullptr;
    unsigned long long* sort_keys[2] = {nullptr, nullptr};
    int* sort_vals[2] = {nullptr, nullptr};
    void* sort_tmp = nullptr;
    size_t sort_tmp_bytes = 0;
    bool order_valid = false;           // `order` was sorted for this upload
    long long packs_since_sort = 0;     // generations packed since the last sort (re-sorted every kResortEvery)
    unsigned int* extent = nullptr;     // kExtentSlots + 4 slots
    double* part = nullptr;             // FP64 partial sums, (1 + part_slots) * 4 * i_cap
    long long i_cap = 0;
    int part_slots = 0;
    unsigned int* tile_counter = nullptr;
    long long counter_cap = 0;
    double* diag_part = nullptr;        // per-CTA diagnostics partials
    long long diag_part_rows = 0;
    double* totals = nullptr;           // device scratch for fixed-order final sums (64 doubles)
    double* h_totals = nullptr;         // pinned host mirror
    int* h_flip = nullptr;              // pinned, written by the kick epilogue: a body's gather-filter outcome changed
    void* l2_scratch = nullptr;         // 256 MiB, nb_flush_l2
    void* staging = nullptr;            // device staging for AoS uploads/downloads
    void* collide_scratch = nullptr;    // nb_count_overlaps: radii, grid arrays, scan temp
    size_t collide_bytes = 0;
    long long overlap_grid_calls = 0;   // calls that took the O(N) grid broad phase
    void* merge_scratch = nullptr;      // nb_resolve_collisions: pair sink, current positions, hit list
    size_t merge_bytes = 0;
    void* replay_scratch = nullptr;     // merge_replay_kernel: sorted keys, heap, snapshots, log
    size_t replay_bytes = 0;
    long long merge_last = 0, merge_initial = 0, merge_events = 0, merge_candidates = 0;  // of the last replay: events popped, bodies scanned
    long long merge_passes = 0;         // calls that found an overlap and replayed merges
    size_t staging_bytes = 0;

    // ---- the reference's solver switch (Physics.hpp:212): Barnes-Hut above bh_threshold bodies ----
    int solver = NB_SOLVER_DIRECT;
    int bh_threshold = 100;             // Constants.hpp:90
    float bh_theta = 0.5f;              // Constants.hpp:91
    void* bh_scratch = nullptr;         // tree arrays, sort buffers
    size_t bh_bytes = 0;
    long long bh_capacity = 0;          // cells allocated
    long long bh_rows = 0;              // body rows the scratch was laid out for
    unsigned int* h_bh_ctl = nullptr;   // pinned mirror of the build's control words + walk statistics
    bool bh_stats = false;
    int bh_coop_ctas_per_sm = 0;
    long long bh_passes = 0, bh_nodes = 0, bh_depth = 0, bh_live = 0;
    unsigned long long bh_walk[3] = {0, 0, 0};  // cells popped, terms, exact accept tests (when bh_stats)

    // ---- ownership along one Hilbert curve of the whole world (multi-GPU contexts, partition.cuh) ----
    bool curve_partition = true;        // NB_CURVE_PARTITION=0: ranks own their upload-order rows (round 1 behaviour)
    bool perm_valid = false;            // perm / inv_perm describe the bodies held now
    long long uploads_since_partition = 0;
    int* perm = nullptr;                // internal row -> upload row, all n rows, identical on every rank
    int* inv_perm = nullptr;            // upload row -> internal row
    double* xb = nullptr;               // exchange buffer: x, v, a, a0 of EVERY row (12 arrays of cap rows)
    float* xb_mass = nullptr;
    uint8_t* xb_pinned = nullptr;
    unsigned long long* part_keys[2] = {nullptr, nullptr};
    int* part_vals[2] = {nullptr, nullptr};
    void* part_tmp = nullptr;
    size_t part_tmp_bytes = 0;
    long long part_cap = 0;             // rows the buffers above were allocated for
    long long partitions = 0;           // times the ownership was recomputed

    int gen = 0;               // generation of pk[gen & 1]
    int mass_exp = 0;          // masses are packed as m * 2^-mass_exp
    bool uniform_mass = false; // every body that can attract has the same mass -> no m*w multiply in the kernel
    bool diag_uniform_mass = false;  // EVERY body has that mass (compute_diagnostics counts all of them)
    double uniform_packed_mass = 1.0;
    bool planar = false;       // 2-D upload: z == 0 forever -> kernels skip the z terms
    bool force_generic = false;
    bool acc_valid = false;    // Acceleration == a(x) for the current positions
    bool packed_valid = false; // pk[gen & 1] matches the current positions
    // Velocity Verlet: every kick epilogue also drifts the NEXT substep into the shadow positions and packs
    // generation gen+1 from them. The next substep commits that (swap x/xs, gen += 1) when it uses the same
    // dt and nothing touched the state in between; otherwise it is discarded.
    bool spec_valid = false;
    double spec_h = 0.0;
    int spec_kind = 0;         // kernel kind (0 force, 1 force + potential) of the pass that prepared it
    bool step_diag_valid = false;
    long long step_diag_rows = 0;

    // ---- launch plans ----
    nb::LaunchPlan plans[3][3];  // [all | own sources | remote sources][force | force + potential | potential only]
    int forced_r = 0, forced_nseg = 0;
    int resort_every = 64;     // generations between two Hilbert sorts of the stream (NB_RESORT_EVERY)

    // ---- measurement ----
    long long kernel_launches = 0;
    long long force_evals = 0;
    long long sorts = 0;
    unsigned long long* near_stats = nullptr;  // device counter of nb_near_stats ([1]: near-path visits)
    long long near_visits = 0;                 // (warp, group) visits launched while the counter was armed
    bool profile_comm = false;           // record ev_comm0/1 around every exchange (overlap evidence)
    cudaEvent_t ev_comm0[2] = {nullptr, nullptr}, ev_comm1[2] = {nullptr, nullptr};  // per generation parity
    cudaEvent_t ev_own0 = nullptr, ev_own1 = nullptr;
    double interactions = 0.0;
    bool profile = false;
    std::vector<cudaEvent_t> prof_events;  // pairs
    size_t prof_used = 0;

    // ---- NCCL ----
    void* nccl_comm = nullptr;

    double* sx(int c) const { return state + (long long)c * cap; }
    double* pos_slot(int sel, int k) const { return sx((sel ? 12 : 0) + k); }
    uint8_t* live() const { return live2[gen & 1]; }
    void drop_speculation() { spec_valid = false; }
    // after any synchronisation: a kick changed the gather filter -> the next frame re-packs and re-evaluates gravity
    void absorb_filter_flip() {
        if (h_flip && *h_flip) {
            *h_flip = 0;
            acc_valid = false;
            packed_valid = false;
            spec_valid = false;
        }
    }
    nb::StateView view() const {
        nb::StateView s;
        for (int k = 0; k < 3; ++k) {
            s.x[k] = pos_slot(x_sel, k);
            s.v[k] = sx(3 + k);
            s.a[k] = sx(6 + k);
            s.a0[k] = sx(9 + k);
        }
        s.mass = mass;
        s.pinned = pinned;
        s.live = live();
        s.n = n;
        s.lo = lo;
        s.hi = hi;
        return s;
    }
};

namespace nb {

// ---- the team behind a handle ---------------------------------------------------------------------
// A plain or NCCL-sharded context is a team of one. The handle of a single-process multi-GPU context
// (nb_create_multi) is member 0 of its group; every operation below enqueues its phases member by member.
struct Team {
    nb_ctx* solo[1];
    nb_ctx** first;
    size_t count;
    explicit Team(nb_ctx* c) {
        solo[0] = c;
        if (c && c->group) { first = c->group->m.data(); count = c->group->m.size(); }
        else { first = solo; count = 1; }
    }
    nb_ctx** begin() const { return first; }
    nb_ctx** end() const { return first + count; }
};

}  // namespace nb
