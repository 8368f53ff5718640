// nbody_b200.cu — C-ABI implementation (include/nbody_b200.h) over the sm_100a kernels.
// Host-side mirror of the reference's frame schedule, src/systems/Physics.hpp:32-64:
// which force evaluations a frame needs, the substep plan, and when cached device state
// (accelerations, the FP32 j-stream) is still valid. No CPU fallback exists in this file.
#include <cuda_runtime.h>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <dlfcn.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <set>
#include <string>
#include <vector>

#include "bh.cuh"
#include "collide.cuh"
#include "merge.cuh"
#include "ctx.cuh"
#include "diag.cuh"
#include "exact.cuh"
#include "force_kernel.cuh"
#include "partition.cuh"
#include "state.cuh"

using namespace nb;

// ---------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------
static thread_local std::string g_last_error;

static int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

#define NB_CUDA(expr)                                                                              \
    do {                                                                                           \
        cudaError_t e_ = (expr);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(NB_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_) + " (" +  \
                                         __FILE__ + ":" + std::to_string(__LINE__) + ")");        \
    } while (0)

#define NB_TRY(expr)             \
    do {                         \
        int r_ = (expr);         \
        if (r_ != NB_OK) return r_; \
    } while (0)

extern "C" const char* nb_last_error(void) { return g_last_error.c_str(); }
extern "C" int nb_partition(int64_t n, int rank, int world, int64_t* lo, int64_t* hi);
static bool use_exact(const nb_ctx* c);
extern "C" int nb_abi_version(void) { return NB_ABI_VERSION; }

// ---------------------------------------------------------------------------------------
// NCCL, resolved at run time
// ---------------------------------------------------------------------------------------
namespace {

struct NcclId {
    char bytes[128];
};
typedef int (*fn_GetUniqueId)(NcclId*);
typedef int (*fn_CommInitRank)(void**, int, NcclId, int);
typedef int (*fn_CommDestroy)(void*);
typedef int (*fn_AllGather)(const void*, void*, size_t, int, void*, cudaStream_t);
typedef int (*fn_AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_Group)(void);
typedef const char* (*fn_ErrStr)(int);

struct Nccl {
    void* handle = nullptr;
    fn_GetUniqueId GetUniqueId = nullptr;
    fn_CommInitRank CommInitRank = nullptr;
    fn_CommDestroy CommDestroy = nullptr;
    fn_AllGather AllGather = nullptr;
    fn_AllReduce AllReduce = nullptr;
    fn_Group GroupStart = nullptr, GroupEnd = nullptr;
    fn_ErrStr GetErrorString = nullptr;
};
Nccl g_nccl;

// ncclDataType_t / ncclRedOp_t values (stable across NCCL 2.x)
constexpr int kNcclFloat32 = 7, kNcclFloat64 = 8, kNcclUint32 = 3, kNcclUint8 = 1, kNcclInt32 = 2, kNcclSum = 0, kNcclMax = 2;

int load_nccl() {
    if (g_nccl.handle) return NB_OK;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);  // torch's copy if already mapped
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return fail(NB_ERR_NCCL, std::string("cannot load libnccl.so.2: ") + dlerror());
    g_nccl.GetUniqueId = (fn_GetUniqueId)dlsym(h, "ncclGetUniqueId");
    g_nccl.CommInitRank = (fn_CommInitRank)dlsym(h, "ncclCommInitRank");
    g_nccl.CommDestroy = (fn_CommDestroy)dlsym(h, "ncclCommDestroy");
    g_nccl.AllGather = (fn_AllGather)dlsym(h, "ncclAllGather");
    g_nccl.AllReduce = (fn_AllReduce)dlsym(h, "ncclAllReduce");
    g_nccl.GroupStart = (fn_Group)dlsym(h, "ncclGroupStart");
    g_nccl.GroupEnd = (fn_Group)dlsym(h, "ncclGroupEnd");
    g_nccl.GetErrorString = (fn_ErrStr)dlsym(h, "ncclGetErrorString");
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllGather || !g_nccl.AllReduce || !g_nccl.GroupStart ||
        !g_nccl.GroupEnd)
        return fail(NB_ERR_NCCL, "libnccl.so.2 lacks a required symbol");
    g_nccl.handle = h;
    return NB_OK;
}

int nccl_fail(int rc, const char* what) {
    return fail(NB_ERR_NCCL,
                std::string(what) + ": " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "nccl error"));
}
#define NB_NCCL(expr)                            \
    do {                                         \
        int rc_ = (expr);                        \
        if (rc_ != 0) return nccl_fail(rc_, #expr); \
    } while (0)

// ---------------------------------------------------------------------------------------
// force-kernel table and launch planning
// ---------------------------------------------------------------------------------------
#ifndef NB_NT
#define NB_NT 128
#endif
constexpr int kNT = NB_NT;  // threads per CTA (developer A/B: -DNB_NT=256)
constexpr int kCtaScale = kNT / 128;
typedef void (*ForceFn)(const StepParams);

struct KernelEntry {
    int r, kind, eqm, d2;  // kind: 0 force, 1 force + potential, 2 potential only
    ForceFn fn;
    int mid;               // 1: the variant for worlds of few source tiles (near path unrolled deeper)
};

// <threads, R, POT, CTAs/SM, flush chunk, EQM uniform mass (no m*w multiply), D2 z == 0 (reference 2-D uploads), PONLY>
#ifndef NB_FORCE_MINB
#define NB_FORCE_MINB 4
#endif
#define NB_KF(R, E, D) {R, 0, E, D, force_kernel<kNT, R, false, ((R) == 4 ? NB_FORCE_MINB : 4) / kCtaScale, kFlushChunk, (E) != 0, (D) != 0, false>}
// the potential accumulators of R >= 3 targets do not fit 128 registers: those variants run 3 CTAs per SM
#define NB_KP(R, E, D) {R, 1, E, D, force_kernel<kNT, R, true, ((R) >= 3 ? 3 : 4) / kCtaScale, kFlushChunk, (E) != 0, (D) != 0, false>}
#define NB_KD(R, E, D) {R, 2, E, D, force_kernel<kNT, R, true, ((R) >= 3 ? 3 : 4) / kCtaScale, kFlushChunk, (E) != 0, (D) != 0, true>}
#define NB_KM(R, E, D) {R, 0, E, D, force_kernel<kNT, R, false, 4 / kCtaScale, kFlushChunk, (E) != 0, (D) != 0, false, 4>, 1}
#define NB_ALL_ED(M, R) M(R, 0, 0), M(R, 1, 0), M(R, 0, 1), M(R, 1, 1)
const KernelEntry kKernels[] = {
    // ("mid" variants — R = 4 with the near path unrolled 4x deeper for worlds of few source tiles — were measured
    //  and dropped: 1.91 vs 1.80 ms at N = 65536, profiles/r02_time_variants.txt; the selection hook stays)
    NB_ALL_ED(NB_KF, 4), NB_ALL_ED(NB_KF, 3), NB_ALL_ED(NB_KF, 2), NB_ALL_ED(NB_KF, 1),
    // with the fused potential (nb_step_with_diagnostics)
    NB_ALL_ED(NB_KP, 4), NB_ALL_ED(NB_KP, 3), NB_ALL_ED(NB_KP, 2), NB_ALL_ED(NB_KP, 1),
    // potential only (nb_compute_diagnostics, hot loop B)
    NB_ALL_ED(NB_KD, 4), NB_ALL_ED(NB_KD, 2), NB_ALL_ED(NB_KD, 1),
};
#undef NB_KF
#undef NB_KM
#undef NB_KP
#undef NB_KD
#undef NB_ALL_ED
constexpr int kNumKernels = sizeof(kKernels) / sizeof(kKernels[0]);
constexpr long long kMidTiles = 256;  // worlds of <= 2^17 bodies take the "mid" variants where they exist

const KernelEntry* find_kernel(int r, int kind, int eqm, int d2, int mid = 0) {
    const KernelEntry* plain = nullptr;
    for (int k = 0; k < kNumKernels; ++k)
        if (kKernels[k].r == r && kKernels[k].kind == kind && kKernels[k].eqm == eqm && kKernels[k].d2 == d2) {
            if (kKernels[k].mid == mid) return &kKernels[k];
            if (!kKernels[k].mid) plain = &kKernels[k];
        }
    return plain;  // no "mid" variant of this shape: the plain one
}

// Choose targets-per-thread and the source split so the grid fills whole waves of SMs.
// kind: 0 force, 1 force + potential, 2 potential only (stand-alone diagnostics stream).
LaunchPlan make_plan(const nb_ctx* c, long long n_targets, long long n_src_tiles, int kind) {
    LaunchPlan best;
    double best_score = -1.0;
    const int r_opts[] = {4, 3, 2, 1};
    const int eqm_k = (kind == 2 ? c->diag_uniform_mass : c->uniform_mass) ? 1 : 0;
    // NB_FORCE_R (developer knob) only binds where that variant exists (the potential-only set has no R = 3)
    const bool honour_forced = c->forced_r && find_kernel(c->forced_r, kind, eqm_k, c->planar ? 1 : 0) != nullptr;
    for (int q = 0; q < 4; ++q) {
        const int r = r_opts[q];
        if (honour_forced && c->forced_r != r) continue;
        // the stand-alone diagnostics count every body: its uniform-mass variant needs ALL masses equal
        const int eqm = (kind == 2 ? c->diag_uniform_mass : c->uniform_mass) ? 1 : 0, d2 = c->planar ? 1 : 0;
        const KernelEntry* ke = find_kernel(r, kind, eqm, d2, c->n_tiles <= kMidTiles ? 1 : 0);
        if (!ke) continue;
        int per_sm = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ke->fn, kNT, 0) != cudaSuccess || per_sm < 1)
            per_sm = 1;
        const long long slots = (long long)c->sm_count * per_sm;
        const long long n_it = std::max<long long>(1, (n_targets + (long long)kNT * r - 1) / ((long long)kNT * r));
        const int seg_hi = (int)std::min<long long>(kMaxSegments, std::max<long long>(1, n_src_tiles));
        for (int nseg = 1; nseg <= seg_hi; ++nseg) {
            if (c->forced_nseg && nseg != c->forced_nseg && !(c->forced_nseg > seg_hi && nseg == seg_hi)) continue;
            const long long units = n_it * nseg;
            // CTAs do not run in lock-step waves: the launch takes (work / slots) plus a tail of about a
            // tenth of one CTA's duration (measured at N = 2^20: 2 -> 13 segments gains 1.3 %,
            // profiles/r01_time_modes.txt), or (work / units) when the grid cannot fill the slots
            const double per_unit = (double)slots / (double)units;
            const double eff = 1.0 / (std::max(1.0, per_unit) + 0.1 * per_unit);
            // larger R amortises shared-memory reads; an extra segment costs one more FP64 partial per target
            // (measured: no visible cost up to 64 segments even with 4 source tiles per CTA, N = 2^17 ... 2^20,
            // profiles/r01_time_modes.txt)
            // worlds of <= 128 source tiles (N <= 65536): a large share of the (warp, group) visits is "near", a warp
            // with fewer targets per thread has a smaller clump and fewer of them, and more, shorter CTAs balance
            // better: R = 1 measured 4-12 % faster than R = 4 there (profiles/r02_time_r_scan.txt); equal above
            const double r_pref = c->n_tiles <= 128 ? (r == 1 ? 1.0 : (r == 2 ? 0.99 : 0.95))
                                                    : (r == 4 ? 1.0 : (r == 3 ? 0.98 : (r == 2 ? 0.97 : 0.90)));
            const double score = eff * r_pref * (1.0 - 0.00003 * (nseg - 1));
            if (score > best_score) {
                best_score = score;
                best.variant = (int)(ke - kKernels);
                best.r = r;
                best.nseg = nseg;
                best.n_itiles = n_it;
                best.ctas_per_sm = per_sm;
            }
        }
    }
    return best;
}

}  // namespace

// ---------------------------------------------------------------------------------------
// AoS (reference component layout) <-> SoA device state
// ---------------------------------------------------------------------------------------
namespace {

// staging: [pos dim*rows doubles][vel ...][acc ...] for state rows [lo, lo + rows)
__global__ void __launch_bounds__(256) unpack_aos_kernel(const double* pos, const double* vel, const double* acc,
                                                         int dim, long long lo, long long rows, StateView s) {
    const long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (q >= rows) return;
    const long long i = lo + q;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        s.x[k][i] = k < dim ? pos[q * dim + k] : 0.0;
        s.v[k][i] = k < dim ? vel[q * dim + k] : 0.0;
        s.a[k][i] = (acc && k < dim) ? acc[q * dim + k] : 0.0;  // fresh entities: Acceleration{0}, Physics.hpp:97
        s.a0[k][i] = 0.0;  // and PrevAcceleration{0}, :98
    }
}

// which: 0 x, 1 v, 2 a, 3 a0 -> out[(i - lo) * dim + k] for rows [lo,hi)
__global__ void __launch_bounds__(256) pack_aos_kernel(StateView s, int which, int dim, double* out) {
    const long long i = s.lo + blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= s.hi) return;
    double* const* src = which == 0 ? s.x : (which == 1 ? s.v : (which == 2 ? s.a : s.a0));
    for (int k = 0; k < dim; ++k) out[(i - s.lo) * dim + k] = src[k][i];
}

// pair records of non-attracting rows: {x0,x1,y0,y1,z0,z1,m0,m1} = {far, far, 0, 0, 0, 0, 0, 0}
__global__ void __launch_bounds__(256) fill_far_kernel(float* pk, long long n_pairs) {
    const long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (q >= n_pairs) return;
    float4* rec = reinterpret_cast<float4*>(pk) + 2 * q;
    rec[0] = make_float4(kFarPacked, kFarPacked, 0.f, 0.f);
    rec[1] = make_float4(0.f, 0.f, 0.f, 0.f);
}

inline unsigned int blocks_for(long long n, int threads) { return (unsigned int)std::max<long long>(1, (n + threads - 1) / threads); }

void free_bodies(nb_ctx* c) {
    cudaFree(c->state); cudaFree(c->mass); cudaFree(c->pinned); cudaFree(c->order);
    cudaFree(c->part); cudaFree(c->tile_counter); cudaFree(c->diag_part); cudaFree(c->staging);
    cudaFree(c->dpk); cudaFree(c->dpk_lo); cudaFree(c->dhdr); cudaFree(c->sort_tmp); cudaFree(c->dtx);
    c->dtx = nullptr;
    for (int g = 0; g < 2; ++g) {
        cudaFree(c->live2[g]); cudaFree(c->pk[g]); cudaFree(c->pk_lo[g]); cudaFree(c->hdr[g]); cudaFree(c->tx[g]);
        c->tx[g] = nullptr;
        cudaFree(c->sort_keys[g]); cudaFree(c->sort_vals[g]);
        c->live2[g] = nullptr; c->pk[g] = c->pk_lo[g] = nullptr; c->hdr[g] = nullptr;
        c->sort_keys[g] = nullptr; c->sort_vals[g] = nullptr;
    }
    c->state = nullptr; c->mass = nullptr; c->pinned = nullptr; c->order = nullptr;
    c->part = nullptr; c->tile_counter = nullptr; c->diag_part = nullptr; c->staging = nullptr;
    c->dpk = c->dpk_lo = nullptr; c->dhdr = nullptr; c->sort_tmp = nullptr; c->sort_tmp_bytes = 0;
    c->cap = 0; c->staging_bytes = 0; c->part_slots = 0; c->i_cap = 0; c->counter_cap = 0; c->diag_part_rows = 0;
}

int ensure_staging(nb_ctx* c, size_t bytes) {
    if (bytes <= c->staging_bytes) return NB_OK;
    if (c->staging) cudaFree(c->staging);
    c->staging = nullptr;
    c->staging_bytes = 0;
    NB_CUDA(cudaMalloc(&c->staging, bytes));
    c->staging_bytes = bytes;
    return NB_OK;
}

// (Re)size every per-body buffer for n bodies and derive the rank's target range and plans.
int size_for(nb_ctx* c, long long n) {
    const long long tiles = std::max<long long>(1, (n + kTileBodies - 1) / kTileBodies);
    c->tiles_per_rank = (tiles + c->world - 1) / c->world;
    c->n_tiles = c->tiles_per_rank * c->world;
    int64_t lo = 0, hi = 0;
    nb_partition(n, c->rank, c->world, &lo, &hi);
    c->lo = lo;
    c->hi = hi;
    c->n = n;
    c->x_sel = 0;
    const long long rows = c->n_tiles * kTileBodies;
    const long long groups = rows / kGroupBodies;
    const bool realloc_rows = rows > c->cap;
    if (realloc_rows) {
        free_bodies(c);
        NB_CUDA(cudaMalloc(&c->state, sizeof(double) * 15 * rows));
        NB_CUDA(cudaMalloc(&c->mass, sizeof(float) * rows));
        NB_CUDA(cudaMalloc(&c->pinned, rows));
        NB_CUDA(cudaMalloc(&c->order, sizeof(int) * rows));
        for (int g = 0; g < 2; ++g) {
            NB_CUDA(cudaMalloc(&c->live2[g], rows));
            NB_CUDA(cudaMalloc(&c->pk[g], sizeof(float) * 4 * rows));
            NB_CUDA(cudaMalloc(&c->pk_lo[g], sizeof(float) * 4 * rows));
            NB_CUDA(cudaMalloc(&c->hdr[g], sizeof(GroupHdr) * groups));
            NB_CUDA(cudaMalloc(&c->tx[g], sizeof(double) * 3 * rows));
        }
        NB_CUDA(cudaMalloc(&c->dpk, sizeof(float) * 4 * rows));
        NB_CUDA(cudaMalloc(&c->dpk_lo, sizeof(float) * 4 * rows));
        NB_CUDA(cudaMalloc(&c->dhdr, sizeof(GroupHdr) * groups));
        NB_CUDA(cudaMalloc(&c->dtx, sizeof(double) * 3 * rows));
        // Hilbert sort of this rank's rows (cub radix sort, double-buffered keys / values)
        const long long per_rank = c->tiles_per_rank * kTileBodies;
        for (int g = 0; g < 2; ++g) {
            NB_CUDA(cudaMalloc(&c->sort_keys[g], sizeof(unsigned long long) * per_rank));
            NB_CUDA(cudaMalloc(&c->sort_vals[g], sizeof(int) * per_rank));
        }
        size_t tmp = 0;
        NB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp, c->sort_keys[0], c->sort_keys[1], c->sort_vals[0],
                                                c->sort_vals[1], (int)per_rank, 0, 63, c->stream));
        NB_CUDA(cudaMalloc(&c->sort_tmp, tmp ? tmp : 16));
        c->sort_tmp_bytes = tmp;
        c->cap = rows;
    }
    // A re-upload of a world of the same size (the per-frame pattern of a caller that edits bodies, and of the
    // bench's end-to-end arm) re-uses the buffers as they are: every row a kernel reads is rewritten by the
    // upload and the packs that follow (own rows incl. padding by pack_kernel, the other ranks' by the exchange).
    const bool same_shape = !realloc_rows && c->n_sized == n;
    if (!same_shape) {
        c->perm_valid = false;
        NB_CUDA(cudaMemsetAsync(c->state, 0, sizeof(double) * 15 * c->cap, c->stream));
        NB_CUDA(cudaMemsetAsync(c->mass, 0, sizeof(float) * c->cap, c->stream));
        // every row starts as a non-attracting row (hi_x = kFarPacked, m = 0) of a group that is never "near";
        // padding rows stay that way
        for (int g = 0; g < 2; ++g) {
            NB_CUDA(cudaMemsetAsync(c->live2[g], 0, c->cap, c->stream));
            NB_CUDA(cudaMemsetAsync(c->pk_lo[g], 0, sizeof(float) * 4 * c->cap, c->stream));
            NB_CUDA(cudaMemsetAsync(c->hdr[g], 0, sizeof(GroupHdr) * (c->cap / kGroupBodies), c->stream));
            fill_far_kernel<<<blocks_for(c->cap / 2, 256), 256, 0, c->stream>>>(c->pk[g], c->cap / 2);
        }
        NB_CUDA(cudaMemsetAsync(c->dpk_lo, 0, sizeof(float) * 4 * c->cap, c->stream));
        NB_CUDA(cudaMemsetAsync(c->dhdr, 0, sizeof(GroupHdr) * (c->cap / kGroupBodies), c->stream));
        fill_far_kernel<<<blocks_for(c->cap / 2, 256), 256, 0, c->stream>>>(c->dpk, c->cap / 2);
        c->kernel_launches += 3;
    }
    c->n_sized = n;
    NB_CUDA(cudaMemsetAsync(c->pinned, 0, c->cap, c->stream));
    iota_order_kernel<<<blocks_for(c->cap, kStateThreads), kStateThreads, 0, c->stream>>>(c->order, 0, c->cap);
    NB_CUDA(cudaGetLastError());
    c->kernel_launches += 1;
    c->order_valid = false;
    c->packs_since_sort = 0;

    // launch plans: all sources in one launch, or (sharded) own sources first, the rest after the exchange
    const long long n_local = c->hi - c->lo;
    int slots = 1;
    for (int kind = 0; kind < 3; ++kind) {
        c->plans[0][kind] = make_plan(c, n_local, c->n_tiles, kind);
        c->plans[1][kind] = make_plan(c, n_local, c->tiles_per_rank, kind);
        c->plans[2][kind] = make_plan(c, n_local, std::max<long long>(1, c->n_tiles - c->tiles_per_rank), kind);
        for (int w = 0; w < 3; ++w) slots = std::max(slots, c->plans[w][kind].nseg);
    }
    const long long i_cap = ((std::max<long long>(n_local, 1) + 1023) / 1024) * 1024;
    if (slots > c->part_slots || i_cap > c->i_cap) {
        cudaFree(c->part);
        c->part = nullptr;
        NB_CUDA(cudaMalloc(&c->part, sizeof(double) * (size_t)(1 + slots) * 4 * i_cap));
        c->part_slots = slots;
        c->i_cap = i_cap;
    }
    const long long max_itiles = (std::max<long long>(n_local, 1) + kNT - 1) / kNT;  // R = 1 upper bound
    if (max_itiles > c->counter_cap) {
        cudaFree(c->tile_counter);
        cudaFree(c->diag_part);
        c->tile_counter = nullptr;
        c->diag_part = nullptr;
        NB_CUDA(cudaMalloc(&c->tile_counter, sizeof(unsigned int) * max_itiles));
        const long long rows_d = std::max<long long>(max_itiles, blocks_for(n_local, kDiagThreads));
        NB_CUDA(cudaMalloc(&c->diag_part, sizeof(double) * kDiagValues * rows_d));
        c->counter_cap = max_itiles;
        c->diag_part_rows = rows_d;
    }
    NB_CUDA(cudaMemsetAsync(c->tile_counter, 0, sizeof(unsigned int) * c->counter_cap, c->stream));
    return NB_OK;
}

}  // namespace

// ---------------------------------------------------------------------------------------
// lifetime and configuration
// ---------------------------------------------------------------------------------------
extern "C" int nb_default_config(nb_config* out) {
    if (!out) return fail(NB_ERR_INVALID, "nb_default_config: null");
    out->g = 6.67430e-11;            // Constants.hpp:87
    out->softening = 4.0f;           // :89
    out->max_speed = 0.0f;           // :90
    out->integrator = NB_INTEGRATOR_VERLET;  // Config.hpp:25
    out->max_substep = 1.0e9f;       // Constants.hpp:99
    out->max_substeps_per_frame = 200;  // :100
    return NB_OK;
}

namespace { void free_partition(nb_ctx* c); }
static void destroy_one(nb_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->comm_stream) cudaStreamSynchronize(c->comm_stream);
    if (c->nccl_comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->nccl_comm);
    free_bodies(c);
    free_partition(c);
    cudaFree(c->extent);
    cudaFree(c->l2_scratch);
    cudaFree(c->collide_scratch);
    cudaFree(c->bh_scratch);
    cudaFree(c->merge_scratch);
    cudaFree(c->replay_scratch);
    cudaFree(c->near_stats);
    if (c->h_bh_ctl) cudaFreeHost(c->h_bh_ctl);
    cudaFree(c->totals);
    if (c->h_totals) cudaFreeHost(c->h_totals);
    if (c->h_flip) cudaFreeHost(c->h_flip);
    for (cudaEvent_t e : c->prof_events) cudaEventDestroy(e);
    cudaEvent_t evs[] = {c->ev_timer0, c->ev_timer1, c->ev_packed, c->ev_gathered, c->ev_comm0[0], c->ev_comm0[1],
                         c->ev_comm1[0], c->ev_comm1[1], c->ev_own0, c->ev_own1, c->ev_pub, c->ev_bar};
    for (cudaEvent_t e : evs)
        if (e) cudaEventDestroy(e);
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->comm_stream) cudaStreamDestroy(c->comm_stream);
    delete c;
}

static int init_ctx(nb_ctx* c, int world) {
    nb_default_config(&c->cfg);
    int lo_prio = 0, hi_prio = 0;
    cudaDeviceGetStreamPriorityRange(&lo_prio, &hi_prio);
    NB_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    NB_CUDA(cudaStreamCreateWithPriority(&c->comm_stream, cudaStreamNonBlocking, hi_prio));
    NB_CUDA(cudaEventCreate(&c->ev_timer0));
    NB_CUDA(cudaEventCreate(&c->ev_timer1));
    NB_CUDA(cudaEventCreateWithFlags(&c->ev_packed, cudaEventDisableTiming));
    NB_CUDA(cudaEventCreateWithFlags(&c->ev_gathered, cudaEventDisableTiming));
    NB_CUDA(cudaEventCreateWithFlags(&c->ev_pub, cudaEventDisableTiming));
    NB_CUDA(cudaEventCreateWithFlags(&c->ev_bar, cudaEventDisableTiming));
    for (int g = 0; g < 2; ++g) {
        NB_CUDA(cudaEventCreate(&c->ev_comm0[g]));
        NB_CUDA(cudaEventCreate(&c->ev_comm1[g]));
    }
    NB_CUDA(cudaEventCreate(&c->ev_own0));
    NB_CUDA(cudaEventCreate(&c->ev_own1));
    NB_CUDA(cudaMalloc(&c->extent, sizeof(unsigned int) * (kExtentSlots + 4)));
    NB_CUDA(cudaMemset(c->extent, 0, sizeof(unsigned int) * (kExtentSlots + 4)));
    // [0,16): this rank's fixed-order sums; [16, 16 + kDiagValues * world): every rank's, all-gathered
    const size_t totals_len = 48 + (size_t)kDiagValues * world + 16;
    NB_CUDA(cudaMalloc(&c->totals, sizeof(double) * totals_len));
    NB_CUDA(cudaMallocHost(&c->h_totals, sizeof(double) * totals_len));
    NB_CUDA(cudaMallocHost(reinterpret_cast<void**>(&c->h_flip), sizeof(int)));
    *c->h_flip = 0;
    if (const char* v = getenv("NB_FORCE_R")) c->forced_r = atoi(v);
    if (const char* v = getenv("NB_FORCE_NSEG")) c->forced_nseg = atoi(v);
    if (const char* v = getenv("NB_RESORT_EVERY")) c->resort_every = std::max(1, atoi(v));
    if (const char* v = getenv("NB_FORCE_GENERIC")) c->force_generic = atoi(v) != 0;
    if (const char* v = getenv("NB_CURVE_PARTITION")) c->curve_partition = atoi(v) != 0;
    return NB_OK;
}

static int create_common(int device, int rank, int world, nb_ctx** out) {
    if (!out) return fail(NB_ERR_INVALID, "nb_create: null out pointer");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count <= 0)
        return fail(NB_ERR_NO_DEVICE, std::string("no CUDA device (this library has no CPU path): ") +
                                          (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0"));
    if (device < 0 || device >= count) return fail(NB_ERR_INVALID, "nb_create: device index out of range");
    NB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    NB_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return fail(NB_ERR_NO_DEVICE, std::string("device '") + prop.name + "' is not sm_100-class; kernels are built for sm_100a only");
    nb_ctx* c = new nb_ctx();
    c->device = device;
    c->rank = rank;
    c->world = world;
    c->sm_count = prop.multiProcessorCount;
    const int rc = init_ctx(c, world);
    if (rc != NB_OK) {  // nothing half-built survives a failed creation
        destroy_one(c);
        return rc;
    }
    *out = c;
    return NB_OK;
}

extern "C" int nb_create(int device, nb_ctx** out) { return create_common(device, 0, 1, out); }

extern "C" int nb_comm_unique_id(void* out_id_128) {
    if (!out_id_128) return fail(NB_ERR_INVALID, "nb_comm_unique_id: null");
    NB_TRY(load_nccl());
    NcclId id;
    NB_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(out_id_128, id.bytes, 128);
    return NB_OK;
}

extern "C" int nb_create_sharded(int device, int rank, int world, const void* nccl_unique_id, nb_ctx** out) {
    if (world < 1 || rank < 0 || rank >= world) return fail(NB_ERR_INVALID, "nb_create_sharded: bad rank/world");
    if (world > 1 && !nccl_unique_id) return fail(NB_ERR_INVALID, "nb_create_sharded: null id");
    NB_TRY(create_common(device, rank, world, out));
    if (world > 1) {
        NB_TRY(load_nccl());
        NcclId id;
        memcpy(id.bytes, nccl_unique_id, 128);
        NB_NCCL(g_nccl.CommInitRank(&(*out)->nccl_comm, world, id, rank));
    }
    return NB_OK;
}

// One process, several GPUs (the reference is one process, src/main.cpp:100-143): member i owns the i-th
// slice of the targets exactly as rank i of a sharded context does (nb_partition), but the members exchange
// the packed stream through peer memory — the packing kernels store straight into every member's buffers
// over NVLink — and order themselves with CUDA events. No NCCL, no torchrun.
extern "C" int nb_create_multi(const int* devices, int n_dev, nb_ctx** out) {
    if (!out || !devices) return fail(NB_ERR_INVALID, "nb_create_multi: null");
    *out = nullptr;
    if (n_dev < 1 || n_dev > kMaxPeers) return fail(NB_ERR_INVALID, "nb_create_multi: 1..8 devices");
    for (int i = 0; i < n_dev; ++i)
        for (int j = 0; j < i; ++j)
            if (devices[i] == devices[j]) return fail(NB_ERR_INVALID, "nb_create_multi: a device is listed twice");
    if (n_dev == 1) return create_common(devices[0], 0, 1, out);
    nb_group* g = new nb_group();
    auto undo = [&]() {
        for (nb_ctx* m : g->m) destroy_one(m);
        delete g;
    };
    for (int i = 0; i < n_dev; ++i) {
        nb_ctx* m = nullptr;
        const int rc = create_common(devices[i], i, n_dev, &m);
        if (rc != NB_OK) {
            undo();
            return rc;
        }
        g->m.push_back(m);
    }
    for (int i = 0; i < n_dev; ++i) {
        cudaSetDevice(devices[i]);
        for (int j = 0; j < n_dev; ++j) {
            if (i == j) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[i], devices[j]);
            if (!can) {
                undo();
                return fail(NB_ERR_NO_DEVICE, "nb_create_multi: devices " + std::to_string(devices[i]) + " and " +
                                                  std::to_string(devices[j]) + " cannot access each other's memory");
            }
            const cudaError_t e = cudaDeviceEnablePeerAccess(devices[j], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                undo();
                return fail(NB_ERR_CUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e));
            }
            cudaGetLastError();  // clear "already enabled"
        }
    }
    for (nb_ctx* m : g->m) m->group = g;
    *out = g->m[0];
    return NB_OK;
}

extern "C" int nb_num_devices(nb_ctx* c) { return c ? (c->group ? (int)c->group->m.size() : 1) : -1; }

extern "C" void nb_destroy(nb_ctx* c) {
    if (!c) return;
    if (c->group) {
        nb_group* g = c->group;
        for (nb_ctx* m : g->m) {
            cudaSetDevice(m->device);
            cudaStreamSynchronize(m->stream);
        }
        for (nb_ctx* m : g->m) destroy_one(m);
        delete g;
        return;
    }
    destroy_one(c);
}

extern "C" int nb_set_config(nb_ctx* c, const nb_config* cfg) {
    if (!c || !cfg) return fail(NB_ERR_INVALID, "nb_set_config: null");
    if (cfg->integrator != NB_INTEGRATOR_EULER && cfg->integrator != NB_INTEGRATOR_VERLET)
        return fail(NB_ERR_INVALID, "nb_set_config: integrator must be 0 (Euler) or 1 (Verlet)");
    const bool repack = cfg->softening != c->cfg.softening || cfg->g != c->cfg.g;
    for (nb_ctx* s : Team(c)) {
        s->cfg = *cfg;
        if (repack) {  // accelerations depend on G and epsilon, and so does the drift prepared with them
            s->acc_valid = false;
            s->drop_speculation();
        }
    }
    return NB_OK;
}

extern "C" int nb_get_config(nb_ctx* c, nb_config* out) {
    if (!c || !out) return fail(NB_ERR_INVALID, "nb_get_config: null");
    *out = c->cfg;
    return NB_OK;
}

extern "C" int64_t nb_num_bodies(nb_ctx* c) { return c ? c->n : -1; }

extern "C" int nb_local_range(nb_ctx* c, int64_t* lo, int64_t* hi) {
    if (!c) return fail(NB_ERR_INVALID, "nb_local_range: null");
    if (lo) *lo = c->group ? 0 : c->lo;  // one process holds every member's rows
    if (hi) *hi = c->group ? c->n : c->hi;
    return NB_OK;
}

extern "C" int nb_partition(int64_t n, int rank, int world, int64_t* lo, int64_t* hi) {
    if (n < 0 || world < 1 || rank < 0 || rank >= world) return fail(NB_ERR_INVALID, "nb_partition: bad arguments");
    const long long tiles = std::max<long long>(1, (n + kTileBodies - 1) / kTileBodies);
    const long long per_rank = (tiles + world - 1) / world;
    if (lo) *lo = std::min<long long>(n, (long long)rank * per_rank * kTileBodies);
    if (hi) *hi = std::min<long long>(n, (long long)(rank + 1) * per_rank * kTileBodies);
    return NB_OK;
}

extern "C" int nb_substep_plan(float dt, float max_substep, int32_t max_substeps_per_frame, float* dt_sub) {
    // Physics.hpp:268-271, float arithmetic throughout
    const float cap = std::max(1e-6f, max_substep);
    int n_steps = static_cast<int>(std::ceil(dt / cap));
    n_steps = std::max(1, std::min(n_steps, std::max(1, (int)max_substeps_per_frame)));
    if (dt_sub) *dt_sub = dt / static_cast<float>(n_steps);
    return n_steps;
}

// ---------------------------------------------------------------------------------------
// exchange between ranks (target-partitioned contexts)
// ---------------------------------------------------------------------------------------
namespace {

// Where member s's packs of generation parity `par` go: its own buffers first, then every peer's.
PeerSet stream_out(nb_ctx* s, int par) {
    PeerSet o;
    memset(&o, 0, sizeof(o));
    o.n = 0;
    auto add = [&](nb_ctx* t) {
        o.pk[o.n] = t->pk[par]; o.lo[o.n] = t->pk_lo[par]; o.hdr[o.n] = t->hdr[par]; o.extent[o.n] = t->extent;
        o.n++;
    };
    add(s);
    if (s->group)
        for (nb_ctx* t : s->group->m)
            if (t != s) add(t);
    return o;
}
PeerSet diag_out(nb_ctx* s) {
    PeerSet o = stream_out(s, 0);
    int q = 0;
    auto set = [&](nb_ctx* t) { o.pk[q] = t->dpk; o.lo[q] = t->dpk_lo; o.hdr[q] = t->dhdr; q++; };
    set(s);
    if (s->group)
        for (nb_ctx* t : s->group->m)
            if (t != s) set(t);
    return o;
}

TxOut tx_out(double* base, long long cap) {
    TxOut t;
    for (int k = 0; k < 3; ++k) t.x[k] = base + (long long)k * cap;
    return t;
}

// Every member's stream waits for everything enqueued so far on every other member's stream.
int team_barrier(nb_ctx* c) {
    if (!c->group) return NB_OK;
    for (nb_ctx* s : Team(c)) {
        NB_CUDA(cudaSetDevice(s->device));
        NB_CUDA(cudaEventRecord(s->ev_bar, s->stream));
    }
    for (nb_ctx* s : Team(c)) {
        NB_CUDA(cudaSetDevice(s->device));
        for (nb_ctx* t : Team(c))
            if (t != s) NB_CUDA(cudaStreamWaitEvent(s->stream, t->ev_bar, 0));
    }
    return NB_OK;
}

// All-gather generation `gen`'s j-stream slices (hi parts, residuals, group headers) in place and
// max-reduce the rotating extent slot, on the comm stream, after everything enqueued so far on the main stream.
// (NCCL-sharded contexts only: the members of a multi-GPU context wrote each other's buffers already.)
int exchange_packed(nb_ctx* c, int gen, bool main_waits) {
    if (c->world <= 1 || c->group) return NB_OK;
    NB_CUDA(cudaEventRecord(c->ev_packed, c->stream));
    NB_CUDA(cudaStreamWaitEvent(c->comm_stream, c->ev_packed, 0));
    const size_t rows = (size_t)c->tiles_per_rank * kTileBodies;
    const size_t count = rows * 4;  // floats per rank
    const size_t hdr_count = rows / kGroupBodies * (sizeof(GroupHdr) / sizeof(float));
    float* buf = c->pk[gen & 1];
    float* lob = c->pk_lo[gen & 1];
    float* hdb = reinterpret_cast<float*>(c->hdr[gen & 1]);
    if (c->profile_comm) NB_CUDA(cudaEventRecord(c->ev_comm0[gen & 1], c->comm_stream));
    NB_NCCL(g_nccl.GroupStart());
    NB_NCCL(g_nccl.AllGather(buf + (size_t)c->rank * count, buf, count, kNcclFloat32, c->nccl_comm, c->comm_stream));
    NB_NCCL(g_nccl.AllGather(lob + (size_t)c->rank * count, lob, count, kNcclFloat32, c->nccl_comm, c->comm_stream));
    NB_NCCL(g_nccl.AllGather(hdb + (size_t)c->rank * hdr_count, hdb, hdr_count, kNcclFloat32, c->nccl_comm,
                             c->comm_stream));
    // only the slot this generation's epilogue wrote: the others are being read by kernels in flight
    unsigned int* slot = c->extent + (gen % kExtentSlots);
    NB_NCCL(g_nccl.AllReduce(slot, slot, 1, kNcclUint32, kNcclMax, c->nccl_comm, c->comm_stream));
    NB_NCCL(g_nccl.GroupEnd());
    if (c->profile_comm) NB_CUDA(cudaEventRecord(c->ev_comm1[gen & 1], c->comm_stream));
    NB_CUDA(cudaEventRecord(c->ev_gathered, c->comm_stream));
    if (main_waits) NB_CUDA(cudaStreamWaitEvent(c->stream, c->ev_gathered, 0));
    return NB_OK;
}

// Multi-GPU context: every member announces that its packing launch is enqueued (ev_pub). Called once per
// generation, AFTER every member's launches of the pass, so a member's wait for the previous generation
// (consume_published) never picks up this newer record and serialises the team.
int publish_all(nb_ctx* c) {
    if (!c->group) return NB_OK;
    for (nb_ctx* s : Team(c)) {
        NB_CUDA(cudaSetDevice(s->device));
        NB_CUDA(cudaEventRecord(s->ev_pub, s->stream));
    }
    return NB_OK;
}
// ... and member s's stream waits for every other member's announcement before it reads their rows.
int consume_published(nb_ctx* s) {
    if (s->group) {
        for (nb_ctx* t : s->group->m)
            if (t != s) NB_CUDA(cudaStreamWaitEvent(s->stream, t->ev_pub, 0));
    } else if (s->world > 1) {
        NB_CUDA(cudaStreamWaitEvent(s->stream, s->ev_gathered, 0));
    }
    return NB_OK;
}

// Order this rank's stream rows along a Hilbert curve of the current positions (extent slot `slot` holds the
// cube). Rows of other ranks keep whatever order their owner chose: only the owner reads `order`.
int sort_stream(nb_ctx* c, int slot) {
    const long long n_local = c->hi - c->lo;
    if (n_local <= 0) return NB_OK;
    if (c->dim == 2)
        curve_key_kernel<2><<<blocks_for(n_local, kStateThreads), kStateThreads, 0, c->stream>>>(
            c->view(), c->extent, slot, c->sort_keys[0], c->sort_vals[0]);
    else
        curve_key_kernel<3><<<blocks_for(n_local, kStateThreads), kStateThreads, 0, c->stream>>>(
            c->view(), c->extent, slot, c->sort_keys[0], c->sort_vals[0]);
    size_t tmp = c->sort_tmp_bytes;
    NB_CUDA(cub::DeviceRadixSort::SortPairs(c->sort_tmp, tmp, c->sort_keys[0], c->sort_keys[1], c->sort_vals[0],
                                            c->order + c->lo, (int)n_local, 0, 63, c->stream));
    c->kernel_launches += 2;  // + cub's own passes
    c->sorts++;
    return NB_OK;
}

bool uses_partition(const nb_ctx* c);
int repartition(nb_ctx* c);

// Measure the extent of the current positions, open a new generation and pack it (the whole team).
int prepare_packed(nb_ctx* c) {
    const Team team(c);
    // multi-GPU contexts: when the stream is due for a re-sort, the ranks also regroup along the curve of the
    // current positions (partition.cuh); every rank takes this decision from the same counters
    if (uses_partition(c) && c->perm_valid && c->order_valid && c->packs_since_sort >= c->resort_every &&
        c->n > kGroupBodies)
        NB_TRY(repartition(c));
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        s->gen += 1;
        s->drop_speculation();
        NB_CUDA(cudaMemsetAsync(s->extent, 0, sizeof(unsigned int) * kExtentSlots, s->stream));
    }
    NB_TRY(team_barrier(c));  // nobody's peer atomics may land before the memset of the slot they hit
    const int g = c->gen;
    const int slot_cur = (g + kExtentSlots - 1) % kExtentSlots;
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        const long long n_local = s->hi - s->lo;
        const unsigned int grid = (unsigned int)std::min<long long>(blocks_for(n_local, kStateThreads), 4 * s->sm_count);
        extent_kernel<<<grid, kStateThreads, 0, s->stream>>>(s->view(), stream_out(s, g & 1), slot_cur, g % kExtentSlots);
        s->kernel_launches++;
    }
    if (c->world > 1 && !c->group) {
        // the pack below needs the GLOBAL extent: reduce it first (blocking the main stream)
        NB_CUDA(cudaEventRecord(c->ev_packed, c->stream));
        NB_CUDA(cudaStreamWaitEvent(c->comm_stream, c->ev_packed, 0));
        NB_NCCL(g_nccl.AllReduce(c->extent, c->extent, kExtentSlots, kNcclUint32, kNcclMax, c->nccl_comm,
                                 c->comm_stream));
        NB_CUDA(cudaEventRecord(c->ev_gathered, c->comm_stream));
        NB_CUDA(cudaStreamWaitEvent(c->stream, c->ev_gathered, 0));
    }
    NB_TRY(team_barrier(c));  // every member's extent kernel has reached every copy of the slots
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        const long long n_local = s->hi - s->lo;
        if (!s->order_valid || s->packs_since_sort >= s->resort_every) {
            // a world of one group gains nothing from an ordering: it keeps upload order (and with it the lane
            // pairing of the reference's mirror-symmetric three-body seed, whose pulls on the centre then cancel exactly)
            if (s->n > kGroupBodies) NB_TRY(sort_stream(s, slot_cur));
            s->order_valid = true;
            s->packs_since_sort = 0;
        }
        if (n_local > 0) {
            const long long groups = (s->tiles_per_rank * kTileBodies) / kGroupBodies;
            pack_kernel<false><<<(unsigned int)groups, kGroupBodies, 0, s->stream>>>(
                s->view(), s->order, stream_out(s, g & 1), tx_out(s->tx[g & 1], s->cap), s->live2[g & 1], s->extent, slot_cur,
                s->mass_exp, s->cfg.g, s->cfg.softening);
            s->kernel_launches++;
        }
        NB_CUDA(cudaGetLastError());
    }
    NB_TRY(exchange_packed(c, g, true));
    NB_TRY(publish_all(c));
    for (nb_ctx* s : team) s->packed_valid = true;
    return NB_OK;
}

int prof_begin(nb_ctx* c) {
    if (!c->profile) return NB_OK;
    if (c->prof_used + 2 > c->prof_events.size()) {
        for (int k = 0; k < 2; ++k) {
            cudaEvent_t e;
            NB_CUDA(cudaEventCreate(&e));
            c->prof_events.push_back(e);
        }
    }
    NB_CUDA(cudaEventRecord(c->prof_events[c->prof_used], c->stream));
    return NB_OK;
}
int prof_end(nb_ctx* c) {
    if (!c->profile) return NB_OK;
    NB_CUDA(cudaEventRecord(c->prof_events[c->prof_used + 1], c->stream));
    c->prof_used += 2;
    return NB_OK;
}

// One launch of the force kernel. which: 0 all sources, 1 own sources (partial only),
// 2 remote sources (+ partial from launch 1, + epilogue). kind: 0 force, 1 force + potential,
// 2 potential only over the stand-alone diagnostics stream.
int launch_force(nb_ctx* c, int which, int mode, int kind, double dt) {
    const long long n_local = c->hi - c->lo;
    if (n_local <= 0) return NB_OK;
    const LaunchPlan& pl = c->plans[which][kind];
    const KernelEntry& ke = kKernels[pl.variant];
    StepParams p;
    const StateView s = c->view();
    for (int k = 0; k < 3; ++k) {
        p.x[k] = s.x[k]; p.v[k] = s.v[k]; p.a[k] = s.a[k]; p.a0[k] = s.a0[k];
        p.xs[k] = c->pos_slot(c->x_sel ^ 1, k);
    }
    p.mass = c->mass; p.pinned = c->pinned;
    p.live = c->live2[c->gen & 1];
    p.live_next = c->live2[(c->gen + 1) & 1];
    p.order = c->order;
    for (int k = 0; k < 3; ++k) {
        p.tx_cur[k] = (kind == 2 ? c->dtx : c->tx[c->gen & 1]) + (long long)k * c->cap;
        p.tx_next[k] = c->tx[(c->gen + 1) & 1] + (long long)k * c->cap;
    }
    if (kind == 2) {
        p.pk_cur = c->dpk; p.lo_cur = c->dpk_lo; p.hdr_cur = c->dhdr;
        p.slot_cur = kDiagExtentSlot; p.slot_nxt = kDiagExtentSlot;
    } else {
        p.pk_cur = c->pk[c->gen & 1]; p.lo_cur = c->pk_lo[c->gen & 1]; p.hdr_cur = c->hdr[c->gen & 1];
        p.slot_cur = (c->gen + kExtentSlots - 1) % kExtentSlots; p.slot_nxt = c->gen % kExtentSlots;
    }
    p.out = stream_out(c, (c->gen + 1) & 1);
    p.extent = c->extent;
    p.part = c->part;
    p.tile_counter = c->tile_counter;
    p.diag_part = c->diag_part;
    p.stats = c->near_stats;
    p.filter_flip = c->world == 1 ? c->h_flip : nullptr;  // single-GPU contexts (ranks would have to agree on it)
    p.n = c->n;
    p.n_tiles = c->n_tiles;
    p.i_lo = c->lo; p.i_hi = c->hi;
    p.i_cap = c->i_cap;
    p.row_cap = c->lo + c->tiles_per_rank * kTileBodies;
    const long long own_lo = (long long)c->rank * c->tiles_per_rank;
    p.hole_lo = c->n_tiles; p.hole_len = 0;
    if (which == 0) { p.j_tile_lo = 0; p.j_tile_hi = c->n_tiles; }
    else if (which == 1) { p.j_tile_lo = own_lo; p.j_tile_hi = own_lo + c->tiles_per_rank; }
    else { p.j_tile_lo = 0; p.j_tile_hi = c->n_tiles - c->tiles_per_rank; p.hole_lo = own_lo; p.hole_len = c->tiles_per_rank; }
    p.nseg = pl.nseg;
    p.gen = c->gen;
    p.mode = which == 1 ? (int)kPartialOnly : mode;
    p.accumulate = which == 2 ? 1 : 0;
    p.mass_exp = c->mass_exp;
    p.softening = c->cfg.softening;
    p.max_speed = c->cfg.max_speed;
    p.g = c->cfg.g;
    p.dt = dt;
    p.half_dt2 = 0.5 * dt * dt;  // Physics.hpp:300
    p.eqm_mass = ke.eqm ? c->uniform_packed_mass : 1.0;
    const long long n_it = (n_local + (long long)kNT * pl.r - 1) / ((long long)kNT * pl.r);
    NB_TRY(prof_begin(c));
    ke.fn<<<(unsigned int)(n_it * pl.nseg), kNT, 0, c->stream>>>(p);
    NB_CUDA(cudaGetLastError());
    NB_TRY(prof_end(c));
    c->kernel_launches++;
    c->interactions += (double)n_local * (double)(p.j_tile_hi - p.j_tile_lo) * kTileBodies;
    if (c->near_stats)  // every warp of every target tile visits every source group of the launch once
        c->near_visits += n_it * (kNT / 32) * (p.j_tile_hi - p.j_tile_lo) * kGroupsPerTile;
    if (which != 1 && kind != 0) {
        c->step_diag_rows = n_it;
        c->step_diag_valid = kind == 1;
    }
    return NB_OK;
}

// Own sources need no communication: in a sharded context a pass starts on them while the exchange of the
// stream it reads may still be in flight, then finishes with the remote ones. The whole team, phase by phase.
int launch_pass(nb_ctx* c, int mode, int kind, double dt) {
    const Team team(c);
    const bool split = c->world > 1 && c->n_tiles > c->tiles_per_rank;
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        if (split) {
            if (s->profile_comm) NB_CUDA(cudaEventRecord(s->ev_own0, s->stream));
            NB_TRY(launch_force(s, 1, mode, kind, dt));
            if (s->profile_comm) NB_CUDA(cudaEventRecord(s->ev_own1, s->stream));
        } else {
            NB_TRY(launch_force(s, 0, mode, kind, dt));
        }
    }
    if (split) {
        for (nb_ctx* s : team) {
            NB_CUDA(cudaSetDevice(s->device));
            NB_TRY(consume_published(s));
            NB_TRY(launch_force(s, 2, mode, kind, dt));
        }
    }
    return NB_OK;
}

// One force evaluation over all sources with epilogue `mode`. kEuler commits the generation its epilogue
// packed; kVerletKickDrift leaves it speculative (commit_drift()).
int force_pass(nb_ctx* c, int mode, bool pot, double dt) {
    if (!c->packed_valid) NB_TRY(prepare_packed(c));
    const Team team(c);
    // a plain re-evaluation at the same positions through the same kernel reproduces Acceleration bit for
    // bit and leaves a prepared drift usable; anything else (another kernel variant, a move) does not
    if (!(mode == kForceOnly && !pot && c->spec_kind == 0))
        for (nb_ctx* s : team) s->drop_speculation();
    NB_TRY(launch_pass(c, mode, pot ? 1 : 0, dt));
    for (nb_ctx* s : team) s->force_evals++;
    if (mode == kEuler) {
        for (nb_ctx* s : team) {
            s->gen += 1;
            s->packs_since_sort++;
            if (s->packs_since_sort >= s->resort_every) s->packed_valid = false;  // next pass re-sorts and re-packs
        }
        NB_TRY(exchange_packed(c, c->gen, false));  // consumed by the next pass after its own-sources launch
        NB_TRY(publish_all(c));
    } else if (mode == kVerletKickDrift) {
        for (nb_ctx* s : team) {
            s->spec_valid = true;
            s->spec_h = dt;
            s->spec_kind = pot ? 1 : 0;
        }
        NB_TRY(exchange_packed(c, c->gen + 1, false));
        NB_TRY(publish_all(c));
    }
    return NB_OK;
}

// The drift the last kick epilogue prepared becomes the state: Position <- shadow positions, generation + 1
// (its j-stream, group headers, gather-filter flags and extent slot were written by that epilogue).
void commit_drift(nb_ctx* c) {
    for (nb_ctx* s : Team(c)) {
        s->x_sel ^= 1;
        s->gen += 1;
        s->packs_since_sort++;
        s->spec_valid = false;
        s->step_diag_valid = false;  // the fused potential described the positions before this drift
        if (s->packs_since_sort >= s->resort_every) s->packed_valid = false;  // this pass re-sorts and re-packs
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------
// ownership along one curve (partition.cuh): the exchange buffer and the permutation
// ---------------------------------------------------------------------------------------
namespace {

bool uses_partition(const nb_ctx* c) { return c->world > 1 && c->curve_partition; }

void free_partition(nb_ctx* c) {
    cudaFree(c->perm); cudaFree(c->inv_perm); cudaFree(c->xb); cudaFree(c->xb_mass); cudaFree(c->xb_pinned);
    cudaFree(c->part_tmp);
    for (int g = 0; g < 2; ++g) { cudaFree(c->part_keys[g]); cudaFree(c->part_vals[g]); c->part_keys[g] = nullptr; c->part_vals[g] = nullptr; }
    c->perm = c->inv_perm = nullptr; c->xb = nullptr; c->xb_mass = nullptr; c->xb_pinned = nullptr; c->part_tmp = nullptr;
    c->part_tmp_bytes = 0; c->part_cap = 0; c->perm_valid = false;
}

// buffers for c->cap rows (called after size_for, with the device current)
int ensure_partition_buffers(nb_ctx* c) {
    if (!uses_partition(c) || c->part_cap >= c->cap) return NB_OK;
    free_partition(c);
    const size_t rows = (size_t)c->cap;
    NB_CUDA(cudaMalloc(&c->perm, sizeof(int) * rows));
    NB_CUDA(cudaMalloc(&c->inv_perm, sizeof(int) * rows));
    NB_CUDA(cudaMalloc(&c->xb, sizeof(double) * kXbArrays * rows));
    NB_CUDA(cudaMalloc(&c->xb_mass, sizeof(float) * rows));
    NB_CUDA(cudaMalloc(&c->xb_pinned, rows));
    for (int g = 0; g < 2; ++g) {
        NB_CUDA(cudaMalloc(&c->part_keys[g], sizeof(unsigned long long) * rows));
        NB_CUDA(cudaMalloc(&c->part_vals[g], sizeof(int) * rows));
    }
    size_t tmp = 0;
    NB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp, c->part_keys[0], c->part_keys[1], c->part_vals[0],
                                            c->part_vals[1], (int)rows, 0, 64, c->stream));
    NB_CUDA(cudaMalloc(&c->part_tmp, tmp ? tmp : 16));
    c->part_tmp_bytes = tmp;
    c->part_cap = c->cap;
    return NB_OK;
}

// every device's exchange buffer as seen from member s: its own first, then the peers' (one process, several GPUs)
XbOut xb_out(nb_ctx* s) {
    XbOut o;
    memset(&o, 0, sizeof(o));
    auto add = [&](nb_ctx* t) { o.st[o.n] = t->xb; o.mass[o.n] = t->xb_mass; o.pinned[o.n] = t->xb_pinned; o.n++; };
    add(s);
    if (s->group)
        for (nb_ctx* t : s->group->m)
            if (t != s) add(t);
    return o;
}

// NCCL contexts: all-gather (in place) the slices [rank * rows, (rank + 1) * rows) of XB arrays [a0, a1) and,
// when asked, of the mass / pinned rows. `src` != nullptr: array a's slice comes from src[a] instead (row 0 of
// the slice). Collective: every rank calls it, rows or not.
int allgather_xb(nb_ctx* c, int a0, int a1, bool with_mass, double* const* src = nullptr) {
    const size_t rows = (size_t)c->tiles_per_rank * kTileBodies;
    // an exchange of the packed stream may still be running on the communication stream: one collective at a time
    NB_CUDA(cudaStreamWaitEvent(c->stream, c->ev_gathered, 0));
    NB_NCCL(g_nccl.GroupStart());
    for (int a = a0; a < a1; ++a) {
        double* arr = c->xb + (size_t)a * c->cap;
        const double* send = src ? src[a] : arr + (size_t)c->rank * rows;
        NB_NCCL(g_nccl.AllGather(send, arr, rows, kNcclFloat64, c->nccl_comm, c->stream));
    }
    if (with_mass) {
        NB_NCCL(g_nccl.AllGather(c->xb_mass + (size_t)c->rank * rows, c->xb_mass, rows, kNcclFloat32, c->nccl_comm, c->stream));
        NB_NCCL(g_nccl.AllGather(c->xb_pinned + (size_t)c->rank * rows, c->xb_pinned, rows, kNcclUint8, c->nccl_comm, c->stream));
    }
    NB_NCCL(g_nccl.GroupEnd());
    return NB_OK;
}

// Every rank's XB <- the rows each rank owns (internal order): state arrays [0, n_arrays) and, when asked, masses
// and pinned flags.
int publish_owned(nb_ctx* c, int n_arrays, bool with_mass) {
    const Team team(c);
    if (c->group) {
        NB_TRY(team_barrier(c));  // nobody still reads the previous contents
        for (nb_ctx* s : team) {
            NB_CUDA(cudaSetDevice(s->device));
            const long long n_local = s->hi - s->lo;
            if (n_local <= 0) continue;
            publish_state_kernel<<<blocks_for(n_local, kStateThreads), kStateThreads, 0, s->stream>>>(
                s->view(), n_arrays, with_mass, s->cap, xb_out(s));
            NB_CUDA(cudaGetLastError());
            s->kernel_launches++;
        }
        return team_barrier(c);
    }
    const size_t row0 = (size_t)c->rank * c->tiles_per_rank * kTileBodies;
    double* src[kXbArrays];
    for (int k = 0; k < 3; ++k) {
        src[k] = c->pos_slot(c->x_sel, k) + row0;
        src[3 + k] = c->sx(3 + k) + row0;
        src[6 + k] = c->sx(6 + k) + row0;
        src[9 + k] = c->sx(9 + k) + row0;
    }
    if (with_mass) {  // in-place slices of the mass / pinned gathers
        const size_t rows = (size_t)c->tiles_per_rank * kTileBodies;
        NB_CUDA(cudaMemcpyAsync(c->xb_mass + row0, c->mass + row0, sizeof(float) * rows, cudaMemcpyDeviceToDevice, c->stream));
        NB_CUDA(cudaMemcpyAsync(c->xb_pinned + row0, c->pinned + row0, rows, cudaMemcpyDeviceToDevice, c->stream));
    }
    return allgather_xb(c, 0, n_arrays, with_mass, src);
}

// perm <- all n bodies along one Hilbert curve of the positions in XB (arrays 0-5: x, v; xb_mass).
// regroup == false: XB rows are upload rows (an upload). regroup == true: XB rows are the internal rows held until
// now (publish_owned): the rows this rank owns from now on are fetched from XB as well, with everything that
// belongs to a body. Every rank computes the same permutation from the same data: no communication.
int compute_perm(nb_ctx* s, bool regroup) {
    const long long n = s->n;
    NB_CUDA(cudaMemsetAsync(s->extent + kPartitionExtentSlot, 0, sizeof(unsigned int), s->stream));
    const unsigned int grid = (unsigned int)std::min<long long>(blocks_for(n, kStateThreads), 4 * s->sm_count);
    xb_extent_kernel<<<grid, kStateThreads, 0, s->stream>>>(s->xb, s->cap, n, s->xb_mass, s->extent);
    const int* row_of = regroup ? s->inv_perm : nullptr;
    if (s->dim == 2)
        xb_curve_key_kernel<2><<<blocks_for(n, kStateThreads), kStateThreads, 0, s->stream>>>(
            s->xb, s->cap, n, s->xb_mass, row_of, s->extent, s->part_keys[0], s->part_vals[0]);
    else
        xb_curve_key_kernel<3><<<blocks_for(n, kStateThreads), kStateThreads, 0, s->stream>>>(
            s->xb, s->cap, n, s->xb_mass, row_of, s->extent, s->part_keys[0], s->part_vals[0]);
    size_t tmp = s->part_tmp_bytes;
    NB_CUDA(cub::DeviceRadixSort::SortPairs(s->part_tmp, tmp, s->part_keys[0], s->part_keys[1], s->part_vals[0],
                                            s->part_vals[1], (int)n, 0, 64, s->stream));
    const int* sorted = s->part_vals[1];  // XB row of the body at each place of the curve
    const int* through = nullptr;
    if (regroup) {
        const long long n_local = s->hi - s->lo;
        if (n_local > 0) {
            gather_owned_kernel<<<blocks_for(n_local, kStateThreads), kStateThreads, 0, s->stream>>>(
                s->xb, s->xb_mass, s->xb_pinned, s->cap, sorted, nullptr, kXbArrays, s->view(), s->mass, s->pinned);
            s->kernel_launches++;
        }
        // XB row -> upload row is the permutation held until now; set_perm_kernel overwrites it in place
        int* held = reinterpret_cast<int*>(s->part_keys[0]);  // the unsorted keys are no longer needed
        NB_CUDA(cudaMemcpyAsync(held, s->perm, sizeof(int) * n, cudaMemcpyDeviceToDevice, s->stream));
        through = held;
    }
    set_perm_kernel<<<blocks_for(n, kStateThreads), kStateThreads, 0, s->stream>>>(sorted, through, n, s->perm, s->inv_perm);
    NB_CUDA(cudaGetLastError());
    s->kernel_launches += 4;
    s->perm_valid = true;
    s->uploads_since_partition = 0;
    s->partitions++;
    return NB_OK;
}

// Long runs: bodies leave the stretch of the curve their rank was given at upload. At every re-sort of the stream
// the ranks publish what they hold, recompute the curve from the current positions and regroup.
int repartition(nb_ctx* c) {
    NB_TRY(publish_owned(c, kXbArrays, true));
    for (nb_ctx* s : Team(c)) {
        NB_CUDA(cudaSetDevice(s->device));
        NB_TRY(compute_perm(s, true));
        s->order_valid = false;  // the rows changed hands: the stream order is rebuilt by the sort that follows
        iota_order_kernel<<<blocks_for(s->cap, kStateThreads), kStateThreads, 0, s->stream>>>(s->order, 0, s->cap);
        NB_CUDA(cudaGetLastError());
    }
    return NB_OK;
}

}  // namespace

// ---------------------------------------------------------------------------------------
// state upload / download
// ---------------------------------------------------------------------------------------
// Multi-GPU contexts that own stretches of one curve (partition.cuh): every rank copies ITS upload rows [lo,hi)
// from the host as before, the ranks exchange them over NVLink, and each keeps the bodies of its stretch.
static int upload_partitioned(nb_ctx* c, const double* pos, const double* vel, const float* mass,
                              const uint8_t* pinned, int dim, const double* acc) {
    const Team team(c);
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        NB_TRY(ensure_partition_buffers(s));
    }
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        const long long rows = s->hi - s->lo;
        if (rows <= 0) continue;
        const size_t vec_bytes = sizeof(double) * (size_t)dim * (size_t)rows;
        const size_t off = (size_t)s->lo * dim;
        NB_TRY(ensure_staging(s, 3 * vec_bytes));
        double* d_pos = static_cast<double*>(s->staging);
        double* d_vel = d_pos + (size_t)dim * rows;
        double* d_acc = d_vel + (size_t)dim * rows;
        NB_CUDA(cudaMemcpyAsync(d_pos, pos + off, vec_bytes, cudaMemcpyHostToDevice, s->stream));
        NB_CUDA(cudaMemcpyAsync(d_vel, vel + off, vec_bytes, cudaMemcpyHostToDevice, s->stream));
        if (acc) NB_CUDA(cudaMemcpyAsync(d_acc, acc + off, vec_bytes, cudaMemcpyHostToDevice, s->stream));
        NB_CUDA(cudaMemcpyAsync(s->xb_mass + s->lo, mass + s->lo, sizeof(float) * rows, cudaMemcpyHostToDevice, s->stream));
        if (pinned)
            NB_CUDA(cudaMemcpyAsync(s->xb_pinned + s->lo, pinned + s->lo, (size_t)rows, cudaMemcpyHostToDevice, s->stream));
        else
            NB_CUDA(cudaMemsetAsync(s->xb_pinned + s->lo, 0, (size_t)rows, s->stream));
        unpack_aos_xb_kernel<<<blocks_for(rows, kStateThreads), kStateThreads, 0, s->stream>>>(
            d_pos, d_vel, acc ? d_acc : nullptr, dim, s->lo, rows, s->cap, xb_out(s));
        NB_CUDA(cudaGetLastError());
        s->kernel_launches++;
    }
    if (c->group) NB_TRY(team_barrier(c));
    else NB_TRY(allgather_xb(c, 0, 9, true));
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        // a re-upload of a world of the same size keeps the ownership (the caller's per-frame pattern: the bodies
        // have moved a step, the curve order is as good as before); it is recomputed every 64th time
        if (!s->perm_valid || ++s->uploads_since_partition >= 64) NB_TRY(compute_perm(s, false));
        const long long n_local = s->hi - s->lo;
        if (n_local > 0) {
            gather_owned_kernel<<<blocks_for(n_local, kStateThreads), kStateThreads, 0, s->stream>>>(
                s->xb, s->xb_mass, s->xb_pinned, s->cap, s->perm, nullptr, 9, s->view(), s->mass, s->pinned);
            NB_CUDA(cudaGetLastError());
            s->kernel_launches++;
        }
        s->acc_valid = acc != nullptr;
    }
    for (nb_ctx* s : team) {  // the borrowed host arrays may be pageable
        NB_CUDA(cudaSetDevice(s->device));
        NB_CUDA(cudaStreamSynchronize(s->stream));
    }
    return NB_OK;
}

static int upload_dim(nb_ctx* c, int64_t n, const double* pos, const double* vel, const float* mass,
                      const uint8_t* pinned, int dim, const double* acc = nullptr) {
    if (!c) return fail(NB_ERR_INVALID, "nb_upload: null context");
    if (n < 0 || (n > 0 && (!pos || !vel || !mass))) return fail(NB_ERR_INVALID, "nb_upload: null array");
    const Team team(c);
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        NB_CUDA(cudaStreamSynchronize(s->stream));
    }
    // kernel specialisations (NB_FORCE_GENERIC=1 disables them): one common mass over every body that can
    // ever pass the gather filter, and z == 0 for reference-layout (2-D) uploads
    // one pass over the masses: largest finite one (packing exponent), and whether all live ones are equal
    float m0 = 0.f, mx = 0.f;
    bool uniform = n > 0, every = true;
    for (int64_t i = 0; i < n; ++i) {
        const float m = mass[i];
        if (!(m > 0.f && m <= 3.4028234663852886e38f)) {  // never live, Physics.hpp:198-199
            every = false;
            continue;
        }
        if (m0 == 0.f) m0 = m;
        uniform = uniform && (m == m0);
        mx = m > mx ? m : mx;
    }
    int mass_exp = 0;
    if (mx > 0.f) frexpf(mx, &mass_exp);  // mx = f * 2^e, f in [0.5, 1): packed masses land in (0, 1)
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        s->dim = dim;
        s->mass_exp = mass_exp;
        s->uniform_mass = uniform && m0 > 0.f && !s->force_generic;
        s->diag_uniform_mass = s->uniform_mass && every;  // compute_diagnostics sums over ALL bodies
        s->uniform_packed_mass = (double)m0 * pow2d(-mass_exp);
        s->planar = dim == 2 && !s->force_generic;
        NB_TRY(size_for(s, n));
        s->acc_valid = false;
        s->packed_valid = false; s->drop_speculation();
        s->step_diag_valid = false;
    }
    if (n == 0) return NB_OK;
    if (uses_partition(c)) return upload_partitioned(c, pos, vel, mass, pinned, dim, acc);
    // a rank only needs the FP64 state of its own targets: the other bodies reach it as FP32 sources
    // through the exchange of the packed stream (prepare_packed), so only rows [lo,hi) are copied
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        const long long rows = s->hi - s->lo;
        if (rows > 0) {
            const size_t vec_bytes = sizeof(double) * (size_t)dim * (size_t)rows;
            const size_t off = (size_t)s->lo * dim;
            NB_TRY(ensure_staging(s, 3 * vec_bytes));
            double* d_pos = static_cast<double*>(s->staging);
            double* d_vel = d_pos + (size_t)dim * rows;
            double* d_acc = d_vel + (size_t)dim * rows;
            NB_CUDA(cudaMemcpyAsync(d_pos, pos + off, vec_bytes, cudaMemcpyHostToDevice, s->stream));
            NB_CUDA(cudaMemcpyAsync(d_vel, vel + off, vec_bytes, cudaMemcpyHostToDevice, s->stream));
            if (acc) NB_CUDA(cudaMemcpyAsync(d_acc, acc + off, vec_bytes, cudaMemcpyHostToDevice, s->stream));
            NB_CUDA(cudaMemcpyAsync(s->mass + s->lo, mass + s->lo, sizeof(float) * rows, cudaMemcpyHostToDevice, s->stream));
            if (pinned)
                NB_CUDA(cudaMemcpyAsync(s->pinned + s->lo, pinned + s->lo, (size_t)rows, cudaMemcpyHostToDevice, s->stream));
            unpack_aos_kernel<<<blocks_for(rows, 256), 256, 0, s->stream>>>(d_pos, d_vel, acc ? d_acc : nullptr, dim, s->lo,
                                                                           rows, s->view());
            NB_CUDA(cudaGetLastError());
            s->kernel_launches++;
        }
        s->acc_valid = acc != nullptr;
    }
    // The j-stream is packed by the first force pass that needs it (prepare_packed), not here: a frame that
    // starts with a drift (Verlet with the caller's accelerations) would only throw this pack away.
    // the borrowed host arrays may be pageable: do not return while a copy could still read them
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        NB_CUDA(cudaStreamSynchronize(s->stream));
    }
    return NB_OK;
}

extern "C" int nb_upload(nb_ctx* c, int64_t n, const double* pos_xy, const double* vel_xy, const float* mass,
                         const uint8_t* pinned) {
    return upload_dim(c, n, pos_xy, vel_xy, mass, pinned, 2);
}
extern "C" int nb_upload3(nb_ctx* c, int64_t n, const double* pos_xyz, const double* vel_xyz, const float* mass,
                          const uint8_t* pinned) {
    return upload_dim(c, n, pos_xyz, vel_xyz, mass, pinned, 3);
}
extern "C" int nb_upload_state(nb_ctx* c, int64_t n, int dim, const double* pos, const double* vel, const double* acc,
                               const float* mass, const uint8_t* pinned) {
    if (dim != 2 && dim != 3) return fail(NB_ERR_INVALID, "nb_upload_state: dim must be 2 or 3");
    return upload_dim(c, n, pos, vel, mass, pinned, dim, acc);
}

static int download_dim(nb_ctx* c, double* pos, double* vel, double* acc, int dim, double* acc_prev = nullptr) {
    if (!c) return fail(NB_ERR_INVALID, "nb_download: null context");
    const Team team(c);
    double* out[4] = {pos, vel, acc, acc_prev};
    if (uses_partition(c) && c->n > 0 && c->perm_valid) {
        // the rows a rank downloads are upload rows [lo,hi); their bodies live wherever the curve put them:
        // publish every rank's owned rows, then read the wanted ones through inv_perm
        int n_arrays = 0;
        for (int w = 0; w < 4; ++w)
            if (out[w]) n_arrays = 3 * (w + 1);
        if (n_arrays > 0) NB_TRY(publish_owned(c, n_arrays, false));
        for (nb_ctx* s : team) {
            NB_CUDA(cudaSetDevice(s->device));
            const long long n_local = s->hi - s->lo;
            if (n_local <= 0) continue;
            const size_t vec_bytes = sizeof(double) * (size_t)dim * (size_t)n_local;
            NB_TRY(ensure_staging(s, 4 * vec_bytes));
            for (int w = 0; w < 4; ++w) {
                if (!out[w]) continue;
                double* d = static_cast<double*>(s->staging) + (size_t)w * dim * n_local;
                pack_aos_xb_kernel<<<blocks_for(n_local, kStateThreads), kStateThreads, 0, s->stream>>>(
                    s->xb, s->cap, s->inv_perm, w, dim, s->lo, s->hi, d);
                s->kernel_launches++;
                NB_CUDA(cudaMemcpyAsync(out[w] + (size_t)s->lo * dim, d, vec_bytes, cudaMemcpyDeviceToHost, s->stream));
            }
        }
        for (nb_ctx* s : team) {
            NB_CUDA(cudaSetDevice(s->device));
            NB_CUDA(cudaStreamSynchronize(s->stream));
            s->absorb_filter_flip();
        }
        return NB_OK;
    }
    for (nb_ctx* s : team) {  // every member fills its own rows of the caller's arrays
        NB_CUDA(cudaSetDevice(s->device));
        const long long n_local = s->hi - s->lo;
        if (n_local <= 0) continue;
        const size_t vec_bytes = sizeof(double) * (size_t)dim * (size_t)n_local;
        NB_TRY(ensure_staging(s, 4 * vec_bytes));
        const StateView v = s->view();
        for (int w = 0; w < 4; ++w) {
            if (!out[w]) continue;
            double* d = static_cast<double*>(s->staging) + (size_t)w * dim * n_local;
            pack_aos_kernel<<<blocks_for(n_local, 256), 256, 0, s->stream>>>(v, w, dim, d);
            s->kernel_launches++;
            NB_CUDA(cudaMemcpyAsync(out[w] + (size_t)s->lo * dim, d, vec_bytes, cudaMemcpyDeviceToHost, s->stream));
        }
    }
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        NB_CUDA(cudaStreamSynchronize(s->stream));
        s->absorb_filter_flip();
    }
    return NB_OK;
}

extern "C" int nb_download(nb_ctx* c, double* pos_xy, double* vel_xy, double* acc_xy) {
    return download_dim(c, pos_xy, vel_xy, acc_xy, 2);
}
extern "C" int nb_download3(nb_ctx* c, double* pos_xyz, double* vel_xyz, double* acc_xyz) {
    return download_dim(c, pos_xyz, vel_xyz, acc_xyz, 3);
}
extern "C" int nb_download_state(nb_ctx* c, int dim, double* pos, double* vel, double* acc, double* acc_prev) {
    if (dim != 2 && dim != 3) return fail(NB_ERR_INVALID, "nb_download_state: dim must be 2 or 3");
    return download_dim(c, pos, vel, acc, dim, acc_prev);
}

// The stream order is part of the summation order: a checkpoint that wants its resumed run to be
// bit-identical to the uninterrupted one carries it (scenarios.save_snapshot / load_snapshot).
extern "C" int nb_stream_order(nb_ctx* c, int32_t* order, int64_t* generations_since_sort) {
    if (!c || !order) return fail(NB_ERR_INVALID, "nb_stream_order: null");
    if (uses_partition(c) && c->perm_valid) {  // stream row -> internal row -> the caller's (upload) row
        for (nb_ctx* s : Team(c)) {
            NB_CUDA(cudaSetDevice(s->device));
            const long long n_local = s->hi - s->lo;
            if (n_local > 0) {
                map_rows_kernel<<<blocks_for(n_local, kStateThreads), kStateThreads, 0, s->stream>>>(
                    s->order, s->perm, s->lo, s->hi, s->part_vals[0]);
                NB_CUDA(cudaMemcpyAsync(order + s->lo, s->part_vals[0] + s->lo, sizeof(int) * n_local, cudaMemcpyDeviceToHost, s->stream));
            }
            NB_CUDA(cudaStreamSynchronize(s->stream));
        }
        if (generations_since_sort) *generations_since_sort = c->packs_since_sort;
        return NB_OK;
    }
    for (nb_ctx* s : Team(c)) {
        NB_CUDA(cudaSetDevice(s->device));
        const long long n_local = s->hi - s->lo;
        if (n_local > 0)
            NB_CUDA(cudaMemcpyAsync(order + s->lo, s->order + s->lo, sizeof(int) * n_local, cudaMemcpyDeviceToHost, s->stream));
        NB_CUDA(cudaStreamSynchronize(s->stream));
    }
    if (generations_since_sort) *generations_since_sort = c->packs_since_sort;
    return NB_OK;
}

// Multi-GPU contexts with ownership along the curve: the stream order IS the ownership (stream row k of the
// whole world belongs to the rank whose range holds k), so restoring it moves bodies between ranks: publish the
// state every rank holds, make the given order the new permutation, fetch the rows this rank now owns.
static int set_stream_order_partitioned(nb_ctx* c, const int32_t* order, int64_t generations_since_sort) {
    const Team team(c);
    const long long n = c->n;
    for (nb_ctx* s : team) {  // entries a context reads: its own range (one process per GPU) or all (one process)
        const long long q0 = s->lo, q1 = s->hi;
        for (long long q = q0; q < q1; ++q)
            if (order[q] < 0 || order[q] >= n) return fail(NB_ERR_INVALID, "nb_set_stream_order: row out of range");
    }
    if (c->group) {
        std::vector<uint8_t> seen((size_t)n, 0);
        for (long long q = 0; q < n; ++q) {
            if (seen[(size_t)order[q]]) return fail(NB_ERR_INVALID, "nb_set_stream_order: not a permutation");
            seen[(size_t)order[q]] = 1;
        }
    }
    NB_TRY(publish_owned(c, kXbArrays, true));  // XB: every body's state, previous internal order
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        const long long n_local = s->hi - s->lo;
        int* fresh = s->part_vals[0];
        if (c->group) {
            NB_CUDA(cudaMemcpyAsync(fresh, order, sizeof(int) * n, cudaMemcpyHostToDevice, s->stream));
        } else {
            const size_t rows = (size_t)s->tiles_per_rank * kTileBodies;
            if (n_local > 0)
                NB_CUDA(cudaMemcpyAsync(fresh + s->lo, order + s->lo, sizeof(int) * n_local, cudaMemcpyHostToDevice, s->stream));
            NB_NCCL(g_nccl.AllGather(fresh + (size_t)s->rank * rows, fresh, rows, kNcclInt32, s->nccl_comm, s->stream));
        }
        if (n_local > 0) {
            gather_owned_kernel<<<blocks_for(n_local, kStateThreads), kStateThreads, 0, s->stream>>>(
                s->xb, s->xb_mass, s->xb_pinned, s->cap, fresh, s->inv_perm, kXbArrays, s->view(), s->mass, s->pinned);
            iota_order_kernel<<<blocks_for(n_local, kStateThreads), kStateThreads, 0, s->stream>>>(s->order, s->lo, s->hi);
        }
        set_perm_kernel<<<blocks_for(n, kStateThreads), kStateThreads, 0, s->stream>>>(fresh, nullptr, n, s->perm, s->inv_perm);
        NB_CUDA(cudaGetLastError());
        s->kernel_launches += 3;
    }
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        NB_CUDA(cudaStreamSynchronize(s->stream));
        s->order_valid = true;
        s->packs_since_sort = generations_since_sort;
        s->packed_valid = false; s->drop_speculation();
        s->step_diag_valid = false;
    }
    return NB_OK;
}

extern "C" int nb_partition_info(nb_ctx* c, int64_t* values) {
    if (!c || !values) return fail(NB_ERR_INVALID, "nb_partition_info: null");
    values[0] = uses_partition(c) ? 1 : 0;
    values[1] = c->partitions;
    values[2] = c->uploads_since_partition;
    values[3] = c->sorts;
    return NB_OK;
}

extern "C" int nb_set_stream_order(nb_ctx* c, const int32_t* order, int64_t generations_since_sort) {
    if (!c || !order) return fail(NB_ERR_INVALID, "nb_set_stream_order: null");
    if (generations_since_sort < 0) return fail(NB_ERR_INVALID, "nb_set_stream_order: negative age");
    if (uses_partition(c) && c->perm_valid) return set_stream_order_partitioned(c, order, generations_since_sort);
    for (nb_ctx* s : Team(c)) {
        const long long n_local = s->hi - s->lo;
        // the rows a context owns must be a permutation of themselves
        std::vector<uint8_t> seen((size_t)std::max<long long>(n_local, 0), 0);
        for (long long q = 0; q < n_local; ++q) {
            const long long b = (long long)order[s->lo + q] - s->lo;
            if (b < 0 || b >= n_local || seen[(size_t)b])
                return fail(NB_ERR_INVALID, "nb_set_stream_order: not a permutation of the context's rows");
            seen[(size_t)b] = 1;
        }
    }
    for (nb_ctx* s : Team(c)) {
        NB_CUDA(cudaSetDevice(s->device));
        const long long n_local = s->hi - s->lo;
        NB_CUDA(cudaStreamSynchronize(s->stream));
        if (n_local > 0)
            NB_CUDA(cudaMemcpyAsync(s->order + s->lo, order + s->lo, sizeof(int) * n_local, cudaMemcpyHostToDevice, s->stream));
        NB_CUDA(cudaStreamSynchronize(s->stream));
        s->order_valid = true;
        s->packs_since_sort = generations_since_sort;
        s->packed_valid = false; s->drop_speculation();
        s->step_diag_valid = false;
    }
    return NB_OK;
}

// ---------------------------------------------------------------------------------------
// FP64 exact path (exact.cuh): unfused, every kernel reads state written by the previous one
// ---------------------------------------------------------------------------------------
static bool use_exact_precision(const nb_ctx* c) {
    if (c->world > 1) return false;
    if (c->precision == NB_PRECISION_FP64_EXACT) return true;
    return c->precision == NB_PRECISION_AUTO && c->n <= NB_EXACT_AUTO_MAX;
}

// The reference's switch (Physics.hpp:212) is honoured and this world is on the tree side of it.
static bool tree_active(const nb_ctx* c) {
    return c->solver == NB_SOLVER_REFERENCE && c->world == 1 && c->dim == 2 && c->n > (long long)c->bh_threshold;
}

// The tree is the reference's 2-D code: a 3-D upload on its side of the switch has no defined result.
static int check_solver(const nb_ctx* c) {
    if (c->solver == NB_SOLVER_REFERENCE && c->dim != 2 && c->n > (long long)c->bh_threshold)
        return fail(NB_ERR_INVALID, "NB_SOLVER_REFERENCE: the reference's Barnes-Hut branch is 2-D; upload with nb_upload "
                                    "or select NB_SOLVER_DIRECT");
    return NB_OK;
}

// Force evaluation and integrator moves in the reference's own FP64 sequence (the tree branch always is).
bool use_exact(const nb_ctx* c) { return tree_active(c) || use_exact_precision(c); }

extern "C" int nb_set_solver(nb_ctx* c, int solver, int32_t bh_threshold, float bh_theta) {
    if (!c) return fail(NB_ERR_INVALID, "nb_set_solver: null context");
    if (solver != NB_SOLVER_DIRECT && solver != NB_SOLVER_REFERENCE)
        return fail(NB_ERR_INVALID, "nb_set_solver: unknown solver");
    if (solver == NB_SOLVER_REFERENCE && c->world > 1)
        return fail(NB_ERR_INVALID, "nb_set_solver: the reference's Barnes-Hut branch needs a single-GPU context");
    if (solver != c->solver || bh_threshold != c->bh_threshold || bh_theta != c->bh_theta) c->acc_valid = false;
    c->solver = solver;
    c->bh_threshold = bh_threshold;
    c->bh_theta = bh_theta;
    return NB_OK;
}

extern "C" int nb_get_solver(nb_ctx* c, int* solver, int32_t* bh_threshold, float* bh_theta) {
    if (!c) return fail(NB_ERR_INVALID, "nb_get_solver: null context");
    if (solver) *solver = c->solver;
    if (bh_threshold) *bh_threshold = c->bh_threshold;
    if (bh_theta) *bh_theta = c->bh_theta;
    return NB_OK;
}

extern "C" int nb_solver_profile(nb_ctx* c, int enable) {
    if (!c) return fail(NB_ERR_INVALID, "nb_solver_profile: null context");
    c->bh_stats = enable != 0;
    return NB_OK;
}

extern "C" int nb_set_precision(nb_ctx* c, int precision) {
    if (!c) return fail(NB_ERR_INVALID, "nb_set_precision: null context");
    if (precision < NB_PRECISION_AUTO || precision > NB_PRECISION_FP64_EXACT)
        return fail(NB_ERR_INVALID, "nb_set_precision: unknown mode");
    if (precision == NB_PRECISION_FP64_EXACT && c->world > 1)
        return fail(NB_ERR_INVALID, "nb_set_precision: FP64_EXACT needs a single-GPU context");
    if (precision != c->precision) c->acc_valid = false;
    c->precision = precision;
    return NB_OK;
}

#include "host_tree.cuh"  // the reference's Barnes-Hut branch: tree_force(), nb_solver_info

static int exact_force(nb_ctx* c) {
    if (tree_active(c)) return tree_force(c);
    return exact_force_direct(c);
}

static int exact_force_direct(nb_ctx* c) {
    exact_force_kernel<<<blocks_for(c->n, kExactThreads), kExactThreads, 0, c->stream>>>(c->view(), c->cfg.g,
                                                                                        c->cfg.softening);
    NB_CUDA(cudaGetLastError());
    c->kernel_launches++;
    c->force_evals++;
    c->interactions += (double)c->n * (double)c->n;
    return NB_OK;
}

static int exact_verlet(nb_ctx* c, double h, long long n_sub) {
    const unsigned int grid = blocks_for(c->n, kStateThreads);
    for (long long k = 0; k < n_sub; ++k) {
        verlet_drift_kernel<<<grid, kStateThreads, 0, c->stream>>>(c->view(), h, 0.5 * h * h);
        NB_TRY(exact_force(c));
        verlet_kick_kernel<<<grid, kStateThreads, 0, c->stream>>>(c->view(), h, c->cfg.max_speed);
        c->kernel_launches += 2;
    }
    NB_CUDA(cudaGetLastError());
    c->packed_valid = false; c->drop_speculation();
    return NB_OK;
}

static int exact_euler_move(nb_ctx* c, double h) {
    euler_move_kernel<<<blocks_for(c->n, kStateThreads), kStateThreads, 0, c->stream>>>(c->view(), h, c->cfg.max_speed);
    NB_CUDA(cudaGetLastError());
    c->kernel_launches++;
    c->packed_valid = false; c->drop_speculation();
    return NB_OK;
}

// ---------------------------------------------------------------------------------------
// the hot path: compute_gravity / integrate / fused frames
// ---------------------------------------------------------------------------------------
static void set_acc_valid(nb_ctx* c, bool v) {
    for (nb_ctx* s : Team(c)) s->acc_valid = v;
}

extern "C" int nb_compute_gravity(nb_ctx* c) {
    if (!c) return fail(NB_ERR_INVALID, "nb_compute_gravity: null context");
    if (c->n == 0) return NB_OK;  // Physics.hpp:208
    NB_CUDA(cudaSetDevice(c->device));
    NB_TRY(check_solver(c));
    if (use_exact(c)) NB_TRY(exact_force(c));
    else NB_TRY(force_pass(c, kForceOnly, false, 0.0));
    set_acc_valid(c, true);
    return NB_OK;
}

// n_sub velocity Verlet substeps (Physics.hpp:296-325) starting from Acceleration == a(x).
// Steady state is ONE launch per substep: the previous kick epilogue already drifted this substep into
// the shadow positions and packed its j-stream (and, sharded, its all-gather is already in flight).
static int verlet_substeps(nb_ctx* c, double h, long long n_sub, bool pot) {
    for (long long k = 0; k < n_sub; ++k) {
        if (c->spec_valid && c->spec_h == h && c->packed_valid) {
            commit_drift(c);
        } else {
            // first substep after an upload / edit / change of dt: stand-alone drift, new generation
            for (nb_ctx* s : Team(c)) {
                NB_CUDA(cudaSetDevice(s->device));
                const long long n_local = s->hi - s->lo;
                if (n_local > 0) {
                    verlet_drift_kernel<<<blocks_for(n_local, kStateThreads), kStateThreads, 0, s->stream>>>(
                        s->view(), h, 0.5 * h * h);
                    s->kernel_launches++;
                }
                s->packed_valid = false; s->drop_speculation();
                // a generation like the ones commit_drift() counts: the stream is re-sorted at the same frame
                // whether this drift was prepared by the last kick or not (a resumed run equals the uninterrupted one)
                if (s->order_valid) s->packs_since_sort++;
            }
        }
        NB_TRY(force_pass(c, kVerletKickDrift, pot && (k + 1 == n_sub), h));
    }
    return NB_OK;
}

extern "C" int nb_integrate(nb_ctx* c, float dt) {
    if (!c) return fail(NB_ERR_INVALID, "nb_integrate: null context");
    if (c->n == 0) return NB_OK;
    NB_CUDA(cudaSetDevice(c->device));
    NB_TRY(check_solver(c));
    float dt_sub;
    const int n_steps = nb_substep_plan(dt, c->cfg.max_substep, c->cfg.max_substeps_per_frame, &dt_sub);
    const double h = (double)dt_sub;
    if (use_exact(c)) {
        if (c->cfg.integrator == NB_INTEGRATOR_EULER) {
            for (int k = 0; k < n_steps; ++k) {
                NB_TRY(exact_euler_move(c, h));
                if (k + 1 < n_steps) NB_TRY(exact_force(c));  // Physics.hpp:292
            }
            set_acc_valid(c, false);
        } else {
            NB_TRY(exact_verlet(c, h, n_steps));
            set_acc_valid(c, true);
        }
        return NB_OK;
    }
    if (c->cfg.integrator == NB_INTEGRATOR_EULER) {
        // kick/drift with the stored acceleration, then (gravity, kick/drift) per further substep
        for (nb_ctx* s : Team(c)) {
            NB_CUDA(cudaSetDevice(s->device));
            const long long n_local = s->hi - s->lo;
            if (n_local > 0) {
                euler_move_kernel<<<blocks_for(n_local, kStateThreads), kStateThreads, 0, s->stream>>>(s->view(), h,
                                                                                                     s->cfg.max_speed);
                s->kernel_launches++;
            }
            s->packed_valid = false; s->drop_speculation();
        }
        set_acc_valid(c, false);
        for (int k = 1; k < n_steps; ++k) NB_TRY(force_pass(c, kEuler, false, h));
    } else {
        NB_TRY(verlet_substeps(c, h, n_steps, false));
        set_acc_valid(c, true);  // Physics.hpp:324
    }
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

static int step_impl(nb_ctx* c, float dt_eff, int64_t n_frames, bool pot) {
    if (!c) return fail(NB_ERR_INVALID, "nb_step: null context");
    if (n_frames < 0) return fail(NB_ERR_INVALID, "nb_step: negative frame count");
    if (c->n == 0 || n_frames == 0) return NB_OK;
    NB_CUDA(cudaSetDevice(c->device));
    NB_TRY(check_solver(c));
    float dt_sub;
    const int n_steps = nb_substep_plan(dt_eff, c->cfg.max_substep, c->cfg.max_substeps_per_frame, &dt_sub);
    const double h = (double)dt_sub;
    const long long total = n_frames * (long long)n_steps;
    c->step_diag_exact = false;
    if (use_exact(c) && c->n <= kSmallWorldMax && !tree_active(c)) {
        // the whole call in one launch (exact.cuh, small_world_steps_kernel)
        const bool verlet = c->cfg.integrator != NB_INTEGRATOR_EULER;
        const int first_force = verlet && !c->acc_valid ? 1 : 0;
        const int threads = (int)std::max<long long>(32, ((c->n + 31) / 32) * 32);
        small_world_steps_kernel<<<1, threads, 0, c->stream>>>(c->view(), c->cfg.g, c->cfg.softening, h, 0.5 * h * h,
                                                              c->cfg.max_speed, verlet ? 1 : 0, first_force, total);
        NB_CUDA(cudaGetLastError());
        c->kernel_launches++;
        c->force_evals += total + first_force;
        c->interactions += (double)(total + first_force) * (double)c->n * (double)c->n;
        set_acc_valid(c, verlet);  // Euler: Acceleration holds a(x) of the positions before the last move
        c->packed_valid = false; c->drop_speculation();
        c->step_diag_exact = pot;
        return NB_OK;
    }
    if (use_exact(c)) {
        if (c->cfg.integrator == NB_INTEGRATOR_EULER) {
            for (long long k = 0; k < total; ++k) {
                NB_TRY(exact_force(c));
                NB_TRY(exact_euler_move(c, h));
            }
            set_acc_valid(c, false);
        } else {
            if (!c->acc_valid) NB_TRY(exact_force(c));
            set_acc_valid(c, true);
            NB_TRY(exact_verlet(c, h, total));
        }
        c->step_diag_exact = pot;
        return NB_OK;
    }
    if (c->cfg.integrator == NB_INTEGRATOR_EULER) {
        // frame = gravity system + n_steps x (kick/drift [+ gravity]): n_steps fused passes
        for (long long k = 0; k < total; ++k) NB_TRY(force_pass(c, kEuler, pot && (k + 1 == total), h));
        set_acc_valid(c, false);  // Acceleration holds a(x) of the positions before the last move
    } else {
        // the gravity system's evaluation is only needed when Acceleration != a(x)
        if (!c->acc_valid) NB_TRY(force_pass(c, kForceOnly, false, 0.0));
        set_acc_valid(c, true);
        NB_TRY(verlet_substeps(c, h, total, pot));
    }
    return NB_OK;
}

extern "C" int nb_step(nb_ctx* c, float dt_eff, int64_t n_frames) { return step_impl(c, dt_eff, n_frames, false); }
extern "C" int nb_step_with_diagnostics(nb_ctx* c, float dt_eff, int64_t n_frames) {
    return step_impl(c, dt_eff, n_frames, true);
}

// ---------------------------------------------------------------------------------------
// diagnostics, zero_net_momentum, reset_scenario
// ---------------------------------------------------------------------------------------
namespace {

// Sum `rows` per-CTA partial rows in fixed order on the device, combine the ranks' totals in
// rank order on the host, and fill the reference's Diagnostics (Physics.hpp:163-176).
int finish_diag(nb_ctx* c, nb_diag* out) {
    memset(out, 0, sizeof(*out));
    const Team team(c);
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        const long long rows = (s->hi - s->lo) > 0 ? s->step_diag_rows : 0;
        for (int k = 0; k < kDiagValues; ++k) s->h_totals[k] = 0.0;
        if (rows > 0) {
            sum_rows_kernel<<<1, 32, 0, s->stream>>>(s->diag_part, (int)rows, 10, kDiagValues, s->totals);
            s->kernel_launches++;
            NB_CUDA(cudaGetLastError());
        } else {
            NB_CUDA(cudaMemsetAsync(s->totals, 0, sizeof(double) * kDiagValues, s->stream));
        }
    }
    double t[kDiagValues];
    for (int k = 0; k < kDiagValues; ++k) t[k] = 0.0;
    if (c->world > 1 && !c->group) {
        // every rank receives every rank's totals and adds them in rank order: bit-identical everywhere
        double* h = c->h_totals;
        double* gathered = c->totals + 16;
        NB_NCCL(g_nccl.AllGather(c->totals, gathered, kDiagValues, kNcclFloat64, c->nccl_comm, c->stream));
        NB_CUDA(cudaMemcpyAsync(h, gathered, sizeof(double) * kDiagValues * c->world, cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaStreamSynchronize(c->stream));
        for (int k = 0; k < kDiagValues; ++k)
            for (int r = 0; r < c->world; ++r) t[k] += h[r * kDiagValues + k];
    } else {
        // one process holds every member: read each member's totals and add them in member order
        for (nb_ctx* s : team) {
            NB_CUDA(cudaSetDevice(s->device));
            NB_CUDA(cudaMemcpyAsync(s->h_totals, s->totals, sizeof(double) * kDiagValues, cudaMemcpyDeviceToHost, s->stream));
        }
        for (nb_ctx* s : team) {
            NB_CUDA(cudaSetDevice(s->device));
            NB_CUDA(cudaStreamSynchronize(s->stream));
            s->absorb_filter_flip();
            for (int k = 0; k < kDiagValues; ++k) t[k] += s->h_totals[k];
        }
    }
    const double M = t[8];
    bool ok = t[9] == 0.0;
    const double vals[10] = {t[0], t[1], t[0] + t[1], t[2], t[3], t[4],
                             M > 0.0 ? t[5] / M : 0.0, M > 0.0 ? t[6] / M : 0.0, M > 0.0 ? t[7] / M : 0.0, M};
    for (int k = 0; k < 10; ++k) ok = ok && std::isfinite(vals[k]);
    if (!ok) {  // the reference returns a zeroed struct with ok = false (Physics.hpp:123,140-145)
        out->ok = 0;
        return NB_OK;
    }
    out->kinetic = vals[0];
    out->potential = vals[1];
    out->energy = vals[2];
    for (int k = 0; k < 3; ++k) {
        out->momentum[k] = vals[3 + k];
        out->com[k] = vals[6 + k];
    }
    out->total_mass = M;
    out->ok = 1;
    return NB_OK;
}

}  // namespace

extern "C" int nb_compute_diagnostics(nb_ctx* c, nb_diag* out) {
    if (!c || !out) return fail(NB_ERR_INVALID, "nb_compute_diagnostics: null");
    NB_CUDA(cudaSetDevice(c->device));
    if (c->n == 0) {  // Physics.hpp:124-127
        memset(out, 0, sizeof(*out));
        out->ok = 1;
        return NB_OK;
    }
    if (use_exact_precision(c)) {
        const long long rows = blocks_for(c->n, kDiagThreads);
        diag_exact_kernel<<<(unsigned int)rows, kDiagThreads, 0, c->stream>>>(c->view(), c->cfg.g, c->cfg.softening,
                                                                               c->diag_part);
        NB_CUDA(cudaGetLastError());
        c->kernel_launches++;
        c->step_diag_valid = false;
        c->step_diag_rows = rows;
        return finish_diag(c, out);
    }
    // hot loop B (Physics.hpp:148-161) through the potential-only instantiation of the force kernel, over a
    // stream packed with this function's own rules (every body with a finite position, raw masses)
    const Team team(c);
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        NB_CUDA(cudaMemsetAsync(s->extent + kDiagExtentSlot, 0, sizeof(unsigned int), s->stream));
    }
    NB_TRY(team_barrier(c));
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        const long long n_local = s->hi - s->lo;
        if (n_local > 0) {
            const StateView v = s->view();
            const unsigned int grid = (unsigned int)std::min<long long>(blocks_for(n_local, kDiagThreads), 4 * s->sm_count);
            diag_extent_kernel<<<grid, kDiagThreads, 0, s->stream>>>(v.x[0], v.x[1], v.x[2], s->lo, s->hi, diag_out(s));
            s->kernel_launches++;
        }
    }
    if (c->world > 1 && !c->group) {
        unsigned int* slot = c->extent + kDiagExtentSlot;
        NB_NCCL(g_nccl.AllReduce(slot, slot, 1, kNcclUint32, kNcclMax, c->nccl_comm, c->stream));
    }
    NB_TRY(team_barrier(c));
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        if (!s->order_valid) {  // diagnostics right after an upload: order the stream now (the extent is known)
            if (s->n > kGroupBodies) NB_TRY(sort_stream(s, kDiagExtentSlot));
            s->order_valid = true;
            s->packs_since_sort = 0;
        }
        if (s->hi - s->lo > 0) {
            const long long groups = (s->tiles_per_rank * kTileBodies) / kGroupBodies;
            pack_kernel<true><<<(unsigned int)groups, kGroupBodies, 0, s->stream>>>(
                s->view(), s->order, diag_out(s), tx_out(s->dtx, s->cap), nullptr, s->extent, kDiagExtentSlot, s->mass_exp, 1.0,
                0.f);
            s->kernel_launches++;
        }
    }
    if (c->world > 1 && !c->group) {
        const size_t rows_r = (size_t)c->tiles_per_rank * kTileBodies;
        const size_t count = rows_r * 4;
        const size_t hdr_count = rows_r / kGroupBodies * (sizeof(GroupHdr) / sizeof(float));
        float* hdb = reinterpret_cast<float*>(c->dhdr);
        NB_NCCL(g_nccl.GroupStart());
        NB_NCCL(g_nccl.AllGather(c->dpk + (size_t)c->rank * count, c->dpk, count, kNcclFloat32, c->nccl_comm, c->stream));
        NB_NCCL(g_nccl.AllGather(c->dpk_lo + (size_t)c->rank * count, c->dpk_lo, count, kNcclFloat32, c->nccl_comm,
                                 c->stream));
        NB_NCCL(g_nccl.AllGather(hdb + (size_t)c->rank * hdr_count, hdb, hdr_count, kNcclFloat32, c->nccl_comm, c->stream));
        NB_NCCL(g_nccl.GroupEnd());
    }
    NB_TRY(team_barrier(c));  // every member's rows are in every member's diagnostics stream
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        NB_TRY(launch_force(s, 0, kDiagOnly, 2, 0.0));
        NB_CUDA(cudaGetLastError());
        s->step_diag_valid = false;  // diag_part was overwritten
    }
    NB_TRY(team_barrier(c));  // nobody re-packs a diagnostics stream a slower member is still reading
    return finish_diag(c, out);
}

extern "C" int nb_step_diagnostics(nb_ctx* c, nb_diag* out) {
    if (!c || !out) return fail(NB_ERR_INVALID, "nb_step_diagnostics: null");
    NB_CUDA(cudaSetDevice(c->device));
    if (c->n == 0) {
        memset(out, 0, sizeof(*out));
        out->ok = 1;
        return NB_OK;
    }
    if (c->step_diag_exact) return nb_compute_diagnostics(c, out);
    for (nb_ctx* s : Team(c))
        if (!s->step_diag_valid && (s->hi - s->lo) > 0)
            return fail(NB_ERR_INVALID, "nb_step_diagnostics: no nb_step_with_diagnostics since the last upload");
    return finish_diag(c, out);
}

extern "C" int nb_zero_net_momentum(nb_ctx* c) {
    if (!c) return fail(NB_ERR_INVALID, "nb_zero_net_momentum: null context");
    if (c->n == 0) return NB_OK;
    const Team team(c);
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        const long long n_local = s->hi - s->lo;
        const int grid = (int)std::min<long long>(blocks_for(n_local, kStateThreads),
                                                  std::min<long long>(s->diag_part_rows, 2 * s->sm_count));
        if (n_local > 0) {
            momentum_partial_kernel<<<grid, kStateThreads, 0, s->stream>>>(s->view(), s->diag_part);
            sum_rows_kernel<<<1, 32, 0, s->stream>>>(s->diag_part, grid, 4, 4, s->totals);
            s->kernel_launches += 2;
        } else {
            NB_CUDA(cudaMemsetAsync(s->totals, 0, sizeof(double) * 4, s->stream));
        }
    }
    if (c->world > 1) {
        // totals are 4 doubles: bring every rank's to the host and add them in rank order
        double t[4] = {0, 0, 0, 0};
        if (!c->group) {
            double* gathered = c->totals + 16;
            NB_NCCL(g_nccl.AllGather(c->totals, gathered, 4, kNcclFloat64, c->nccl_comm, c->stream));
            NB_CUDA(cudaMemcpyAsync(c->h_totals, gathered, sizeof(double) * 4 * c->world, cudaMemcpyDeviceToHost, c->stream));
            NB_CUDA(cudaStreamSynchronize(c->stream));
            for (int r = 0; r < c->world; ++r)
                for (int k = 0; k < 4; ++k) t[k] += c->h_totals[r * 4 + k];
        } else {
            for (nb_ctx* s : team) {
                NB_CUDA(cudaSetDevice(s->device));
                NB_CUDA(cudaMemcpyAsync(s->h_totals, s->totals, sizeof(double) * 4, cudaMemcpyDeviceToHost, s->stream));
            }
            for (nb_ctx* s : team) {
                NB_CUDA(cudaSetDevice(s->device));
                NB_CUDA(cudaStreamSynchronize(s->stream));
                for (int k = 0; k < 4; ++k) t[k] += s->h_totals[k];
            }
        }
        for (nb_ctx* s : team) {
            NB_CUDA(cudaSetDevice(s->device));
            for (int k = 0; k < 4; ++k) s->h_totals[32 + k] = t[k];  // pinned: stays valid for the async copy
            NB_CUDA(cudaMemcpyAsync(s->totals, s->h_totals + 32, sizeof(double) * 4, cudaMemcpyHostToDevice, s->stream));
            NB_CUDA(cudaStreamSynchronize(s->stream));
        }
    }
    for (nb_ctx* s : team) {
        NB_CUDA(cudaSetDevice(s->device));
        const long long n_local = s->hi - s->lo;
        if (n_local > 0) {
            momentum_apply_kernel<<<blocks_for(n_local, kStateThreads), kStateThreads, 0, s->stream>>>(s->view(), s->totals);
            s->kernel_launches++;
        }
        NB_CUDA(cudaGetLastError());
        s->step_diag_valid = false;
        // velocities changed: a prepared drift is void, and the gather filter (finite velocities, Physics.hpp:198-199)
        // may now pass bodies it dropped (a pinned body's NaN velocity was just reset) -> re-pack and re-evaluate
        // like the reference's next gravity system does
        s->drop_speculation();
        s->packed_valid = false;
        s->acc_valid = false;
    }
    return NB_OK;
}

#include "host_collide.cuh"  // nb_count_overlaps, nb_resolve_collisions
#include "host_scenario.cuh"  // nb_scenario_write / nb_scenario_read
#include "host_noalloc.cuh"   // nb_diagnostics_noalloc (static device scratch, no allocation)

extern "C" int nb_reset_scenario(nb_ctx* c) {
    if (!c) return fail(NB_ERR_INVALID, "nb_reset_scenario: null context");
    // Physics.hpp:87-115 with the constants of Constants.hpp:47-51; values pass through float
    // (raylib::Vector2) before widening (Math.hpp:12)
    const double central = 5.972e24, small = 7.342e22, radius = 3.844e8;
    const float v = static_cast<float>(std::sqrt(c->cfg.g * central / radius));
    const double pos[6] = {0.0, 0.0, (double)static_cast<float>(radius), 0.0, (double)static_cast<float>(-radius), 0.0};
    const double vel[6] = {0.0, 0.0, 0.0, (double)v, 0.0, (double)-v};
    const float mass[3] = {static_cast<float>(central), static_cast<float>(small), static_cast<float>(small)};
    const uint8_t pinned[3] = {0, 0, 0};
    NB_TRY(nb_upload(c, 3, pos, vel, mass, pinned));
    return nb_zero_net_momentum(c);
}

// ---------------------------------------------------------------------------------------
// synchronisation and measurement
// ---------------------------------------------------------------------------------------
extern "C" int nb_synchronize(nb_ctx* c) {
    if (!c) return fail(NB_ERR_INVALID, "nb_synchronize: null context");
    for (nb_ctx* s : Team(c)) {
        NB_CUDA(cudaSetDevice(s->device));
        NB_CUDA(cudaStreamSynchronize(s->stream));
        NB_CUDA(cudaStreamSynchronize(s->comm_stream));
        s->absorb_filter_flip();
    }
    return NB_OK;
}

extern "C" int nb_timer_begin(nb_ctx* c) {
    if (!c) return fail(NB_ERR_INVALID, "nb_timer_begin: null context");
    for (nb_ctx* s : Team(c)) {
        NB_CUDA(cudaSetDevice(s->device));
        NB_CUDA(cudaEventRecord(s->ev_timer0, s->stream));
    }
    return NB_OK;
}

extern "C" int nb_timer_end(nb_ctx* c, float* elapsed_ms) {
    if (!c || !elapsed_ms) return fail(NB_ERR_INVALID, "nb_timer_end: null");
    // the exchange a step leaves in flight belongs to the step: an NCCL-sharded context waits for its
    // all-gather, the members of a multi-GPU context for each other's packing launches
    if (c->world > 1)
        for (nb_ctx* s : Team(c)) {
            NB_CUDA(cudaSetDevice(s->device));
            NB_TRY(consume_published(s));
        }
    for (nb_ctx* s : Team(c)) {
        NB_CUDA(cudaSetDevice(s->device));
        NB_CUDA(cudaEventRecord(s->ev_timer1, s->stream));
    }
    float worst = 0.f;
    for (nb_ctx* s : Team(c)) {  // device time, maximum over the members
        NB_CUDA(cudaSetDevice(s->device));
        NB_CUDA(cudaEventSynchronize(s->ev_timer1));
        NB_CUDA(cudaStreamSynchronize(s->comm_stream));
        float ms = 0.f;
        NB_CUDA(cudaEventElapsedTime(&ms, s->ev_timer0, s->ev_timer1));
        worst = std::max(worst, ms);
    }
    *elapsed_ms = worst;
    return NB_OK;
}

extern "C" int nb_device_info(nb_ctx* c, int* sm_count, int* cc_major, int* cc_minor, int* sm_clock_khz) {
    if (!c) return fail(NB_ERR_INVALID, "nb_device_info: null context");
    int v = 0;
    if (sm_count) *sm_count = c->sm_count;
    if (cc_major) { NB_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMajor, c->device)); *cc_major = v; }
    if (cc_minor) { NB_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMinor, c->device)); *cc_minor = v; }
    if (sm_clock_khz) { NB_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrClockRate, c->device)); *sm_clock_khz = v; }
    return NB_OK;
}

extern "C" int nb_flush_l2(nb_ctx* c) {
    if (!c) return fail(NB_ERR_INVALID, "nb_flush_l2: null context");
    const size_t bytes = (size_t)256 << 20;
    for (nb_ctx* s : Team(c)) {
        NB_CUDA(cudaSetDevice(s->device));
        if (!s->l2_scratch) NB_CUDA(cudaMalloc(&s->l2_scratch, bytes));
        NB_CUDA(cudaMemsetAsync(s->l2_scratch, 0x5a, bytes, s->stream));
    }
    return NB_OK;
}

extern "C" int nb_alloc_pinned(size_t bytes, void** out) {
    if (!out) return fail(NB_ERR_INVALID, "nb_alloc_pinned: null");
    NB_CUDA(cudaMallocHost(out, bytes ? bytes : 1));
    return NB_OK;
}

extern "C" int nb_free_pinned(void* p) {
    if (p) NB_CUDA(cudaFreeHost(p));
    return NB_OK;
}

extern "C" int nb_counters(nb_ctx* c, int64_t* kernel_launches, int64_t* force_evals, double* interactions) {
    if (!c) return fail(NB_ERR_INVALID, "nb_counters: null context");
    long long k = 0;
    double x = 0.0;
    for (nb_ctx* s : Team(c)) {
        k += s->kernel_launches;
        x += s->interactions;
    }
    if (kernel_launches) *kernel_launches = k;
    if (force_evals) *force_evals = c->force_evals;
    if (interactions) *interactions = x;
    return NB_OK;
}

extern "C" int nb_near_stats(nb_ctx* c, int enable, int64_t* visits, int64_t* near_visits) {
    if (!c) return fail(NB_ERR_INVALID, "nb_near_stats: null context");
    NB_CUDA(cudaSetDevice(c->device));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    if (c->near_stats) {
        unsigned long long h[2] = {0, 0};
        NB_CUDA(cudaMemcpy(h, c->near_stats, sizeof(h), cudaMemcpyDeviceToHost));
        if (visits) *visits = (int64_t)c->near_visits;
        if (near_visits) *near_visits = (int64_t)h[1];
    } else {
        if (visits) *visits = 0;
        if (near_visits) *near_visits = 0;
    }
    if (enable && !c->near_stats) NB_CUDA(cudaMalloc(&c->near_stats, 2 * sizeof(unsigned long long)));
    if (c->near_stats) NB_CUDA(cudaMemset(c->near_stats, 0, 2 * sizeof(unsigned long long)));
    c->near_visits = 0;
    if (!enable && c->near_stats) {
        cudaFree(c->near_stats);
        c->near_stats = nullptr;
    }
    return NB_OK;
}

extern "C" int nb_overlap_profile(nb_ctx* c, int enable, float* ms4) {
    if (!c) return fail(NB_ERR_INVALID, "nb_overlap_profile: null context");
    NB_CUDA(cudaSetDevice(c->device));
    if (ms4) {
        ms4[0] = ms4[1] = ms4[2] = ms4[3] = 0.f;
        if (c->profile_comm && c->world > 1) {
            NB_CUDA(cudaStreamSynchronize(c->stream));
            NB_CUDA(cudaStreamSynchronize(c->comm_stream));
            // the last Verlet step read generation `gen`: its own-sources launch (ev_own0/1) ran while the
            // exchange of that generation (issued by the step before) was in flight on the communication stream
            const int g = c->gen & 1;
            NB_CUDA(cudaEventElapsedTime(&ms4[0], c->ev_own0, c->ev_own1));      // own-sources launch
            NB_CUDA(cudaEventElapsedTime(&ms4[1], c->ev_own0, c->ev_comm0[g]));  // exchange start - launch start
            NB_CUDA(cudaEventElapsedTime(&ms4[2], c->ev_own0, c->ev_comm1[g]));  // exchange end   - launch start
            NB_CUDA(cudaEventElapsedTime(&ms4[3], c->ev_comm0[g], c->ev_comm1[g]));
        }
    }
    c->profile_comm = enable != 0;
    return NB_OK;
}

extern "C" int nb_profile_forces(nb_ctx* c, int enable) {
    if (!c) return fail(NB_ERR_INVALID, "nb_profile_forces: null context");
    for (nb_ctx* s : Team(c)) {
        s->profile = enable != 0;
        s->prof_used = 0;
    }
    return NB_OK;
}

extern "C" int nb_profile_read(nb_ctx* c, int64_t* n_launches, double* total_ms) {
    if (!c) return fail(NB_ERR_INVALID, "nb_profile_read: null context");
    NB_CUDA(cudaSetDevice(c->device));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    double total = 0.0;
    for (size_t k = 0; k + 1 < c->prof_used; k += 2) {
        float ms = 0.f;
        NB_CUDA(cudaEventElapsedTime(&ms, c->prof_events[k], c->prof_events[k + 1]));
        total += ms;
    }
    if (n_launches) *n_launches = (int64_t)(c->prof_used / 2);
    if (total_ms) *total_ms = total;
    c->prof_used = 0;
    return NB_OK;
}
