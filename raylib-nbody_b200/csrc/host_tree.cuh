// host_tree.cuh — host side of the reference's Barnes-Hut branch (kernels: bh.cuh): scratch layout, the
// build / order / walk sequence of one compute_gravity pass, nb_solver_info. A fragment of nbody_b200.cu
// (included there once, after the context helpers it uses: fail(), NB_CUDA/NB_TRY, blocks_for()).
#pragma once

// ---- the reference's Barnes-Hut branch (bh.cuh) ---------------------------------------------------
namespace {

inline size_t bh_align(size_t v) { return (v + 255) & ~(size_t)255; }

struct BhLayout {
    size_t bpos, centre, cell, link, count, cur, key, key_out, order_in, order_out, level, ctl, visits, trace, cub, total;
    size_t cub_bytes;
};

BhLayout bh_layout(long long rows, long long cells) {
    BhLayout L{};
    size_t o = 0;
    auto take = [&](size_t bytes) { const size_t at = o; o += bh_align(bytes); return at; };
    L.ctl = take(sizeof(unsigned int) * kBhCtlWords);
    L.visits = take(sizeof(unsigned long long) * 4);
    L.level = take(sizeof(int) * (kBhMaxLevels + 2));
    L.trace = take(sizeof(unsigned long long) * (2 * kBhMaxLevels + 16));
    L.bpos = take(sizeof(float2) * (size_t)rows);
    L.cur = take(sizeof(int) * (size_t)rows);
    L.key = take(sizeof(unsigned int) * (size_t)rows);
    L.key_out = take(sizeof(unsigned int) * (size_t)rows);
    L.order_in = take(sizeof(int) * (size_t)rows);
    L.order_out = take(sizeof(int) * (size_t)rows);
    L.cub_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, L.cub_bytes, (unsigned int*)nullptr, (unsigned int*)nullptr, (int*)nullptr,
                                    (int*)nullptr, (int)rows);
    L.cub = take(L.cub_bytes);
    L.centre = take(sizeof(float2) * (size_t)cells);
    L.cell = take(sizeof(float4) * (size_t)cells);
    L.link = take(sizeof(int2) * (size_t)cells);
    L.count = take(sizeof(int) * (size_t)cells);
    L.total = o;
    return L;
}

int ensure_bh(nb_ctx* c, long long cells) {
    if (!c->h_bh_ctl) NB_CUDA(cudaMallocHost(reinterpret_cast<void**>(&c->h_bh_ctl), 256));
    if (c->bh_scratch && c->bh_rows >= c->n && c->bh_capacity >= cells) return NB_OK;
    NB_CUDA(cudaStreamSynchronize(c->stream));
    if (c->bh_scratch) cudaFree(c->bh_scratch);
    c->bh_scratch = nullptr;
    c->bh_rows = c->bh_capacity = 0;
    const long long rows = std::max<long long>(c->n, c->bh_rows);
    const BhLayout L = bh_layout(rows, cells);
    if (cudaMalloc(&c->bh_scratch, L.total) != cudaSuccess) {
        cudaGetLastError();
        return fail(NB_ERR_ALLOC, "Barnes-Hut scratch: out of device memory");
    }
    c->bh_bytes = L.total;
    c->bh_rows = rows;
    c->bh_capacity = cells;
    return NB_OK;
}

BhTree bh_view(const nb_ctx* c, const BhLayout& L) {
    char* b = static_cast<char*>(c->bh_scratch);
    BhTree t;
    t.bpos = reinterpret_cast<float2*>(b + L.bpos);
    t.centre = reinterpret_cast<float2*>(b + L.centre);
    t.cell = reinterpret_cast<float4*>(b + L.cell);
    t.link = reinterpret_cast<int2*>(b + L.link);
    t.count = reinterpret_cast<int*>(b + L.count);
    t.cur = reinterpret_cast<int*>(b + L.cur);
    t.key = reinterpret_cast<unsigned int*>(b + L.key);
    t.order_in = reinterpret_cast<int*>(b + L.order_in);
    t.level_begin = reinterpret_cast<int*>(b + L.level);
    t.ctl = reinterpret_cast<unsigned int*>(b + L.ctl);
    t.capacity = (int)c->bh_capacity;
    static const bool trace = getenv("NB_BH_TRACE") != nullptr;  // developer knob: per-phase times on stderr
    t.trace = trace ? reinterpret_cast<unsigned long long*>(b + L.trace) : nullptr;
    return t;
}

}  // namespace

static int exact_force_direct(nb_ctx* c);

// compute_gravity with more than bh_threshold bodies (Physics.hpp:212-230): build, order, walk.
static int tree_force(nb_ctx* c) {
    if (c->n > (1ll << 30)) return fail(NB_ERR_INVALID, "Barnes-Hut branch: at most 2^30 bodies");
    if (c->bh_coop_ctas_per_sm == 0) {
        int coop = 0;
        NB_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c->device));
        if (!coop) return fail(NB_ERR_CUDA, "device does not support cooperative launches");
        NB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c->bh_coop_ctas_per_sm, bh_build_kernel, kBhBuildThreads, 0));
        if (c->bh_coop_ctas_per_sm < 1) return fail(NB_ERR_CUDA, "bh_build_kernel does not fit on an SM");
    }
    long long cells = std::max<long long>(c->bh_capacity, 6 * c->n + 4096);
    for (int attempt = 0;; ++attempt) {
        if (cells > (1ll << 30)) return fail(NB_ERR_ALLOC, "Barnes-Hut tree needs more than 2^30 cells");
        NB_TRY(ensure_bh(c, cells));
        const BhLayout L = bh_layout(c->bh_rows, c->bh_capacity);
        BhTree t = bh_view(c, L);
        StateView s = c->view();
        // resident CTAs only (grid-wide barriers); small worlds do not need the whole device
        // (developer knob; measured: fewer, fatter CTAs make every level's latency-bound sweep longer — 1024 bodies per
        //  CTA costs +10-15 % at N = 4096 ... 65536, profiles/r01_bh_sizes.txt)
        static const long long per_cta = getenv("NB_BH_BODIES_PER_CTA") ? atoll(getenv("NB_BH_BODIES_PER_CTA")) : kBhBuildThreads;
        const long long want = std::max<long long>(1, (std::max<long long>(c->n, 1) + per_cta - 1) / per_cta);
        const unsigned int grid = (unsigned int)std::min<long long>(want, (long long)c->bh_coop_ctas_per_sm * c->sm_count);
        void* args[] = {&s, &t};
        NB_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(bh_build_kernel), dim3(grid), dim3(kBhBuildThreads), args, 0,
                                            c->stream));
        c->kernel_launches++;
        NB_CUDA(cudaMemcpyAsync(c->h_bh_ctl, t.ctl, sizeof(unsigned int) * kBhCtlWords, cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaStreamSynchronize(c->stream));
        const unsigned int err = c->h_bh_ctl[kBhError];
        if (t.trace) {
            std::vector<unsigned long long> tr(2 * kBhMaxLevels + 16);
            NB_CUDA(cudaMemcpy(tr.data(), t.trace, tr.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
            fprintf(stderr, "[bh trace] %d stamps, us since start:", (int)tr[0] - 1);
            for (unsigned long long k = 2; k < tr[0] && k < tr.size(); ++k) fprintf(stderr, " %.1f", (double)(tr[k] - tr[1]) * 1e-3);
            fprintf(stderr, "\n");
        }
        if (err == kBhOverflow) {  // more cells than allocated: the count the build reached is a lower bound
            cells = std::max<long long>(2 * c->bh_capacity, (long long)c->h_bh_ctl[kBhNodes] + 4096);
            continue;
        }
        if (err == kBhTooDeep)
            return fail(NB_ERR_INVALID,
                        "Barnes-Hut branch: two bodies share their float coordinates (the reference's quadtree never "
                        "separates them, SpatialPartition.hpp:109-156)");
        c->bh_live = c->h_bh_ctl[kBhLive];
        if (c->bh_live <= (long long)c->bh_threshold) return exact_force_direct(c);  // :212 counts gathered bodies
        c->bh_nodes = c->h_bh_ctl[kBhNodes];
        c->bh_depth = c->h_bh_ctl[kBhDepth];
        char* b = static_cast<char*>(c->bh_scratch);
        unsigned int* key_out = reinterpret_cast<unsigned int*>(b + L.key_out);
        int* order_out = reinterpret_cast<int*>(b + L.order_out);
        size_t cub_bytes = L.cub_bytes;
        NB_CUDA(cub::DeviceRadixSort::SortPairs(b + L.cub, cub_bytes, t.key, key_out, t.order_in, order_out, (int)c->n, 0, 32,
                                                c->stream));
        unsigned long long* visits = reinterpret_cast<unsigned long long*>(b + L.visits);
        const unsigned int wgrid = blocks_for(c->n, kBhWalkThreads);
        if (c->bh_stats) {
            NB_CUDA(cudaMemsetAsync(visits, 0, sizeof(unsigned long long) * 4, c->stream));
            bh_walk_kernel<true><<<wgrid, kBhWalkThreads, 0, c->stream>>>(s, t, order_out, c->cfg.g, c->cfg.softening,
                                                                          c->bh_theta, visits);
            NB_CUDA(cudaMemcpyAsync(c->h_bh_ctl + 16, visits, sizeof(unsigned long long) * 3, cudaMemcpyDeviceToHost,
                                    c->stream));
        } else {
            bh_walk_kernel<false><<<wgrid, kBhWalkThreads, 0, c->stream>>>(s, t, order_out, c->cfg.g, c->cfg.softening,
                                                                           c->bh_theta, visits);
        }
        NB_CUDA(cudaGetLastError());
        c->kernel_launches += 3;  // sort (histogram + onesweep passes counted as one library call) + walk
        c->force_evals++;
        c->bh_passes++;
        return NB_OK;
    }
}

extern "C" int nb_solver_info(nb_ctx* c, int64_t* values) {
    if (!c || !values) return fail(NB_ERR_INVALID, "nb_solver_info: null");
    NB_CUDA(cudaSetDevice(c->device));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    values[0] = c->bh_passes; values[1] = c->bh_nodes; values[2] = c->bh_depth; values[3] = c->bh_live;
    const unsigned long long* w = reinterpret_cast<const unsigned long long*>(c->h_bh_ctl ? c->h_bh_ctl + 16 : nullptr);
    for (int k = 0; k < 3; ++k) values[4 + k] = (c->bh_stats && w) ? (int64_t)w[k] : 0;
    values[7] = 0;
    return NB_OK;
}
