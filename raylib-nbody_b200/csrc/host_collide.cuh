// host_collide.cuh — host side of the collision pass (kernels: collide.cuh): the detection driver shared by
// nb_count_overlaps and nb_resolve_collisions, the grid planner, and the reference-ordered replay of the merges.
// A fragment of nbody_b200.cu (included there once, after download_dim() and the context helpers it uses).
#pragma once

namespace {

int ensure_collide_scratch(nb_ctx* c, size_t bytes) {
    if (bytes <= c->collide_bytes) return NB_OK;
    if (c->collide_scratch) cudaFree(c->collide_scratch);
    c->collide_scratch = nullptr;
    c->collide_bytes = 0;
    NB_CUDA(cudaMalloc(&c->collide_scratch, bytes));
    c->collide_bytes = bytes;
    return NB_OK;
}

inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

// Grid for the broad phase, or cells[0] == 0 when the all-pairs sweep is the better (or only) choice.
GridSpec plan_grid(int dim, const double lo[3], const double hi[3], double r_max, long long n) {
    GridSpec g{};
    g.dim = dim;
    double ext[3] = {hi[0] - lo[0], hi[1] - lo[1], dim == 3 ? hi[2] - lo[2] : 0.0};
    const double ext_max = std::max(ext[0], std::max(ext[1], ext[2]));
    if (!(ext_max > 0.0) || !std::isfinite(ext_max) || !(r_max >= 0.0) || !std::isfinite(r_max)) return g;
    double h = 2.0 * r_max * (1.0 + 1e-9);
    h = std::max(h, ext_max * 9.5367431640625e-07);  // at most 2^20 cells per axis
    const double cap = (double)std::max<long long>(4 * n, 1 << 16);
    for (int it = 0; it < 8; ++it) {
        double prod = 1.0;
        for (int k = 0; k < dim; ++k) prod *= std::floor(ext[k] / h) + 1.0;
        if (prod <= cap) break;
        h *= std::pow(prod / cap, 1.0 / dim) * 1.01;
    }
    long long prod = 1;
    for (int k = 0; k < 3; ++k) {
        g.cells[k] = k < dim ? (int)std::floor(ext[k] / h) + 1 : 1;
        g.origin[k] = lo[k];
        prod *= g.cells[k];
    }
    g.inv_h = 1.0 / h;
    // a neighbourhood that covers most of the world is an all-pairs sweep with extra steps
    if ((double)prod > cap * 1.05 || std::max(g.cells[0], g.cells[1]) < 8) g.cells[0] = 0;
    return g;
}

}  // namespace

// Detection over all pairs (all-pairs sweep or grid broad phase). d_pairs (optional, device): sink for the
// overlapping pairs' keys i * n + j; *n_emitted = how many the kernels tried to append.
struct GridUse {  // what the broad phase built (valid until the next call that touches collide_scratch)
    GridSpec g{};
    const unsigned int* cell_start = nullptr;
    const unsigned int* sorted_idx = nullptr;
    double r_max = 0.0;
};

static int overlaps_impl(nb_ctx* c, const double* radius, int64_t* n_overlaps, int64_t* first_pair,
                         unsigned long long* d_pairs, unsigned long long pair_cap, unsigned long long* n_emitted,
                         GridUse* used = nullptr) {
    NB_CUDA(cudaSetDevice(c->device));
    const long long n = c->n;
    const StateView s = c->view();
    double r_max = 0.0;
    for (long long i = 0; i < n; ++i) r_max = std::fabs(radius[i]) > r_max || std::isnan(radius[i]) ? std::fabs(radius[i]) : r_max;
    const char* mode = getenv("NB_OVERLAP_MODE");  // developer/test knob: "pairs" or "grid"
    const bool want_grid = !(mode && !strcmp(mode, "pairs")) && (n >= 2048 || (mode && !strcmp(mode, "grid")));

    // scratch layout: header (bounds[6], result[2]) | radius | grid arrays
    const size_t off_radius = 256, off_grid = off_radius + align256(sizeof(double) * (size_t)n);
    NB_TRY(ensure_collide_scratch(c, off_grid));
    char* base = static_cast<char*>(c->collide_scratch);
    unsigned long long* d_bounds = reinterpret_cast<unsigned long long*>(base);
    unsigned long long* d_res = d_bounds + 6;
    double* d_radius = reinterpret_cast<double*>(base + off_radius);
    const unsigned long long init[9] = {~0ull, 0ull, ~0ull, 0ull, ~0ull, 0ull, 0ull, ~0ull, 0ull};
    NB_CUDA(cudaMemcpyAsync(d_bounds, init, sizeof(init), cudaMemcpyHostToDevice, c->stream));
    NB_CUDA(cudaMemcpyAsync(d_radius, radius, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, c->stream));

    GridSpec g{};
    if (want_grid) {
        const unsigned int grid = (unsigned int)std::min<long long>(blocks_for(n, 256), 4 * c->sm_count);
        collide_bounds_kernel<<<grid, 256, 0, c->stream>>>(s.x[0], s.x[1], s.x[2], c->mass, n, d_bounds);
        c->kernel_launches++;
        unsigned long long hb[6];
        NB_CUDA(cudaMemcpyAsync(hb, d_bounds, sizeof(hb), cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaStreamSynchronize(c->stream));
        if (hb[0] <= hb[1]) {  // at least one body passed the filter
            double lo[3], hi[3];
            for (int k = 0; k < 3; ++k) { lo[k] = dbl_from_order_key(hb[2 * k]); hi[k] = dbl_from_order_key(hb[2 * k + 1]); }
            g = plan_grid(c->dim, lo, hi, r_max, n);
        }
    }
    if (g.cells[0] > 0) {
        const size_t n_cells = (size_t)g.cells[0] * g.cells[1] * g.cells[2];
        size_t cub_bytes = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, cub_bytes, (unsigned int*)nullptr, (unsigned int*)nullptr, (int)(n_cells + 1), c->stream);
        const size_t off_count = off_grid, off_start = off_count + align256(4 * (n_cells + 1)),
                     off_fill = off_start + align256(4 * (n_cells + 1)), off_sorted = off_fill + align256(4 * n_cells),
                     off_cub = off_sorted + align256(4 * (size_t)n), total = off_cub + align256(cub_bytes);
        if (total > c->collide_bytes) {  // growing frees the old block: re-stage the header and radii
            NB_CUDA(cudaStreamSynchronize(c->stream));
            NB_TRY(ensure_collide_scratch(c, total));
            base = static_cast<char*>(c->collide_scratch);
            d_bounds = reinterpret_cast<unsigned long long*>(base);
            d_res = d_bounds + 6;
            d_radius = reinterpret_cast<double*>(base + off_radius);
            NB_CUDA(cudaMemcpyAsync(d_bounds, init, sizeof(init), cudaMemcpyHostToDevice, c->stream));
            NB_CUDA(cudaMemcpyAsync(d_radius, radius, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
        }
        unsigned int* cell_count = reinterpret_cast<unsigned int*>(base + off_count);
        unsigned int* cell_start = reinterpret_cast<unsigned int*>(base + off_start);
        unsigned int* cell_fill = reinterpret_cast<unsigned int*>(base + off_fill);
        unsigned int* sorted_idx = reinterpret_cast<unsigned int*>(base + off_sorted);
        NB_CUDA(cudaMemsetAsync(cell_count, 0, 4 * (n_cells + 1), c->stream));
        NB_CUDA(cudaMemsetAsync(cell_fill, 0, 4 * n_cells, c->stream));
        grid_bin_kernel<<<blocks_for(n, 256), 256, 0, c->stream>>>(s.x[0], s.x[1], s.x[2], c->mass, n, g, cell_count, nullptr,
                                                                  nullptr, nullptr);
        NB_CUDA(cub::DeviceScan::ExclusiveSum(base + off_cub, cub_bytes, cell_count, cell_start, (int)(n_cells + 1), c->stream));
        grid_bin_kernel<<<blocks_for(n, 256), 256, 0, c->stream>>>(s.x[0], s.x[1], s.x[2], c->mass, n, g, nullptr, cell_start,
                                                                  cell_fill, sorted_idx);
        grid_overlap_kernel<<<blocks_for(n, 256), 256, 0, c->stream>>>(s.x[0], s.x[1], s.x[2], c->mass, d_radius, n, g,
                                                                      cell_start, sorted_idx, d_res, d_pairs, pair_cap);
        c->kernel_launches += 4;
        c->overlap_grid_calls++;
        if (used) {
            used->g = g;
            used->cell_start = cell_start;
            used->sorted_idx = sorted_idx;
        }
    } else {
        if (c->dim == 2)
            overlap_kernel<true><<<blocks_for(n, kOverlapThreads), kOverlapThreads, 0, c->stream>>>(
                s.x[0], s.x[1], s.x[2], c->mass, d_radius, n, d_res, d_pairs, pair_cap);
        else
            overlap_kernel<false><<<blocks_for(n, kOverlapThreads), kOverlapThreads, 0, c->stream>>>(
                s.x[0], s.x[1], s.x[2], c->mass, d_radius, n, d_res, d_pairs, pair_cap);
        c->kernel_launches++;
    }
    NB_CUDA(cudaGetLastError());
    if (used) used->r_max = r_max;
    unsigned long long res[3];
    NB_CUDA(cudaMemcpyAsync(res, d_res, sizeof(res), cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    *n_overlaps = (int64_t)res[0];
    if (first_pair && res[0]) {
        first_pair[0] = (int64_t)(res[1] / (unsigned long long)n);
        first_pair[1] = (int64_t)(res[1] % (unsigned long long)n);
    }
    if (n_emitted) *n_emitted = res[2];
    return NB_OK;
}

extern "C" int nb_count_overlaps(nb_ctx* c, const double* radius, int64_t* n_overlaps, int64_t* first_pair) {
    if (!c || !n_overlaps) return fail(NB_ERR_INVALID, "nb_count_overlaps: null");
    if (c->world > 1) return fail(NB_ERR_INVALID, "nb_count_overlaps: single-GPU contexts only (FP64 positions are not replicated)");
    *n_overlaps = 0;
    if (first_pair) first_pair[0] = first_pair[1] = -1;
    if (c->n < 2) return NB_OK;  // Collision.hpp:70
    if (!radius) return fail(NB_ERR_INVALID, "nb_count_overlaps: null radius array");
    return overlaps_impl(c, radius, n_overlaps, first_pair, nullptr, 0, nullptr);
}

// Collision::resolve in its default merge mode (Collision.hpp:49-132), entirely on the device: the kernels of
// collide.cuh detect the initially overlapping pairs, merge_replay_kernel (merge.cuh) replays the merges in the
// reference's visiting order as an event queue — re-tests, merge arithmetic (incl. the radius' cbrt) and the
// neighbourhood scans after every merge — with no host round trip per merge. The host only stages the caller's
// radii/densities, sorts the initial pair keys (cub) and reads the outcome back.
extern "C" int nb_resolve_collisions(nb_ctx* c, const double* radius, const double* density, int64_t* n_merges,
                                     uint8_t* alive, uint8_t* changed, double* pos_xy, double* vel_xy, float* mass,
                                     uint8_t* pinned, double* radius_out, int64_t* merge_log, int64_t merge_log_cap) {
    if (!c || !n_merges || !alive || !changed) return fail(NB_ERR_INVALID, "nb_resolve_collisions: null");
    if (c->world > 1) return fail(NB_ERR_INVALID, "nb_resolve_collisions: single-GPU contexts only");
    if (c->dim != 2) return fail(NB_ERR_INVALID, "nb_resolve_collisions: the reference's collision pass is 2-D (nb_upload)");
    *n_merges = 0;
    const long long n = c->n;
    for (long long i = 0; i < n; ++i) { alive[i] = 1; changed[i] = 0; }
    if (n < 2) return NB_OK;  // Collision.hpp:70
    if (!radius || !density || !pos_xy || !vel_xy || !mass || !pinned || !radius_out)
        return fail(NB_ERR_INVALID, "nb_resolve_collisions: null array");
    NB_CUDA(cudaSetDevice(c->device));
    if (merge_log_cap < 0 || !merge_log) merge_log_cap = 0;

    // ---- detection: every overlapping pair's key i * n + j into the head of the merge scratch -----------
    unsigned long long pair_cap = (unsigned long long)std::max<long long>(4 * n, 1 << 16);
    int64_t overlaps = 0;
    unsigned long long emitted = 0;
    GridUse used;
    for (;;) {
        const size_t need = align256(sizeof(unsigned long long) * pair_cap);
        if (need > c->merge_bytes) {
            NB_CUDA(cudaStreamSynchronize(c->stream));
            if (c->merge_scratch) cudaFree(c->merge_scratch);
            c->merge_scratch = nullptr;
            c->merge_bytes = 0;
            NB_CUDA(cudaMalloc(&c->merge_scratch, need));
            c->merge_bytes = need;
        }
        used = GridUse{};
        NB_TRY(overlaps_impl(c, radius, &overlaps, nullptr, static_cast<unsigned long long*>(c->merge_scratch), pair_cap,
                             &emitted, &used));
        if (emitted <= pair_cap) break;
        pair_cap = emitted + 1024;  // more overlapping pairs than the sink held: once more with room for all
    }
    if (overlaps == 0) return NB_OK;  // the common frame: resolve() changes nothing

    // ---- replay scratch (its own allocation: the pair sink above stays where it is) ------------------------
    unsigned long long heap_cap = (unsigned long long)std::max<long long>(8 * n, 1 << 20);
    for (int attempt = 0;; ++attempt) {
        size_t cub_bytes = 0;
        NB_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, cub_bytes, (unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                               (int)emitted, 0, 64, c->stream));
        size_t off = 0;
        auto take = [&](size_t bytes) { const size_t at = off; off += align256(bytes); return at; };
        const size_t o_sorted = take(8 * (size_t)emitted), o_cub = take(cub_bytes), o_heap = take(8 * (size_t)heap_cap),
                     o_cx = take(8 * (size_t)n), o_cy = take(8 * (size_t)n), o_vx = take(8 * (size_t)n),
                     o_vy = take(8 * (size_t)n), o_rho = take(8 * (size_t)n), o_cm = take(4 * (size_t)n),
                     o_hits = take(4 * (size_t)n), o_pin = take((size_t)n), o_alive = take((size_t)n),
                     o_live = take((size_t)n), o_changed = take((size_t)n), o_log = take(16 * (size_t)std::max<int64_t>(merge_log_cap, 1)),
                     o_out = take(64);
        if (off > c->replay_bytes) {
            NB_CUDA(cudaStreamSynchronize(c->stream));
            if (c->replay_scratch) cudaFree(c->replay_scratch);
            c->replay_scratch = nullptr;
            c->replay_bytes = 0;
            NB_CUDA(cudaMalloc(&c->replay_scratch, off));
            c->replay_bytes = off;
        }
        char* rb = static_cast<char*>(c->replay_scratch);
        unsigned long long* d_sorted = reinterpret_cast<unsigned long long*>(rb + o_sorted);
        NB_CUDA(cub::DeviceRadixSort::SortKeys(rb + o_cub, cub_bytes, static_cast<unsigned long long*>(c->merge_scratch), d_sorted,
                                               (int)emitted, 0, 64, c->stream));
        MergeParams mp{};
        mp.n = n;
        mp.cx = reinterpret_cast<double*>(rb + o_cx); mp.cy = reinterpret_cast<double*>(rb + o_cy);
        mp.vx = reinterpret_cast<double*>(rb + o_vx); mp.vy = reinterpret_cast<double*>(rb + o_vy);
        mp.cr = reinterpret_cast<double*>(static_cast<char*>(c->collide_scratch) + 256);  // the radii overlaps_impl staged
        mp.cm = reinterpret_cast<float*>(rb + o_cm);
        mp.pin = reinterpret_cast<unsigned char*>(rb + o_pin);
        mp.alive = reinterpret_cast<unsigned char*>(rb + o_alive);
        mp.live = reinterpret_cast<unsigned char*>(rb + o_live);
        mp.changed = reinterpret_cast<unsigned char*>(rb + o_changed);
        double* d_rho = reinterpret_cast<double*>(rb + o_rho);
        mp.density = d_rho;
        const StateView sv = c->view();
        mp.x_init = sv.x[0]; mp.y_init = sv.x[1];
        mp.initial = d_sorted; mp.n_initial = emitted;
        mp.heap = reinterpret_cast<unsigned long long*>(rb + o_heap); mp.heap_cap = heap_cap;
        mp.hits = reinterpret_cast<unsigned int*>(rb + o_hits);
        mp.g = used.g; mp.cell_start = used.cell_start; mp.sorted_idx = used.sorted_idx; mp.r_max = used.r_max;
        mp.merge_log = reinterpret_cast<long long*>(rb + o_log); mp.merge_log_cap = merge_log_cap;
        mp.out = reinterpret_cast<long long*>(rb + o_out);
        NB_CUDA(cudaMemcpyAsync(d_rho, density, 8 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
        if (attempt > 0)  // a retry starts from the caller's radii again (the first replay grew the survivors')
            NB_CUDA(cudaMemcpyAsync(mp.cr, radius, 8 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
        merge_snapshot_kernel<<<blocks_for(n, 256), 256, 0, c->stream>>>(sv.x[0], sv.x[1], sv.x[2], sv.v[0], sv.v[1], c->mass,
                                                                        c->pinned, n, mp);
        merge_replay_kernel<<<1, kMergeThreads, 0, c->stream>>>(mp);
        NB_CUDA(cudaGetLastError());
        c->kernel_launches += 2;
        long long out[4] = {0, 0, 0, 0};
        NB_CUDA(cudaMemcpyAsync(out, mp.out, sizeof(out), cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaStreamSynchronize(c->stream));
        if (out[1] != 0) {  // the event heap overflowed: replay again with a larger one
            if (attempt >= 6) return fail(NB_ERR_INVALID, "nb_resolve_collisions: event queue kept overflowing");
            heap_cap *= 4;
            continue;
        }
        const int64_t merges = out[0];
        c->merge_last = out[0];
        c->merge_initial = (long long)emitted;
        c->merge_events = out[2];
        c->merge_candidates = out[3];
        // ---- the outcome: dead bodies, survivors' new components, the merges in order ----------------------------
        std::vector<double> hx((size_t)n), hy((size_t)n), hvx((size_t)n), hvy((size_t)n), hr((size_t)n);
        NB_CUDA(cudaMemcpyAsync(alive, mp.alive, (size_t)n, cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaMemcpyAsync(changed, mp.changed, (size_t)n, cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaMemcpyAsync(hx.data(), mp.cx, 8 * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaMemcpyAsync(hy.data(), mp.cy, 8 * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaMemcpyAsync(hvx.data(), mp.vx, 8 * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaMemcpyAsync(hvy.data(), mp.vy, 8 * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaMemcpyAsync(hr.data(), mp.cr, 8 * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
        std::vector<float> hm((size_t)n);
        std::vector<uint8_t> hpin((size_t)n);
        NB_CUDA(cudaMemcpyAsync(hm.data(), mp.cm, 4 * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaMemcpyAsync(hpin.data(), mp.pin, (size_t)n, cudaMemcpyDeviceToHost, c->stream));
        const int64_t logged = std::min<int64_t>(merges, merge_log_cap);
        if (logged > 0)
            NB_CUDA(cudaMemcpyAsync(merge_log, mp.merge_log, 16 * (size_t)logged, cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaStreamSynchronize(c->stream));
        for (long long i = 0; i < n; ++i) {
            if (!changed[i]) continue;
            pos_xy[2 * i] = hx[(size_t)i]; pos_xy[2 * i + 1] = hy[(size_t)i];
            vel_xy[2 * i] = hvx[(size_t)i]; vel_xy[2 * i + 1] = hvy[(size_t)i];
            mass[i] = hm[(size_t)i];
            pinned[i] = hpin[(size_t)i];
            radius_out[i] = hr[(size_t)i];
        }
        *n_merges = merges;
        c->merge_passes++;
        return NB_OK;
    }
}

// What the last nb_resolve_collisions replay did: {merges, events popped and re-tested, bodies visited by the
// scans after the merges, initially overlapping pairs}.
extern "C" int nb_merge_info(nb_ctx* c, int64_t* values) {
    if (!c || !values) return fail(NB_ERR_INVALID, "nb_merge_info: null");
    values[0] = c->merge_last;
    values[1] = c->merge_events;
    values[2] = c->merge_candidates;
    values[3] = c->merge_initial;
    return NB_OK;
}

// glibc's cbrt restated for the device (merge.cuh), for the test that pins it against libm.
extern "C" int nb_debug_cbrt(nb_ctx* c, const double* in, double* out, int64_t n) {
    if (!c || !in || !out || n < 0) return fail(NB_ERR_INVALID, "nb_debug_cbrt: bad arguments");
    if (n == 0) return NB_OK;
    NB_CUDA(cudaSetDevice(c->device));
    double* d = nullptr;
    NB_CUDA(cudaMalloc(&d, 16 * (size_t)n));
    NB_CUDA(cudaMemcpyAsync(d, in, 8 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    cbrt_probe_kernel<<<blocks_for(n, 256), 256, 0, c->stream>>>(d, d + n, (int)n);
    NB_CUDA(cudaMemcpyAsync(out, d + n, 8 * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    cudaFree(d);
    return NB_OK;
}
