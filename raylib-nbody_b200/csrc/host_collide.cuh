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
static int overlaps_impl(nb_ctx* c, const double* radius, int64_t* n_overlaps, int64_t* first_pair,
                         unsigned long long* d_pairs, unsigned long long pair_cap, unsigned long long* n_emitted) {
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

// Collision::resolve in its default merge mode (Collision.hpp:49-132). The pair tests run on the device; the
// merges are replayed on the host in the reference's visiting order as an event queue:
//   * pairs whose two bodies are untouched so far behave as the initial detection says;
//   * every merge changes exactly one snapshot (the survivor's): the device re-tests that body against all
//     living bodies (collide_rescan_kernel) and the hits that still lie AHEAD in the (i, j) order are queued;
//   * an event is re-tested with the current snapshots when its turn comes.
// Any pair that overlaps when the reference would test it was detected either initially or by the rescan that
// followed the last change of one of its bodies, so the replay performs exactly the reference's merges.
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

    // scratch of the merge phase: pair sink | cx | cy | hits | live | hit counter
    unsigned long long pair_cap = (unsigned long long)std::max<long long>(4 * n, 1 << 16);
    int64_t overlaps = 0;
    unsigned long long emitted = 0;
    constexpr long long kMergeFirstHits = 64;  // count + this many hits ride in the first readback
    size_t off_cx = 0, off_cy = 0, off_hits = 0, off_live = 0, off_cnt = 0;
    for (;;) {
        off_cx = align256(sizeof(unsigned long long) * pair_cap);
        off_cy = off_cx + align256(sizeof(double) * (size_t)n);
        off_hits = off_cy + align256(sizeof(double) * (size_t)n);
        off_live = off_hits + align256(sizeof(unsigned int) * ((size_t)n + 1));
        off_cnt = off_live + align256((size_t)n);
        const size_t total = off_cnt + 256;
        if (total > c->merge_bytes) {
            NB_CUDA(cudaStreamSynchronize(c->stream));
            if (c->merge_scratch) cudaFree(c->merge_scratch);
            c->merge_scratch = nullptr;
            c->merge_bytes = 0;
            NB_CUDA(cudaMalloc(&c->merge_scratch, total));
            c->merge_bytes = total;
        }
        NB_TRY(overlaps_impl(c, radius, &overlaps, nullptr, static_cast<unsigned long long*>(c->merge_scratch), pair_cap,
                             &emitted));
        if (emitted <= pair_cap) break;
        pair_cap = emitted + 1024;  // more overlapping pairs than the sink held: once more with room for all
    }
    if (overlaps == 0) return NB_OK;  // the common frame: resolve() changes nothing

    char* mb = static_cast<char*>(c->merge_scratch);
    double* d_cx = reinterpret_cast<double*>(mb + off_cx);
    double* d_cy = reinterpret_cast<double*>(mb + off_cy);
    unsigned int* d_hits = reinterpret_cast<unsigned int*>(mb + off_hits);  // [0]: running hit total, [1..]: this launch's hits
    unsigned char* d_live = reinterpret_cast<unsigned char*>(mb + off_live);
    double* d_cr = reinterpret_cast<double*>(static_cast<char*>(c->collide_scratch) + 256);  // the radii overlaps_impl staged

    // host snapshots (Collision.hpp:51-67): everything the merges read and write
    std::vector<unsigned long long> keys((size_t)emitted);
    NB_CUDA(cudaMemcpyAsync(keys.data(), c->merge_scratch, sizeof(unsigned long long) * (size_t)emitted, cudaMemcpyDeviceToHost,
                            c->stream));
    const StateView sv = c->view();
    collide_snapshot_kernel<<<blocks_for(n, 256), 256, 0, c->stream>>>(sv.x[0], sv.x[1], sv.x[2], c->mass, n, d_cx, d_cy, d_live);
    NB_CUDA(cudaMemsetAsync(d_hits, 0, sizeof(unsigned int), c->stream));
    c->kernel_launches++;
    if (!c->h_merge) NB_CUDA(cudaMallocHost(reinterpret_cast<void**>(&c->h_merge), sizeof(unsigned int) * kMergeFirstHits));
    std::vector<double> hp(2 * (size_t)n), hv(2 * (size_t)n), hr(radius, radius + n);
    std::vector<float> hm((size_t)n);
    std::vector<uint8_t> hpin((size_t)n), hlive((size_t)n);
    NB_TRY(download_dim(c, hp.data(), hv.data(), nullptr, 2));
    NB_CUDA(cudaMemcpyAsync(hm.data(), c->mass, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaMemcpyAsync(hpin.data(), c->pinned, (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaMemcpyAsync(hlive.data(), d_live, (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));

    std::set<unsigned long long> events(keys.begin(), keys.end());
    std::vector<unsigned int> hits((size_t)n);
    const unsigned long long un = (unsigned long long)n;
    const double pi = 3.141592653589793238462643383279502884;  // std::numbers::pi
    long long cur_row = -1;
    bool row_alive = false;
    int64_t merges = 0;
    unsigned int hit_base = 0;
    while (!events.empty()) {
        const unsigned long long key = *events.begin();
        events.erase(events.begin());
        const long long a = (long long)(key / un), b = (long long)(key % un);
        if (a != cur_row) {  // alive[i] is looked at when row i starts and not again (:76)
            cur_row = a;
            row_alive = alive[a] != 0;
        }
        if (!row_alive || !alive[b]) continue;  // :76, :78
        const double dx = hp[2 * b] - hp[2 * a];
        const double dy = hp[2 * b + 1] - hp[2 * a + 1];
        const double rsum = hr[a] + hr[b];
        const double dist2 = dx * dx + dy * dy;
        if (dist2 > rsum * rsum) continue;  // :85, with the snapshots as they are NOW
        const bool a_survives = (hm[a] >= hm[b]) || hpin[b];  // :94
        const long long sv_i = a_survives ? a : b, dd = a_survives ? b : a;
        if (hpin[dd] && hpin[sv_i]) continue;  // :97-104
        const double ms = (double)hm[sv_i], md = (double)hm[dd];
        const double M = ms + md;                                            // :107
        const double px = ms * hv[2 * sv_i] + md * hv[2 * dd];               // :108-109
        const double py = ms * hv[2 * sv_i + 1] + md * hv[2 * dd + 1];
        const double xx = (ms * hp[2 * sv_i] + md * hp[2 * dd]) / M;         // :111-112
        const double xy = (ms * hp[2 * sv_i + 1] + md * hp[2 * dd + 1]) / M;
        hm[sv_i] = (float)M;                                                 // :115,128
        hv[2 * sv_i] = px / M;                                               // :110,116
        hv[2 * sv_i + 1] = py / M;
        hp[2 * sv_i] = xx;
        hp[2 * sv_i + 1] = xy;
        hpin[sv_i] = (hpin[sv_i] || hpin[dd]) ? 1 : 0;                       // :120,131
        hr[sv_i] = std::cbrt((3.0 * std::max(1.0, (double)hm[sv_i])) / (4.0 * pi * density[sv_i]));  // :37-46,132
        changed[sv_i] = 1;
        const bool newly_dead = alive[dd] != 0;
        alive[dd] = 0;                                                       // :124-125
        if (merge_log && merges < merge_log_cap) {
            merge_log[2 * merges] = sv_i;
            merge_log[2 * merges + 1] = dd;
        }
        ++merges;
        // which living bodies does the survivor's new state reach? One launch, one readback of the count and
        // the first hits into pinned memory; the (rare) long hit lists take a second copy.
        collide_rescan_kernel<<<blocks_for(n, 256), 256, 0, c->stream>>>(d_cx, d_cy, d_cr, d_live, n, sv_i, xx, xy, hr[sv_i],
                                                                       newly_dead ? dd : -1, d_hits, hit_base);
        c->kernel_launches++;
        const size_t first = (size_t)std::min<long long>(kMergeFirstHits, n + 1);
        NB_CUDA(cudaMemcpyAsync(c->h_merge, d_hits, sizeof(unsigned int) * first, cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaStreamSynchronize(c->stream));
        const unsigned int n_hits = c->h_merge[0] - hit_base;
        hit_base = c->h_merge[0];
        const unsigned int* hit = c->h_merge + 1;
        if ((size_t)n_hits + 1 > first) {
            NB_CUDA(cudaMemcpyAsync(hits.data(), d_hits + 1, sizeof(unsigned int) * n_hits, cudaMemcpyDeviceToHost, c->stream));
            NB_CUDA(cudaStreamSynchronize(c->stream));
            hit = hits.data();
        }
        for (unsigned int h = 0; h < n_hits; ++h) {
            const unsigned long long k = hit[h];
            const unsigned long long lo = std::min<unsigned long long>(k, (unsigned long long)sv_i);
            const unsigned long long hi = std::max<unsigned long long>(k, (unsigned long long)sv_i);
            const unsigned long long nk = lo * un + hi;
            if (nk > key) events.insert(nk);  // only tests the reference has not made yet
        }
    }
    NB_CUDA(cudaGetLastError());
    for (long long i = 0; i < n; ++i) {
        if (!changed[i]) continue;
        pos_xy[2 * i] = hp[2 * i]; pos_xy[2 * i + 1] = hp[2 * i + 1];
        vel_xy[2 * i] = hv[2 * i]; vel_xy[2 * i + 1] = hv[2 * i + 1];
        mass[i] = hm[i];
        pinned[i] = hpin[i];
        radius_out[i] = hr[i];
    }
    *n_merges = merges;
    c->merge_passes++;
    return NB_OK;
}
