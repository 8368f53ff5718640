// host_scenario.cuh — on-disk scenario / checkpoint files shared by every binding (host-only, no device work).
//
// The reference keeps scenarios in memory (src/core/Scenario.hpp:14-128: BodySnapshot{pos, vel, mass, pinned} +
// the Config subset that replays them); nothing is ever written to disk. This is the one file format behind
// scenarios.py::save_snapshot/load_snapshot (".nbs" paths) and PhysicsB200::save_scenario/load_scenario:
//   header  : magic "NBS1", n, dim, flags, frame, order_age, nb_scenario_config
//   arrays  : pos[n*dim] f64, vel[n*dim] f64, mass[n] f32, pinned[n] u8,
//             then, when their flag is set: acc[n*dim] f64, acc_prev[n*dim] f64, order[n] i32
// Little-endian, no padding between arrays; floats round-trip exactly. A fragment of nbody_b200.cu.
#pragma once

namespace {

struct ScenarioHeader {
    char magic[4];
    int32_t dim;
    int64_t n;
    int32_t flags;
    int32_t reserved;
    int64_t frame;
    int64_t order_age;
    nb_scenario_config cfg;
};

int scenario_io_fail(const char* what, const char* path) {
    return fail(NB_ERR_INVALID, std::string("nb_scenario: ") + what + " '" + (path ? path : "(null)") + "'");
}

}  // namespace

extern "C" int nb_scenario_default_config(nb_scenario_config* out) {
    if (!out) return fail(NB_ERR_INVALID, "nb_scenario_default_config: null");
    nb_default_config(&out->physics);
    out->bh_threshold = 100;     // Constants.hpp:90
    out->bh_theta = 0.5f;        // :91
    out->use_fixed_dt = 0;       // Config.hpp / Scenario.hpp:34
    out->fixed_dt = 1.0f / 120.0f;
    out->time_scale = 1.0e6f;
    return NB_OK;
}

extern "C" int nb_scenario_write(const char* path, int64_t n, int dim, const double* pos, const double* vel,
                                 const double* acc, const double* acc_prev, const float* mass, const uint8_t* pinned,
                                 const nb_scenario_config* cfg, const int32_t* order, int64_t order_age, int64_t frame) {
    if (!path || n < 0 || (dim != 2 && dim != 3) || (n > 0 && (!pos || !vel || !mass)))
        return fail(NB_ERR_INVALID, "nb_scenario_write: bad arguments");
    FILE* f = fopen(path, "wb");
    if (!f) return scenario_io_fail("cannot create", path);
    ScenarioHeader h;
    memset(&h, 0, sizeof(h));
    memcpy(h.magic, "NBS1", 4);
    h.dim = dim;
    h.n = n;
    h.flags = (acc ? NB_SCENARIO_HAS_ACC : 0) | (acc_prev ? NB_SCENARIO_HAS_ACC_PREV : 0) | (order ? NB_SCENARIO_HAS_ORDER : 0);
    h.frame = frame;
    h.order_age = order ? order_age : 0;
    if (cfg) h.cfg = *cfg;
    else nb_scenario_default_config(&h.cfg);
    const size_t un = (size_t)n, v = un * (size_t)dim;
    std::vector<uint8_t> zeros;
    if (!pinned) zeros.assign(un, 0);
    bool ok = fwrite(&h, sizeof(h), 1, f) == 1;
    ok = ok && (v == 0 || (fwrite(pos, 8, v, f) == v && fwrite(vel, 8, v, f) == v));
    ok = ok && (un == 0 || (fwrite(mass, 4, un, f) == un && fwrite(pinned ? pinned : zeros.data(), 1, un, f) == un));
    if (acc) ok = ok && (v == 0 || fwrite(acc, 8, v, f) == v);
    if (acc_prev) ok = ok && (v == 0 || fwrite(acc_prev, 8, v, f) == v);
    if (order) ok = ok && (un == 0 || fwrite(order, 4, un, f) == un);
    ok = (fclose(f) == 0) && ok;
    return ok ? NB_OK : scenario_io_fail("short write to", path);
}

extern "C" int nb_scenario_read_header(const char* path, int64_t* n, int* dim, int* flags, nb_scenario_config* cfg,
                                       int64_t* order_age, int64_t* frame) {
    if (!path) return fail(NB_ERR_INVALID, "nb_scenario_read_header: null path");
    FILE* f = fopen(path, "rb");
    if (!f) return scenario_io_fail("cannot open", path);
    ScenarioHeader h;
    const bool ok = fread(&h, sizeof(h), 1, f) == 1;
    fclose(f);
    if (!ok || memcmp(h.magic, "NBS1", 4) != 0 || h.n < 0 || (h.dim != 2 && h.dim != 3))
        return scenario_io_fail("not a scenario file:", path);
    if (n) *n = h.n;
    if (dim) *dim = h.dim;
    if (flags) *flags = h.flags;
    if (cfg) *cfg = h.cfg;
    if (order_age) *order_age = h.order_age;
    if (frame) *frame = h.frame;
    return NB_OK;
}

// Arrays sized from nb_scenario_read_header; any output pointer may be null (that array is skipped).
extern "C" int nb_scenario_read(const char* path, double* pos, double* vel, double* acc, double* acc_prev, float* mass,
                                uint8_t* pinned, int32_t* order) {
    if (!path) return fail(NB_ERR_INVALID, "nb_scenario_read: null path");
    FILE* f = fopen(path, "rb");
    if (!f) return scenario_io_fail("cannot open", path);
    ScenarioHeader h;
    if (fread(&h, sizeof(h), 1, f) != 1 || memcmp(h.magic, "NBS1", 4) != 0 || h.n < 0 || (h.dim != 2 && h.dim != 3)) {
        fclose(f);
        return scenario_io_fail("not a scenario file:", path);
    }
    const size_t un = (size_t)h.n, v = un * (size_t)h.dim;
    auto take = [&](void* dst, size_t elem, size_t count) {
        if (count == 0) return true;
        if (dst) return fread(dst, elem, count, f) == count;
        return fseek(f, (long)(elem * count), SEEK_CUR) == 0;
    };
    bool ok = take(pos, 8, v) && take(vel, 8, v) && take(mass, 4, un) && take(pinned, 1, un);
    if (h.flags & NB_SCENARIO_HAS_ACC) ok = ok && take(acc, 8, v);
    if (h.flags & NB_SCENARIO_HAS_ACC_PREV) ok = ok && take(acc_prev, 8, v);
    if (h.flags & NB_SCENARIO_HAS_ORDER) ok = ok && take(order, 4, un);
    fclose(f);
    return ok ? NB_OK : scenario_io_fail("truncated file:", path);
}
