// oracle/dropin_check.cpp — TEST INFRASTRUCTURE (run by tests/test_gpu_dropin.py on a B200).
//
// Proves the drop-in claim end to end: two flecs worlds with the reference's own components and
// Config; world R runs the UNMODIFIED reference systems (Physics::register_systems,
// src/systems/Physics.hpp:32-64, compiled from /root/reference where it lies), world B runs
// include/reference_side/PhysicsB200.hpp (gather -> libnbody_b200.so -> scatter). Same inputs,
// same world.progress() calls, same external edits; the components are compared afterwards.
// Built by oracle/Makefile into oracle/_ref/dropin_check (needs the reference's headers).
#include <algorithm>
#include <chrono>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <numbers>
#include <random>
#include <tuple>
#include <vector>

#include <flecs.h>
#include <raylib-cpp.hpp>
#include <raymath.h>

#include "reference_side/PhysicsB200.hpp"

using nbody::Physics;
using nbody::PhysicsB200;

namespace {

struct Snapshot {
    std::vector<double> pos, vel, acc, acc_prev;
};

Snapshot snap(const flecs::world& w) {
    Snapshot s;
    w.each([&](const Position& p, const Velocity& v, const Acceleration& a, const PrevAcceleration& a0) {
        s.pos.push_back(p.value.x); s.pos.push_back(p.value.y);
        s.vel.push_back(v.value.x); s.vel.push_back(v.value.y);
        s.acc.push_back(a.value.x); s.acc.push_back(a.value.y);
        s.acc_prev.push_back(a0.value.x); s.acc_prev.push_back(a0.value.y);
    });
    return s;
}

bool bitwise(const std::vector<double>& a, const std::vector<double>& b) {
    return a.size() == b.size() && std::memcmp(a.data(), b.data(), a.size() * sizeof(double)) == 0;
}

double max_rel(const std::vector<double>& a, const std::vector<double>& b) {
    if (a.size() != b.size()) return INFINITY;
    double scale = 0.0, diff = 0.0;
    for (size_t i = 0; i < a.size(); ++i) {
        scale = std::max(scale, std::fabs(b[i]));
        diff = std::max(diff, std::fabs(a[i] - b[i]));
    }
    return scale > 0.0 ? diff / scale : diff;
}

void make_config(flecs::world& w, int integrator) {
    // flecs' C++ layer caches component ids per type across worlds: register every component in the
    // same order in both worlds so the twin worlds agree on them
    w.component<Config>(); w.component<Position>(); w.component<Velocity>(); w.component<Acceleration>();
    w.component<PrevAcceleration>(); w.component<Mass>(); w.component<Pinned>(); w.component<Radius>();
    w.component<Density>(); w.component<Tint>(); w.component<Trail>(); w.component<Selectable>();
    w.component<Selected>(); w.component<Draggable>(); w.component<Physics::Diagnostics>();
    w.component<PhysicsB200::Device>();
    Config cfg{};
    cfg.bh_threshold = INT_MAX;  // reference world: force the direct-sum branch (Physics.hpp:212)
    cfg.use_fixed_dt = true;
    cfg.draw_trails = false;
    cfg.integrator = integrator;
    w.set<Config>(cfg);
}

void add_disk(flecs::world& w, int n, unsigned seed, float softening, bool physical_radius = false) {
    std::mt19937_64 rng(seed);
    std::uniform_real_distribution<double> u(0.0, 1.0);
    const double R = 3.844e8;
    for (int i = 0; i < n; ++i) {
        const double r = R * std::sqrt(u(rng)), th = 2.0 * std::numbers::pi * u(rng);
        const float x = (float)(r * std::cos(th)), y = (float)(r * std::sin(th));
        const float vx = (float)(-300.0 * std::sin(th)), vy = (float)(300.0 * std::cos(th));
        auto e = w.entity()
                     .set<Position>({DVec2{x, y}})
                     .set<Velocity>({DVec2{vx, vy}})
                     .set<Acceleration>({DVec2{0.0, 0.0}})
                     .set<PrevAcceleration>({DVec2{0.0, 0.0}})
                     .set<Mass>({7.342e22f})
                     .set<Pinned>({i % 97 == 0});
        // Radius{0}: Collision::resolve never merges (Collision.hpp:29,83-85); without the component
        // the radius is derived from mass and density (1.47e6 m here, Collision.hpp:30-34)
        if (!physical_radius) e.set<Radius>({0.0});
    }
    w.get_mut<Config>()->softening = softening;
}

void neutralise_collisions(flecs::world& w) {
    std::vector<flecs::entity> es;
    w.each([&](flecs::entity e, Position&) { es.push_back(e); });
    for (auto& e : es) e.set<Radius>({0.0});
}

int failures = 0;
void expect(bool ok, const char* what) {
    std::printf("  [%s] %s\n", ok ? "ok" : "FAIL", what);
    if (!ok) ++failures;
}

}  // namespace

int main() {
    const char* only = std::getenv("NB_DROPIN_ONLY");  // "F": just the multi-GPU scenario
    if (!(only && only[0] == 'F')) {
    // ---- A: config 1, the built-in three-body seed; n <= 512 runs the FP64 exact path ----------
    for (int integrator : {1, 0}) {
        flecs::world wr, wb;
        make_config(wr, integrator);
        make_config(wb, integrator);
        Physics::register_systems(wr);
        PhysicsB200::register_systems(wb, 0);
        Physics::reset_scenario(wr);
        Physics::reset_scenario(wb);  // the seed itself is host-side scenario code (same call on both)
        neutralise_collisions(wr);
        neutralise_collisions(wb);
        const float dt = wr.get<Config>()->fixed_dt;
        for (int f = 0; f < 2000; ++f) {
            wr.progress(dt);
            wb.progress(dt);
        }
        const Snapshot r = snap(wr), b = snap(wb);
        std::printf("three-body seed, integrator %d, 2000 frames\n", integrator);
        expect(bitwise(r.pos, b.pos) && bitwise(r.vel, b.vel), "positions and velocities bit-identical");
        expect(bitwise(r.acc, b.acc), "accelerations bit-identical");
        if (integrator == 1) expect(bitwise(r.acc_prev, b.acc_prev), "previous accelerations bit-identical");
        const auto* dr = wr.get<Physics::Diagnostics>();
        const auto* db = wb.get<Physics::Diagnostics>();
        expect(dr && db && dr->ok && db->ok && std::fabs(dr->energy - db->energy) <= 1e-12 * std::fabs(dr->energy) &&
                   dr->totalMass == db->totalMass,
               "Diagnostics singleton matches (energy to 1e-12)");
        const auto* dev = wb.get<PhysicsB200::Device>();
        std::printf("  uploads %lld over %lld frames\n", dev->uploads, dev->frames);
        expect(dev->uploads == 1, "state stayed device-resident (one upload)");
        PhysicsB200::shutdown(wb);
    }

    // ---- B: 2048-body disk on the FP32 path, external edits, integrator switch ---------------------
    {
        flecs::world wr, wb;
        make_config(wr, 1);
        make_config(wb, 1);
        Physics::register_systems(wr);
        PhysicsB200::register_systems(wb, 0);
        add_disk(wr, 2048, 99, 2.0e6f);
        add_disk(wb, 2048, 99, 2.0e6f);
        for (flecs::world* w : {&wr, &wb}) w->get_mut<Config>()->time_scale = 3.0e4f;  // dtEff = 250 s
        const float dt = wr.get<Config>()->fixed_dt;
        auto run = [&](int frames) {
            for (int f = 0; f < frames; ++f) {
                wr.progress(dt);
                wb.progress(dt);
            }
        };
        run(5);
        // external edit, as the UI does (UI.hpp:195-210): same edit on both worlds
        for (flecs::world* w : {&wr, &wb}) {
            int k = 0;
            w->each([&](Velocity& v, const Mass&) {
                if (k++ == 7) v.value = DVec2{1234.5, -321.0};
            });
        }
        run(3);
        for (flecs::world* w : {&wr, &wb}) w->get_mut<Config>()->integrator = 0;
        run(3);
        const Snapshot r = snap(wr), b = snap(wb);
        std::printf("2048-body disk, Verlet x5, edit, Verlet x3, Euler x3\n");
        const double ep = max_rel(b.pos, r.pos), ev = max_rel(b.vel, r.vel), ea = max_rel(b.acc, r.acc);
        std::printf("  max |dpos|/max|pos| %.3e  max |dvel|/max|vel| %.3e  max |dacc|/max|acc| %.3e\n", ep, ev, ea);
        // once positions have moved off the float grid the FP32 j-stream carries them to 2^-24 of the
        // extent (32 m here) against a softening of 2e6 m: pair forces agree to ~1.6e-5
        expect(ep < 1e-6 && ev < 5e-5 && ea < 1e-4, "trajectories within the FP32-path bounds");
        const auto* dr = wr.get<Physics::Diagnostics>();
        const auto* db = wb.get<Physics::Diagnostics>();
        expect(dr && db && dr->ok && db->ok && std::fabs(dr->energy - db->energy) <= 2e-6 * std::fabs(dr->energy),
               "Diagnostics energy within 2e-6");
        const auto* dev = wb.get<PhysicsB200::Device>();
        std::printf("  uploads %lld over %lld frames, frame-start diagnostics taken from the previous step %lld times\n",
                    dev->uploads, dev->frames, dev->diag_reused);
        expect(dev->uploads == 2, "one upload at start, one after the external edit, none otherwise");
        expect(dev->diag_reused == 6, "Verlet frames after the first / after the edit reuse the fused diagnostics (4 + 2)");

        // non-finite body: the reference pauses itself (Physics.hpp:140-145); so must the binding
        for (flecs::world* w : {&wr, &wb}) {
            int k = 0;
            w->each([&](Position& p, const Mass&) {
                if (k++ == 3) p.value.x = NAN;
            });
        }
        run(1);
        expect(wr.get<Config>()->paused && wb.get<Config>()->paused, "non-finite state auto-pauses both worlds");
        expect(!wb.get<Physics::Diagnostics>()->ok, "Diagnostics.ok is false");
        PhysicsB200::shutdown(wb);
    }
    // ---- C: real radii: the whole collision pass (Collision.hpp:49-132) through nb_resolve_collisions ------
    for (int dense = 0; dense < 2; ++dense) {
        flecs::world wr, wb;
        make_config(wr, 1);
        make_config(wb, 1);
        Physics::register_systems(wr);
        PhysicsB200::register_systems(wb, 0);
        const int n0 = dense ? 500 : 400, frames = dense ? 15 : 40;  // <= 512 bodies: the FP64-exact gravity path, so states can be compared bitwise
        add_disk(wr, n0, 7, 2.0e6f, true);
        add_disk(wb, n0, 7, 2.0e6f, true);
        for (flecs::world* w : {&wr, &wb}) {
            w->get_mut<Config>()->time_scale = 3.0e4f;
            if (dense) {  // puffy bodies (Density component, Collision.hpp:31-32): chains of merges within one frame
                std::vector<flecs::entity> es;
                w->each([&](flecs::entity e, Position&) { es.push_back(e); });
                for (auto& e : es) e.set<Density>({0.5});
            }
        }
        const float dt = wr.get<Config>()->fixed_dt;
        double worst_diag = 0.0;
        bool diag_ok = true;
        for (int f = 0; f < frames; ++f) {
            wr.progress(dt);
            wb.progress(dt);
            // frame-start diagnostics, merge frames included (there the destructed bodies are still counted, in
            // both worlds, because destruct() is deferred to the end of the system)
            const auto* dr = wr.get<Physics::Diagnostics>();
            const auto* db = wb.get<Physics::Diagnostics>();
            if (getenv("NB_DROPIN_DEBUG")) {
                const Snapshot sr = snap(wr), sb = snap(wb);
                const auto* dv = wb.get<PhysicsB200::Device>();
                std::printf("    frame %d: bodies %zu / %zu, merges so far %lld, pos %s vel %s acc %s, E %.17g / %.17g\n", f,
                            sr.pos.size() / 2, sb.pos.size() / 2, dv->merges, bitwise(sr.pos, sb.pos) ? "same" : "DIFF",
                            bitwise(sr.vel, sb.vel) ? "same" : "DIFF", bitwise(sr.acc, sb.acc) ? "same" : "DIFF",
                            dr ? dr->energy : 0.0, db ? db->energy : 0.0);
                if (sr.pos.size() == sb.pos.size() && !bitwise(sr.pos, sb.pos)) {
                    int shown = 0;
                    for (size_t k = 0; k < sr.pos.size() / 2 && shown < 4; ++k)
                        if (sr.pos[2 * k] != sb.pos[2 * k] || sr.pos[2 * k + 1] != sb.pos[2 * k + 1]) {
                            std::printf("      row %zu: ref (%.17g, %.17g) binding (%.17g, %.17g)\n", k, sr.pos[2 * k], sr.pos[2 * k + 1],
                                        sb.pos[2 * k], sb.pos[2 * k + 1]);
                            ++shown;
                        }
                }
            }
            diag_ok = diag_ok && dr && db && dr->ok == db->ok;
            if (dr && db && dr->ok && db->ok)
                worst_diag = std::max(worst_diag, std::fabs(db->energy - dr->energy) / std::fabs(dr->energy));
        }
        const Snapshot r = snap(wr), b = snap(wb);
        const auto* dev = wb.get<PhysicsB200::Device>();
        std::printf("%d-body disk with %s radii, %d frames: %zu bodies left (reference %zu); device collision passes %lld, "
                    "frames with a merge %lld, merges %lld; max |dE|/|E| of the frame diagnostics %.2e\n",
                    n0, dense ? "puffy (Density 0.5)" : "physical", frames, b.pos.size() / 2, r.pos.size() / 2,
                    dev->collision_checks, dev->collision_frames, dev->merges, worst_diag);
        expect(r.pos.size() == b.pos.size() && (int)r.pos.size() / 2 < n0, "same bodies survive in both worlds");
        expect(bitwise(r.pos, b.pos) && bitwise(r.vel, b.vel) && bitwise(r.acc, b.acc), "state bit-identical after merges");
        expect(diag_ok && worst_diag <= 2e-6, "frame diagnostics agree on every frame, merge frames included");
        expect(dev->collision_checks == frames && dev->collision_frames >= 1 && dev->merges >= (long long)(n0 - r.pos.size() / 2),
               "every frame's pass ran on the device; the reference's CPU pass never ran in the binding's world");
        PhysicsB200::shutdown(wb);
    }
    // ---- D: what the drop-in buys per frame (whole world.progress(): collisions, diagnostics, gravity,
    //         integrate, gather/scatter), 16 384 bodies with physical radii, Velocity Verlet ------------
    {
        flecs::world wr, wb;
        make_config(wr, 1);
        make_config(wb, 1);
        Physics::register_systems(wr);
        PhysicsB200::register_systems(wb, 0);
        add_disk(wr, 16384, 5, 2.0e6f, false);
        add_disk(wb, 16384, 5, 2.0e6f, false);
        for (flecs::world* w : {&wr, &wb}) w->get_mut<Config>()->time_scale = 3.0e4f;
        const float dt = wr.get<Config>()->fixed_dt;
        auto seconds = [&](flecs::world& w, int frames) {
            const auto t0 = std::chrono::steady_clock::now();
            for (int f = 0; f < frames; ++f) w.progress(dt);
            return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() / frames;
        };
        wb.progress(dt);  // first frame uploads
        wr.progress(dt);
        const double tr = seconds(wr, 2), tb = seconds(wb, 20);
        const Snapshot r = snap(wr), b = snap(wb);
        std::printf("16384-body frame (flecs world.progress, Verlet): reference %.3f s, binding %.6f s  (x%.0f)\n", tr, tb, tr / tb);
        expect(tb < tr, "the binding's frame is faster than the reference's");
        PhysicsB200::shutdown(wb);
    }
    // ---- E: the reference at its DEFAULT solver settings (bh_threshold = 100, bh_theta = 0.5): above 100
    //         bodies compute_gravity is the float Barnes-Hut quadtree (Physics.hpp:212-230). The binding with
    //         NB_SOLVER_REFERENCE reproduces those frames bit for bit.
    for (int integrator : {1, 0}) {
        std::printf("E. 4096-body disk, reference default solver (Barnes-Hut above %d bodies, theta %.2f), %s\n",
                    nbody::constants::default_bh_threshold, (double)nbody::constants::default_bh_theta,
                    integrator ? "Velocity Verlet" : "Semi-Implicit Euler");
        flecs::world wr, wb;
        make_config(wr, integrator);
        make_config(wb, integrator);
        Physics::register_systems(wr);
        PhysicsB200::register_systems(wb, 0, NB_SOLVER_REFERENCE);
        add_disk(wr, 4096, 21, 2.0e6f);
        add_disk(wb, 4096, 21, 2.0e6f);
        for (flecs::world* w : {&wr, &wb}) {
            Config* cfg = w->get_mut<Config>();
            cfg->bh_threshold = nbody::constants::default_bh_threshold;
            cfg->bh_theta = nbody::constants::default_bh_theta;
            cfg->time_scale = 3.0e4f;
            cfg->max_substep = 100.0f;  // 3 substeps per frame
        }
        const float dt = wr.get<Config>()->fixed_dt;
        for (int f = 0; f < 12; ++f) { wr.progress(dt); wb.progress(dt); }
        Snapshot r = snap(wr), b = snap(wb);
        expect(bitwise(r.pos, b.pos) && bitwise(r.vel, b.vel), "positions and velocities bit-identical after 12 frames");
        expect(bitwise(r.acc, b.acc) && (integrator == 0 || bitwise(r.acc_prev, b.acc_prev)),
               "accelerations (float-valued tree sums) bit-identical");
        // an external edit and a change of the opening angle mid-run
        for (flecs::world* w : {&wr, &wb}) {
            int k = 0;
            w->each([&](Velocity& v) { if (k++ == 9) v.value.x += 50.0; });
            w->get_mut<Config>()->bh_theta = 0.8f;
        }
        for (int f = 0; f < 8; ++f) { wr.progress(dt); wb.progress(dt); }
        r = snap(wr); b = snap(wb);
        expect(bitwise(r.pos, b.pos) && bitwise(r.vel, b.vel) && bitwise(r.acc, b.acc),
               "still bit-identical after an edit and bh_theta = 0.8");
        const auto* dr = wr.get<Physics::Diagnostics>();
        const auto* db = wb.get<Physics::Diagnostics>();
        expect(dr && db && dr->ok && db->ok && std::fabs(db->energy - dr->energy) <= 1e-6 * std::fabs(dr->energy),
               "frame diagnostics agree to 1e-6 (direct-sum potential in both)");
        expect(wb.get<PhysicsB200::Device>()->uploads == 2, "two uploads: the first frame and the edit");
        PhysicsB200::shutdown(wb);
    }
    {
        flecs::world wr, wb;
        make_config(wr, 1);
        make_config(wb, 1);
        Physics::register_systems(wr);
        PhysicsB200::register_systems(wb, 0, NB_SOLVER_REFERENCE);
        add_disk(wr, 16384, 5, 2.0e6f, false);
        add_disk(wb, 16384, 5, 2.0e6f, false);
        for (flecs::world* w : {&wr, &wb}) {
            w->get_mut<Config>()->bh_threshold = nbody::constants::default_bh_threshold;
            w->get_mut<Config>()->time_scale = 3.0e4f;
        }
        const float dt = wr.get<Config>()->fixed_dt;
        auto seconds = [&](flecs::world& w, int frames) {
            const auto t0 = std::chrono::steady_clock::now();
            for (int f = 0; f < frames; ++f) w.progress(dt);
            return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() / frames;
        };
        wb.progress(dt);
        wr.progress(dt);
        const double tr = seconds(wr, 2), tb = seconds(wb, 20);
        std::printf("16384-body frame at the reference's default settings (Barnes-Hut gravity, O(N^2) collision test and "
                    "diagnostics): reference %.3f s, binding %.6f s  (x%.0f)\n", tr, tb, tr / tb);
        expect(tb < tr, "the binding's default-settings frame is faster than the reference's");
        PhysicsB200::shutdown(wb);
    }
    // ---- G: a scenario written to disk by the binding and applied to a fresh world (Scenario.hpp:53-128) ----------
    {
        flecs::world wa, wb;
        make_config(wa, 1);
        make_config(wb, 0);
        PhysicsB200::register_systems(wa, 0);
        PhysicsB200::register_systems(wb, 0);
        add_disk(wa, 600, 77, 3.0e6f);
        wa.get_mut<Config>()->time_scale = 2.0e4f;
        {   // snapshot_from_world gathers bodies that carry a Tint (Scenario.hpp:77); not inside an iteration
            std::vector<flecs::entity> es;
            wa.each([&](flecs::entity e, const Position&) { es.push_back(e); });
            for (auto& e : es) e.set<Tint>({WHITE});
        }
        const float dt = wa.get<Config>()->fixed_dt;
        for (int f = 0; f < 4; ++f) wa.progress(dt);
        const char* path = "/tmp/nb_dropin_scenario.nbs";
        PhysicsB200::save_scenario(wa, path);
        const nbody::Scenario sc = PhysicsB200::load_scenario(path);
        nbody::apply_scenario_to_world(wb, sc);  // the reference's own apply: bodies + config subset
        neutralise_collisions(wa);
        neutralise_collisions(wb);
        wb.get_mut<Config>()->bh_threshold = INT_MAX;
        const Snapshot a0 = snap(wa), b0 = snap(wb);
        std::printf("600-body scenario: binding -> .nbs file -> reference apply_scenario_to_world\n");
        expect(sc.bodies.size() == 600 && bitwise(a0.pos, b0.pos) && bitwise(a0.vel, b0.vel),
               "bodies round-trip bit for bit");
        expect(wb.get<Config>()->integrator == 1 && wb.get<Config>()->softening == 3.0e6f &&
                   wb.get<Config>()->time_scale == 2.0e4f,
               "config subset applied (integrator, softening, time scale)");
        for (int f = 0; f < 3; ++f) {
            wa.progress(dt);
            wb.progress(dt);
        }
        const Snapshot a1 = snap(wa), b1 = snap(wb);
        // wb restarts with Acceleration = 0 like any applied scenario (Scenario.hpp:96): same frames from there
        expect(max_rel(b1.pos, a1.pos) < 1e-9, "both worlds run on from the scenario");
        PhysicsB200::shutdown(wa);
        PhysicsB200::shutdown(wb);
    }
    }  // scenarios A-E
    // ---- F: several GPUs behind the same flecs systems (nb_create_multi); devices from NB_DROPIN_DEVICES="0,1,.." -
    if (const char* list = std::getenv("NB_DROPIN_DEVICES")) {
        std::vector<int> devices;
        for (const char* q = list; *q;) {
            devices.push_back(std::atoi(q));
            while (*q && *q != ',') ++q;
            if (*q == ',') ++q;
        }
        {
            flecs::world wr, wb;
            make_config(wr, 1);
            make_config(wb, 1);
            Physics::register_systems(wr);
            PhysicsB200::register_systems(wb, devices);
            add_disk(wr, 16384, 5, 4.0f);   // the reference's default softening
            add_disk(wb, 16384, 5, 4.0f);
            // At a softening of 4 m a random disk of this size has pairs a few 10^4 m apart whose accelerations are
            // ~10^4 m/s^2: with the 250 s frames of the other scenarios their motion is numerical noise in both worlds.
            // dtEff = 0.83 s keeps the three frames a comparison of force evaluations and integrator arithmetic.
            for (flecs::world* w : {&wr, &wb}) w->get_mut<Config>()->time_scale = 100.0f;
            const float dt = wr.get<Config>()->fixed_dt;
            for (int f = 0; f < 3; ++f) {
                wr.progress(dt);
                wb.progress(dt);
            }
            const Snapshot r = snap(wr), b = snap(wb);
            std::printf("16384-body disk on %zu GPUs in one process, softening 4 m, Verlet x3\n", devices.size());
            const double ep = max_rel(b.pos, r.pos), ev = max_rel(b.vel, r.vel), ea = max_rel(b.acc, r.acc);
            std::printf("  max |dpos|/max|pos| %.3e  max |dvel|/max|vel| %.3e  max |dacc|/max|acc| %.3e\n", ep, ev, ea);
            expect(ep < 1e-9 && ev < 1e-5 && ea < 1e-5, "multi-GPU trajectories within the FP32-path bounds");
            const auto* dr = wr.get<Physics::Diagnostics>();
            const auto* db = wb.get<Physics::Diagnostics>();
            expect(dr && db && dr->ok && db->ok && std::fabs(dr->energy - db->energy) <= 2e-6 * std::fabs(dr->energy),
                   "Diagnostics energy within 2e-6");
            expect(wb.get<PhysicsB200::Device>()->uploads == 1, "state stayed device-resident (one upload)");
            PhysicsB200::shutdown(wb);
        }
        // whole-frame time of the binding at 2^17 bodies: one GPU beside all of them (no reference twin: one of its
        // frames takes minutes at this size)
        double t_one = 0.0, t_all = 0.0;
        for (int pass = 0; pass < 2; ++pass) {
            flecs::world wb;
            make_config(wb, 1);
            if (pass == 0) PhysicsB200::register_systems(wb, std::vector<int>{devices[0]});
            else PhysicsB200::register_systems(wb, devices);
            add_disk(wb, 131072, 6, 1.0e6f);
            wb.get_mut<Config>()->time_scale = 3.0e4f;
            const float dt = wb.get<Config>()->fixed_dt;
            for (int f = 0; f < 3; ++f) wb.progress(dt);
            const auto t0 = std::chrono::steady_clock::now();
            for (int f = 0; f < 10; ++f) wb.progress(dt);
            const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() / 10;
            (pass == 0 ? t_one : t_all) = sec;
            PhysicsB200::shutdown(wb);
        }
        std::printf("131072-body frame through the binding (gather, step + diagnostics, scatter): 1 GPU %.4f s, %zu GPUs %.4f s\n",
                    t_one, devices.size(), t_all);
    }
    std::printf(failures ? "DROPIN CHECK FAILED (%d)\n" : "DROPIN CHECK PASSED\n", failures);
    return failures ? 1 : 0;
}
