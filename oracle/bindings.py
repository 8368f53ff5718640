"""oracle/bindings.py — TEST INFRASTRUCTURE, not product code.

ctypes bindings for the two CPU checkers:
  * ``Oracle``    -> oracle/liboracle.so       (plain-C restatement, oracle/nbody_oracle.c)
  * ``Reference`` -> oracle/_ref/libnbref.so   (the unmodified reference, oracle/ref_shim.cpp)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
import this module. The product package never does.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libnbref.so")

_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)
_bp = C.POINTER(C.c_uint8)


def _d(a):
    return None if a is None else a.ctypes.data_as(_dp)


def _f(a):
    return None if a is None else a.ctypes.data_as(_fp)


def _b(a):
    return None if a is None else a.ctypes.data_as(_bp)


class OrcConfig(C.Structure):
    _fields_ = [("g", C.c_double), ("softening", C.c_float), ("max_speed", C.c_float),
                ("integrator", C.c_int), ("max_substep", C.c_float), ("max_substeps_per_frame", C.c_int),
                ("bh_threshold", C.c_int), ("bh_theta", C.c_float)]


# Reference defaults (src/core/Constants.hpp:86-100, src/core/Config.hpp:25).
DEFAULTS = dict(g=6.67430e-11, softening=4.0, max_speed=0.0, integrator=1, fixed_dt=float(np.float32(1.0) / np.float32(120)),
                time_scale=1e6, max_substep=1.0e9, max_substeps_per_frame=200)
# Solver switch (Config.hpp:17-18). The reference's own defaults are REF_BH_THRESHOLD / REF_BH_THETA
# (Constants.hpp:90-91); the checkers default to "never" so the direct sum — the north-star path — is
# what a test gets unless it asks for the tree.
BODY_DENSITY = 5510.0  # Constants.hpp:34
BH_NEVER = 2**31 - 1
REF_BH_THRESHOLD = 100
REF_BH_THETA = 0.5
DEFAULTS.update(bh_threshold=BH_NEVER, bh_theta=REF_BH_THETA)


def make_config(**kw):
    d = dict(DEFAULTS)
    d.update(kw)
    return OrcConfig(d["g"], d["softening"], d["max_speed"], d["integrator"], d["max_substep"],
                     d["max_substeps_per_frame"], d["bh_threshold"], d["bh_theta"])


def dt_eff(fixed_dt=DEFAULTS["fixed_dt"], time_scale=DEFAULTS["time_scale"]):
    """dtEff as the integrate system computes it in float (Physics.hpp:54-55)."""
    return float(np.float32(fixed_dt) * max(np.float32(0.0), np.float32(time_scale)))


class Oracle:
    """Plain-C restatement on flat numpy arrays."""

    def __init__(self, path=ORACLE_SO):
        self.lib = C.CDLL(path)
        L = self.lib
        L.orc_substep_plan.restype = C.c_int
        L.orc_substep_plan.argtypes = [C.c_float, C.c_float, C.c_int, _fp]
        L.orc_compute_diagnostics.restype = C.c_int
        L.orc_compute_diagnostics3.restype = C.c_int

    @staticmethod
    def _dim(pos):
        return pos.shape[1]

    def compute_gravity(self, pos, vel, mass, pinned=None, g=DEFAULTS["g"], softening=DEFAULTS["softening"], acc=None):
        n, dim = pos.shape
        pos = np.ascontiguousarray(pos, np.float64)
        vel = np.ascontiguousarray(vel, np.float64)
        mass = np.ascontiguousarray(mass, np.float32)
        pinned = None if pinned is None else np.ascontiguousarray(pinned, np.uint8)
        acc = np.zeros((n, dim)) if acc is None else np.ascontiguousarray(acc, np.float64)
        fn = self.lib.orc_compute_gravity if dim == 2 else self.lib.orc_compute_gravity3
        fn(C.c_int64(n), _d(pos), _d(vel), _f(mass), _b(pinned), C.c_double(g), C.c_float(softening), _d(acc))
        return acc

    def compute_gravity_cfg(self, pos, vel, mass, pinned, cfg, acc=None):
        """compute_gravity with the solver switch of Physics.hpp:212 (cfg.bh_threshold / cfg.bh_theta), 2-D."""
        n = pos.shape[0]
        out = np.zeros((n, 2), np.float64) if acc is None else acc
        self.lib.orc_compute_gravity_cfg(C.c_int64(n), _d(pos), _d(vel), _f(mass), _b(pinned), C.byref(cfg), _d(out))
        return out

    def bh_gravity(self, pos, vel, mass, pinned=None, g=DEFAULTS["g"], softening=DEFAULTS["softening"],
                   theta=REF_BH_THETA, acc=None, with_stats=False):
        """Barnes-Hut branch of compute_gravity (Physics.hpp:212-230), 2-D. Raises if two gathered bodies
        share their float coordinates (the reference does not terminate on such input)."""
        n = pos.shape[0]
        out = np.zeros((n, 2), np.float64) if acc is None else acc
        stats = np.zeros(4, np.int64)
        self.lib.orc_bh_gravity.restype = C.c_int64
        rc = self.lib.orc_bh_gravity(C.c_int64(n), _d(pos), _d(vel), _f(mass), _b(pinned), C.c_double(g),
                                     C.c_float(softening), C.c_float(theta), _d(out),
                                     stats.ctypes.data_as(C.POINTER(C.c_int64)))
        if rc < 0:
            raise ValueError("coincident float positions: the reference's quadtree does not terminate")
        return (out, dict(nodes=int(stats[0]), depth=int(stats[1]), popped=int(stats[2]), terms=int(stats[3]))) \
            if with_stats else out

    def substep_plan(self, dt, max_substep=DEFAULTS["max_substep"], max_substeps=DEFAULTS["max_substeps_per_frame"]):
        out = C.c_float()
        n = self.lib.orc_substep_plan(C.c_float(dt), C.c_float(max_substep), C.c_int(max_substeps), C.byref(out))
        return n, out.value

    def integrate(self, pos, vel, acc, acc_prev, mass, pinned, cfg, dt):
        """In place on float64 C-contiguous arrays."""
        n, dim = pos.shape
        fn = self.lib.orc_integrate if dim == 2 else self.lib.orc_integrate3
        fn(C.c_int64(n), _d(pos), _d(vel), _d(acc), _d(acc_prev), _f(mass), _b(pinned), C.byref(cfg), C.c_float(dt))

    def physics_frames(self, pos, vel, acc, acc_prev, mass, pinned, cfg, dt, n_frames=1):
        n, dim = pos.shape
        fn = self.lib.orc_physics_frames if dim == 2 else self.lib.orc_physics_frames3
        fn(C.c_int64(n), _d(pos), _d(vel), _d(acc), _d(acc_prev), _f(mass), _b(pinned), C.byref(cfg),
           C.c_float(dt), C.c_int64(n_frames))

    def compute_diagnostics(self, pos, vel, mass, g=DEFAULTS["g"], softening=DEFAULTS["softening"]):
        n, dim = pos.shape
        pos = np.ascontiguousarray(pos, np.float64)
        vel = np.ascontiguousarray(vel, np.float64)
        mass = np.ascontiguousarray(mass, np.float32)
        eps2 = float(np.float64(np.float32(softening)) * np.float64(np.float32(softening)))
        out = np.zeros(8 if dim == 2 else 10)
        fn = self.lib.orc_compute_diagnostics if dim == 2 else self.lib.orc_compute_diagnostics3
        ok = fn(C.c_int64(n), _d(pos), _d(vel), _f(mass), C.c_double(g), C.c_double(eps2), _d(out))
        return out, bool(ok)

    def count_overlaps(self, pos, mass, radius):
        """Detection phase of Collision::resolve (Collision.hpp:63-84): (n_pairs, first_pair or None)."""
        first = (C.c_int64 * 2)()
        self.lib.orc_count_overlaps.restype = C.c_int64
        n = self.lib.orc_count_overlaps(C.c_int64(pos.shape[0]), _d(np.ascontiguousarray(pos, np.float64)),
                                        _f(np.ascontiguousarray(mass, np.float32)),
                                        _d(np.ascontiguousarray(radius, np.float64)), first)
        return int(n), ((first[0], first[1]) if n else None)

    def radius_from_mass(self, mass, density=None):
        """Collision::radius_of for a body without a Radius component (Collision.hpp:30-34)."""
        n = len(mass)
        rho = np.full(n, BODY_DENSITY) if density is None else np.ascontiguousarray(density, np.float64)
        out = np.zeros(n)
        self.lib.orc_radius_from_mass(C.c_int64(n), _f(np.ascontiguousarray(mass, np.float32)), _d(rho), _d(out))
        return out

    def resolve_collisions(self, pos, vel, mass, pinned, radius, density=None, with_log=False):
        """Collision::resolve, merge mode (Collision.hpp:49-132). Returns (merges, alive, changed, pos, vel, mass,
        pinned, radius) with the survivors' new values; inputs are not modified."""
        n = pos.shape[0]
        p, v = np.array(pos, np.float64, order="C"), np.array(vel, np.float64, order="C")
        m, pin = np.array(mass, np.float32), np.array(pinned if pinned is not None else np.zeros(n), np.uint8)
        r = np.array(radius, np.float64)
        rho = np.full(n, BODY_DENSITY) if density is None else np.array(density, np.float64)
        alive, changed = np.ones(n, np.uint8), np.zeros(n, np.uint8)
        self.lib.orc_resolve_collisions.restype = C.c_int64
        log = np.zeros((8 * max(n, 1), 2), np.int64)
        k = self.lib.orc_resolve_collisions(C.c_int64(n), _d(p), _d(v), _f(m), _b(pin), _d(r), _d(rho), _b(alive), _b(changed),
                                            log.ctypes.data_as(C.POINTER(C.c_int64)), C.c_int64(log.shape[0]))
        if with_log:
            return int(k), alive, changed, p, v, m, pin, r, log[:int(k)]
        return int(k), alive, changed, p, v, m, pin, r

    def cbrt(self, x):
        """libm's cbrt (what std::cbrt is in the reference's build), element-wise."""
        x = np.ascontiguousarray(x, np.float64)
        out = np.zeros_like(x)
        self.lib.orc_cbrt(C.c_int64(x.size), _d(x), _d(out))
        return out

    def zero_net_momentum(self, vel, mass, pinned=None):
        n = vel.shape[0]
        self.lib.orc_zero_net_momentum(C.c_int64(n), _d(vel), _f(mass), _b(pinned))

    def seed_three_body(self, g=DEFAULTS["g"]):
        pos = np.zeros((3, 2))
        vel = np.zeros((3, 2))
        mass = np.zeros(3, np.float32)
        pinned = np.zeros(3, np.uint8)
        self.lib.orc_seed_three_body(C.c_double(g), _d(pos), _d(vel), _f(mass), _b(pinned))
        return pos, vel, mass, pinned


class Reference:
    """The unmodified reference physics (2-D), through oracle/ref_shim.cpp."""

    def __init__(self, path=REF_SO):
        self.lib = C.CDLL(path)
        L = self.lib
        L.ref_world_create.restype = C.c_void_p
        L.ref_reset_scenario.restype = C.c_int64
        L.ref_num_bodies.restype = C.c_int64
        L.ref_dt_eff.restype = C.c_float
        L.ref_compute_diagnostics.restype = C.c_int
        L.ref_read_frame_diagnostics.restype = C.c_int
        L.ref_paused.restype = C.c_int
        self.h = C.c_void_p(L.ref_world_create())

    def close(self):
        if self.h:
            self.lib.ref_world_destroy(self.h)
            self.h = None

    def set_config(self, **kw):
        d = dict(DEFAULTS)
        d.update(kw)
        self.lib.ref_set_config(self.h, C.c_double(d["g"]), C.c_float(d["softening"]), C.c_float(d["max_speed"]),
                                C.c_int(d["integrator"]), C.c_float(d["fixed_dt"]), C.c_float(d["time_scale"]),
                                C.c_float(d["max_substep"]), C.c_int(d["max_substeps_per_frame"]))

    def set_solver(self, bh_threshold=REF_BH_THRESHOLD, bh_theta=REF_BH_THETA):
        self.lib.ref_set_solver(self.h, C.c_int(int(bh_threshold)), C.c_float(bh_theta))

    def solver_defaults(self):
        t, th = C.c_int(), C.c_float()
        self.lib.ref_get_solver_defaults(C.byref(t), C.byref(th))
        return t.value, th.value

    def defaults(self):
        g = C.c_double()
        s, ms, fd, ts, sub = C.c_float(), C.c_float(), C.c_float(), C.c_float(), C.c_float()
        it, nsub = C.c_int(), C.c_int()
        self.lib.ref_get_defaults(C.byref(g), C.byref(s), C.byref(ms), C.byref(it), C.byref(fd), C.byref(ts),
                                  C.byref(sub), C.byref(nsub))
        return dict(g=g.value, softening=s.value, max_speed=ms.value, integrator=it.value, fixed_dt=fd.value,
                    time_scale=ts.value, max_substep=sub.value, max_substeps_per_frame=nsub.value)

    def load_bodies(self, pos, vel, mass, pinned=None):
        n = pos.shape[0]
        pos = np.ascontiguousarray(pos, np.float64)
        vel = np.ascontiguousarray(vel, np.float64)
        mass = np.ascontiguousarray(mass, np.float32)
        pinned = None if pinned is None else np.ascontiguousarray(pinned, np.uint8)
        self.lib.ref_load_bodies(self.h, C.c_int64(n), _d(pos), _d(vel), _f(mass), _b(pinned))

    def reset_scenario(self):
        return self.lib.ref_reset_scenario(self.h)

    def collision_resolve(self, radius):
        """Collision::resolve with explicit Radius components; returns the number of bodies left."""
        self.lib.ref_collision_resolve.restype = C.c_int64
        return int(self.lib.ref_collision_resolve(self.h, _d(np.ascontiguousarray(radius, np.float64))))

    def collision_resolve_full(self, radius, density=None):
        """Collision::resolve with per-body results in load order. radius < 0: no Radius component (derived from
        mass and density). Returns (alive, pos, vel, acc, mass, pinned, radius_after, visit_order); visit_order[k] is
        the load index of the k-th body the pass visits (flecs iterates table by table)."""
        n = self.num_bodies()
        alive = np.zeros(n, np.uint8)
        pos, vel, acc = np.zeros((n, 2)), np.zeros((n, 2)), np.zeros((n, 2))
        mass, pinned, rout = np.zeros(n, np.float32), np.zeros(n, np.uint8), np.zeros(n)
        rin = np.ascontiguousarray(radius, np.float64)
        rho = None if density is None else np.ascontiguousarray(density, np.float64)
        self.lib.ref_collision_resolve_full.restype = C.c_int64
        order = np.zeros(n, np.int64)
        self.lib.ref_collision_resolve_full(self.h, _d(rin), _d(rho), _b(alive), _d(pos), _d(vel), _d(acc), _f(mass),
                                            _b(pinned), _d(rout), order.ctypes.data_as(C.POINTER(C.c_int64)))
        return alive, pos, vel, acc, mass, pinned, rout, order

    def num_bodies(self):
        return self.lib.ref_num_bodies(self.h)

    def read_bodies(self):
        n = self.num_bodies()
        pos, vel, acc = np.zeros((n, 2)), np.zeros((n, 2)), np.zeros((n, 2))
        mass, pinned = np.zeros(n, np.float32), np.zeros(n, np.uint8)
        self.lib.ref_read_bodies(self.h, _d(pos), _d(vel), _d(acc), _f(mass), _b(pinned))
        return pos, vel, acc, mass, pinned

    def zero_net_momentum(self):
        self.lib.ref_zero_net_momentum(self.h)

    def compute_gravity(self):
        self.lib.ref_compute_gravity(self.h)

    def integrate(self, dt):
        self.lib.ref_integrate(self.h, C.c_float(dt))

    def dt_eff(self):
        return self.lib.ref_dt_eff(self.h)

    def physics_frame(self):
        self.lib.ref_physics_frame(self.h)

    def progress(self):
        self.lib.ref_progress(self.h)

    def paused(self):
        return bool(self.lib.ref_paused(self.h))

    def compute_diagnostics(self):
        out = np.zeros(8)
        ok = self.lib.ref_compute_diagnostics(self.h, _d(out))
        return out, bool(ok)

    def frame_diagnostics(self):
        out = np.zeros(8)
        ok = self.lib.ref_read_frame_diagnostics(self.h, _d(out))
        return out, ok


def have_reference():
    return os.path.exists(REF_SO)
