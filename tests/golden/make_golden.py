"""Generates tests/golden/*.npz|json by running the UNMODIFIED reference (oracle/_ref/libnbref.so,
built from /root/reference by oracle/Makefile). Run in the build container only:

    python tests/golden/make_golden.py

The fixtures pin the oracle (tests/test_oracle.py) and the CUDA path (tests/test_gpu_parity.py)
on machines where /root/reference does not exist.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
from oracle import bindings as ob  # noqa: E402

sc = ge.load_package().scenarios


def mixed_bodies(n=256, seed=99):
    """Random bodies exercising every filter of Physics.hpp:198-199 and the pinned rules."""
    rng = np.random.default_rng(seed)
    pos = (rng.standard_normal((n, 2)) * 1e8).astype(np.float32).astype(np.float64)
    vel = (rng.standard_normal((n, 2)) * 1e3).astype(np.float32).astype(np.float64)
    mass = rng.uniform(1e22, 1e24, n).astype(np.float32)
    pinned = (rng.uniform(size=n) < 0.06).astype(np.uint8)
    mass[5] = 0.0
    mass[17] = -3.0e22
    pos[7, 0] = np.nan
    vel[11, 1] = np.inf
    pinned[7] = 1
    mass[23] = np.float32(np.inf)
    return pos, vel, mass, pinned


def main():
    r = ob.Reference()
    with open(os.path.join(HERE, "defaults.json"), "w") as f:
        json.dump(r.defaults(), f, indent=1, sort_keys=True)

    # ---- C1: built-in three-body seed, 10k frames, both integrators --------------------
    out = {}
    for integ, name in ((1, "verlet"), (0, "euler")):
        r.set_config(integrator=integ)
        r.reset_scenario()
        p, v, a, m, pin = r.read_bodies()
        out["seed_pos"], out["seed_vel"], out["seed_mass"] = p, v, m
        d, ok = r.compute_diagnostics()
        out[f"{name}_diag0"] = d
        done = 0
        emax = 0.0
        for upto in (100, 1000, 10000):
            while done < upto:
                r.physics_frame()
                done += 1
                dd, _ = r.compute_diagnostics()
                emax = max(emax, abs((dd[2] - d[2]) / d[2]))
            p, v, a, m, pin = r.read_bodies()
            out[f"{name}_pos_{upto}"], out[f"{name}_vel_{upto}"], out[f"{name}_acc_{upto}"] = p, v, a
            out[f"{name}_diag_{upto}"] = r.compute_diagnostics()[0]
            out[f"{name}_max_drift_{upto}"] = np.array(emax)
    out["dt_eff"] = np.array(r.dt_eff(), dtype=np.float32)
    # the same 10 frames through the reference's own flecs pipeline (collision + diagnostics +
    # gravity + integrate): identical to physics_frame when no body can merge
    r.set_config(integrator=1)
    r.reset_scenario()
    for _ in range(10):
        r.progress()
    p, v, a, m, pin = r.read_bodies()
    out["progress_pos_10"], out["progress_vel_10"] = p, v
    out["progress_diag_10"] = r.frame_diagnostics()[0]
    np.savez_compressed(os.path.join(HERE, "c1_three_body.npz"), **out)

    # ---- C2: 4096-body uniform disk, reference accelerations ----------------------------
    pos, vel, mass = sc.uniform_disk(4096, seed=1234)
    r.set_config(integrator=0)
    r.load_bodies(pos, vel, mass)
    r.compute_gravity()
    _, _, acc, _, _ = r.read_bodies()
    d, ok = r.compute_diagnostics()
    np.savez_compressed(os.path.join(HERE, "c2_disk_4096.npz"), pos=pos, mass=mass, acc=acc, diag=d)

    # ---- mixed: pins, non-finite, m <= 0, speed cap, substeps, both integrators ----------
    pos, vel, mass, pinned = mixed_bodies()
    out = dict(pos=pos, vel=vel, mass=mass, pinned=pinned)
    knobs = dict(max_substep=3000.0, max_speed=2000.0)
    for integ, name in ((1, "verlet"), (0, "euler")):
        r.set_config(integrator=integ, **knobs)
        r.load_bodies(pos, vel, mass, pinned)
        r.compute_gravity()
        out[f"{name}_acc0"] = r.read_bodies()[2]
        for _ in range(3):
            r.physics_frame()
        p, v, a, _, _ = r.read_bodies()
        out[f"{name}_pos"], out[f"{name}_vel"], out[f"{name}_acc"] = p, v, a
        d, ok = r.compute_diagnostics()
        out[f"{name}_diag_ok"] = np.array(ok)
    # zero_net_momentum with pins
    r.load_bodies(pos, vel, mass, pinned)
    r.zero_net_momentum()
    out["zeroed_vel"] = r.read_bodies()[1]
    np.savez_compressed(os.path.join(HERE, "mixed_256.npz"), **out)

    # ---- diagnostics on a clean 2-D Plummer projection ---------------------------------
    pos, vel, mass = sc.plummer(1024, seed=5, dim=2)
    r.set_config()
    r.load_bodies(pos, vel, mass)
    d, ok = r.compute_diagnostics()
    r.compute_gravity()
    acc = r.read_bodies()[2]
    np.savez_compressed(os.path.join(HERE, "plummer2d_1024.npz"), pos=pos, vel=vel, mass=mass, diag=d, acc=acc)
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
