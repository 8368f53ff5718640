"""Pins oracle/bh_oracle.c — the plain-C restatement of the reference's Barnes-Hut branch
(Physics.hpp:212-230, SpatialPartition.hpp:36-220) — bit-for-bit against fixtures the unmodified
reference generated with its DEFAULT solver settings (tests/golden/make_golden_bh.py) and, where the
compiled reference exists, against the live library on fresh inputs."""
import numpy as np
import pytest

from conftest import golden

STRIDE = 16


@pytest.fixture(scope="module")
def g():
    return golden("bh_reference_solver.npz")


def test_reference_solver_defaults(g, ob):
    assert int(g["bh_threshold"]) == ob.REF_BH_THRESHOLD == 100
    assert float(g["bh_theta"]) == ob.REF_BH_THETA == 0.5


def test_disk_4096_default_solver_bitexact(g, oracle_lib, pkg):
    pos, vel, mass = pkg.scenarios.uniform_disk(4096, seed=1234)
    acc, st = oracle_lib.bh_gravity(pos, vel, mass, with_stats=True)
    assert np.array_equal(acc, g["disk4096_acc"])
    assert st["nodes"] % 4 == 1 and st["terms"] < 4096 * 4095 // 4  # a tree walk, not the all-pairs sum
    # what the approximation costs against the direct sum the north-star path evaluates
    direct = golden("c2_disk_4096.npz")["acc"]
    e = np.linalg.norm(acc - direct, axis=1) / np.linalg.norm(direct, axis=1)
    assert 1e-3 < np.median(e) < 5e-2


@pytest.mark.parametrize("theta", [0.3, 1.0, 0.0])
def test_opening_angles_bitexact(g, oracle_lib, pkg, theta):
    pos, vel, mass = pkg.scenarios.uniform_disk(2048, seed=77)
    acc, st = oracle_lib.bh_gravity(pos, vel, mass, theta=theta, with_stats=True)
    assert np.array_equal(acc, g[f"disk2048_theta{theta}_acc"])
    if theta == 0.0:
        # nothing is ever accepted: every body meets every other leaf, i.e. the direct sum with the
        # terms rounded to float and added in walk order
        assert st["terms"] == 2048 * 2047
        direct = oracle_lib.compute_gravity(pos, vel, mass)
        e = np.linalg.norm(acc - direct, axis=1) / np.linalg.norm(direct, axis=1)
        assert e.max() < 1e-4


def test_switch_at_bh_threshold_bitexact(g, oracle_lib, ob, pkg):
    pos, vel, mass = pkg.scenarios.uniform_disk(101, seed=9)
    cfg = ob.make_config(bh_threshold=100)
    for n in (100, 101):
        acc = oracle_lib.compute_gravity_cfg(pos[:n], vel[:n], mass[:n], None, cfg)
        assert np.array_equal(acc, g[f"disk{n}_acc"])
    # 100 bodies: the direct sum (FP64 terms); 101: float-rounded terms
    assert np.array_equal(g["disk100_acc"], oracle_lib.compute_gravity(pos[:100], vel[:100], mass[:100]))
    a101 = g["disk101_acc"]
    assert np.array_equal(a101, a101.astype(np.float32).astype(np.float64))


@pytest.mark.parametrize("integ,name", [(1, "verlet"), (0, "euler")])
def test_mixed_bodies_three_frames_bitexact(g, oracle_lib, ob, integ, name):
    m = golden("mixed_256.npz")
    pos, vel, mass, pinned = m["pos"].copy(), m["vel"].copy(), m["mass"], m["pinned"]
    n = pos.shape[0]
    cfg = ob.make_config(integrator=integ, max_substep=3000.0, max_speed=2000.0, bh_threshold=100)
    acc0 = oracle_lib.compute_gravity_cfg(pos, vel, mass, pinned, cfg)
    assert np.array_equal(acc0, g[f"mixed_{name}_acc0"])
    assert not acc0[5].any() and not acc0[7].any() and not acc0[pinned == 1].any()
    acc, accp = np.zeros((n, 2)), np.zeros((n, 2))
    oracle_lib.physics_frames(pos, vel, acc, accp, mass, pinned, cfg, ob.dt_eff(), 3)
    assert np.array_equal(pos, g[f"mixed_{name}_pos"], equal_nan=True)
    assert np.array_equal(vel, g[f"mixed_{name}_vel"], equal_nan=True)
    assert np.array_equal(acc, g[f"mixed_{name}_acc"], equal_nan=True)


def test_plummer_16k_three_verlet_frames_bitexact(g, oracle_lib, ob, pkg):
    pos, vel, mass = pkg.scenarios.plummer(16384, seed=5, dim=2)
    n = pos.shape[0]
    cfg = ob.make_config(integrator=1, softening=3.844e8 / 128.0, bh_threshold=100)
    dt = ob.dt_eff(1.0 / 120.0, 3.0e5)
    assert np.float32(dt) == g["plummer16k_dt_eff"]
    acc, accp = np.zeros((n, 2)), np.zeros((n, 2))
    oracle_lib.physics_frames(pos, vel, acc, accp, mass, None, cfg, dt, 3)
    assert np.array_equal(pos[::STRIDE], g["plummer16k_pos"])
    assert np.array_equal(vel[::STRIDE], g["plummer16k_vel"])
    assert np.array_equal(acc[::STRIDE], g["plummer16k_acc"])


def test_coincident_float_positions_are_reported(oracle_lib, pkg):
    pos, vel, mass = pkg.scenarios.uniform_disk(200, seed=3)
    pos[17] = pos[3] + 1e-3  # distinct doubles, identical floats
    assert np.array_equal(pos[17].astype(np.float32), pos[3].astype(np.float32))
    with pytest.raises(ValueError):
        oracle_lib.bh_gravity(pos, vel, mass)


def test_against_live_reference(oracle_lib, ob):
    if not ob.have_reference():
        pytest.skip("oracle/_ref/libnbref.so not present (no /root/reference on this machine)")
    r = ob.Reference()
    rng = np.random.default_rng(4242)
    for n, theta, thr in ((150, 0.5, 100), (3000, 0.7, 100), (3000, 0.5, 2999), (3000, 0.5, 3000)):
        pos = (rng.standard_normal((n, 2)) * 2e8 * rng.uniform(0.01, 1.0, (n, 1))).astype(np.float32).astype(np.float64)
        pos += rng.uniform(-1, 1, (n, 2))  # doubles that are NOT float-representable: the cast of :217 matters
        vel = rng.standard_normal((n, 2)) * 500.0
        mass = rng.uniform(1e21, 1e25, n).astype(np.float32)
        pinned = (rng.uniform(size=n) < 0.1).astype(np.uint8)
        mass[3] = 0.0
        vel[9, 0] = np.nan
        for integ in (0, 1):
            r.set_config(integrator=integ, softening=1.0e5, max_substep=5000.0)
            r.set_solver(thr, theta)
            r.load_bodies(pos, vel, mass, pinned)
            for _ in range(3):
                r.physics_frame()
            P, V, A, _, _ = r.read_bodies()
            p, v = pos.copy(), vel.copy()
            a, a0 = np.zeros((n, 2)), np.zeros((n, 2))
            cfg = ob.make_config(integrator=integ, softening=1.0e5, max_substep=5000.0, bh_threshold=thr, bh_theta=theta)
            oracle_lib.physics_frames(p, v, a, a0, mass, pinned, cfg, ob.dt_eff(), 3)
            assert np.array_equal(P, p, equal_nan=True) and np.array_equal(V, v, equal_nan=True)
            assert np.array_equal(A, a, equal_nan=True)
    r.close()


def test_tiny_and_degenerate_worlds_against_live_reference(oracle_lib, ob):
    if not ob.have_reference():
        pytest.skip("oracle/_ref/libnbref.so not present (no /root/reference on this machine)")
    rng = np.random.default_rng(17)
    r = ob.Reference()
    for n in (1, 2, 3, 5, 40):
        pos = (rng.standard_normal((n, 2)) * 1e8).astype(np.float32).astype(np.float64)
        vel = np.zeros((n, 2))
        mass = rng.uniform(1e22, 1e24, n).astype(np.float32)
        for pinned in (None, np.ones(n, np.uint8)):
            r.set_config(softening=0.0)
            r.set_solver(0, 0.5)
            r.load_bodies(pos, vel, mass, pinned)
            r.compute_gravity()
            assert np.array_equal(r.read_bodies()[2], oracle_lib.bh_gravity(pos, vel, mass, pinned, softening=0.0))
    # all bodies on one vertical line: the root square comes from the y extent alone; and one on a single point's
    # neighbourhood with a huge offset (float spacing >> separation in double)
    n = 64
    pos = np.zeros((n, 2)); pos[:, 1] = np.linspace(-1e8, 1e8, n)
    pos = pos.astype(np.float32).astype(np.float64)
    r.set_solver(10, 0.5)
    r.load_bodies(pos, np.zeros((n, 2)), np.full(n, 1e23, np.float32))
    r.compute_gravity()
    assert np.array_equal(r.read_bodies()[2], oracle_lib.bh_gravity(pos, np.zeros((n, 2)), np.full(n, 1e23, np.float32)))
    r.close()
