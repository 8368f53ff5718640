"""The reference's solver switch on the device (NB_SOLVER_REFERENCE, raylib-nbody_b200/csrc/bh.cuh):
above bh_threshold bodies compute_gravity is the reference's float Barnes-Hut quadtree
(Physics.hpp:212-230, SpatialPartition.hpp:36-220). The bar is BIT-IDENTICAL accelerations and
trajectories: against fixtures the unmodified reference generated with its default solver settings
(tests/golden/bh_reference_solver.npz), against the oracle (oracle/bh_oracle.c, pinned to the
reference by tests/test_oracle_bh.py) on fresh inputs up to N = 2^20, and the tree's shape and the
walks' visit counts against the oracle's."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu
STRIDE = 16


@pytest.fixture(scope="module")
def g():
    return golden("bh_reference_solver.npz")


@pytest.fixture()
def ph(pkg):
    p = pkg.Physics(0)
    p.set_solver(pkg.SOLVER_REFERENCE, 100, 0.5)
    yield p
    p.close()


def gravity(ph, pos, vel, mass, pinned=None):
    ph.upload(pos, vel, mass, pinned)
    ph.compute_gravity()
    return ph.download()[2]


def test_disk_4096_default_solver_bitexact(g, pkg, ph):
    pos, vel, mass = pkg.scenarios.uniform_disk(4096, seed=1234)
    assert np.array_equal(gravity(ph, pos, vel, mass), g["disk4096_acc"])
    info = ph.solver_info()
    assert info["passes"] == 1 and info["bodies"] == 4096 and info["cells"] % 4 == 1


@pytest.mark.parametrize("theta", [0.3, 1.0, 0.0])
def test_opening_angles_bitexact(g, pkg, ph, theta):
    pos, vel, mass = pkg.scenarios.uniform_disk(2048, seed=77)
    ph.set_solver(pkg.SOLVER_REFERENCE, 100, theta)
    ph.solver_profile(True)
    assert np.array_equal(gravity(ph, pos, vel, mass), g[f"disk2048_theta{theta}_acc"])
    if theta == 0.0:
        assert ph.solver_info()["terms"] == 2048 * 2047


def test_switch_at_bh_threshold_bitexact(g, pkg, ph):
    pos, vel, mass = pkg.scenarios.uniform_disk(101, seed=9)
    for n in (100, 101):
        assert np.array_equal(gravity(ph, pos[:n], vel[:n], mass[:n]), g[f"disk{n}_acc"])
    assert ph.solver_info()["passes"] == 1  # only the 101-body world took the tree
    # the default solver of this library ignores the threshold: always the direct sum
    ph.set_solver(pkg.SOLVER_DIRECT, 100, 0.5)
    a = gravity(ph, pos, vel, mass)
    e = np.linalg.norm(a - g["disk101_acc"], axis=1) / np.linalg.norm(a, axis=1)
    assert not np.array_equal(a, g["disk101_acc"]) and np.median(e) < 0.1


@pytest.mark.parametrize("integ,name", [(1, "verlet"), (0, "euler")])
def test_mixed_bodies_three_frames_bitexact(g, pkg, ph, ob, integ, name):
    m = golden("mixed_256.npz")
    pos, vel, mass, pinned = m["pos"], m["vel"], m["mass"], m["pinned"]
    ph.set_config(integrator=integ, max_substep=3000.0, max_speed=2000.0)
    acc0 = gravity(ph, pos, vel, mass, pinned)
    assert np.array_equal(acc0, g[f"mixed_{name}_acc0"])
    ph.upload(pos, vel, mass, pinned)
    ph.step(ob.dt_eff(), 3)
    p, v, a = ph.download()
    assert np.array_equal(p, g[f"mixed_{name}_pos"], equal_nan=True)
    assert np.array_equal(v, g[f"mixed_{name}_vel"], equal_nan=True)
    assert np.array_equal(a, g[f"mixed_{name}_acc"], equal_nan=True)
    # frame by frame through the two reference statics gives the same state
    ph.upload(pos, vel, mass, pinned)
    for _ in range(3):
        ph.compute_gravity()
        ph.integrate(ob.dt_eff())
    p2, v2, a2 = ph.download()
    assert np.array_equal(p, p2, equal_nan=True) and np.array_equal(v, v2, equal_nan=True)
    assert np.array_equal(a, a2, equal_nan=True)


def test_plummer_16k_three_verlet_frames_bitexact(g, pkg, ph, ob):
    pos, vel, mass = pkg.scenarios.plummer(16384, seed=5, dim=2)
    ph.set_config(integrator=1, softening=3.844e8 / 128.0)
    ph.upload(pos, vel, mass)
    ph.step(ob.dt_eff(1.0 / 120.0, 3.0e5), 3)
    p, v, a = ph.download()
    assert np.array_equal(p[::STRIDE], g["plummer16k_pos"])
    assert np.array_equal(v[::STRIDE], g["plummer16k_vel"])
    assert np.array_equal(a[::STRIDE], g["plummer16k_acc"])


@pytest.mark.parametrize("n", [1000, 65536, 1 << 18])
def test_fresh_inputs_against_oracle_bitexact(pkg, ph, oracle_lib, n):
    rng = np.random.default_rng(n)
    pos, vel, mass = pkg.scenarios.plummer(n, seed=21, dim=2)
    pos = pos + rng.uniform(-3.0, 3.0, pos.shape)  # not float-representable: the cast of Physics.hpp:217 matters
    mass = (mass * rng.uniform(0.25, 4.0, n)).astype(np.float32)
    pinned = (rng.random(n) < 0.02).astype(np.uint8)
    mass[5] = 0.0
    vel[11, 1] = np.inf
    ph.solver_profile(True)
    acc = gravity(ph, pos, vel, mass, pinned)
    ref, st = oracle_lib.bh_gravity(pos, vel, mass, pinned, with_stats=True)
    assert np.array_equal(acc, ref)
    assert not acc[5].any() and not acc[11].any() and not acc[pinned == 1].any()
    info = ph.solver_info()
    assert (info["cells"], info["depth"], info["bodies"]) == (st["nodes"], st["depth"], n - 2)
    assert (info["visited"], info["terms"]) == (st["popped"], st["terms"])
    # the FP32 accept test is a shortcut, never a decision of its own: a handful of FP64 re-evaluations at most
    assert info["exact_tests"] < 1e-4 * info["visited"] + 10


def test_gathered_body_count_decides_the_switch(pkg, ph, oracle_lib):
    """Physics.hpp:207,212 compare the number of bodies that PASSED the filter with bh_threshold."""
    pos, vel, mass = pkg.scenarios.uniform_disk(600, seed=4)
    mass[:50] = 0.0  # 550 gathered bodies
    ph.set_precision(pkg.PRECISION_FP64_EXACT)
    for thr, tree in ((549, True), (550, False)):
        ph.set_solver(pkg.SOLVER_REFERENCE, thr, 0.5)
        acc = gravity(ph, pos, vel, mass)
        direct = oracle_lib.compute_gravity(pos, vel, mass)
        assert np.array_equal(acc, oracle_lib.bh_gravity(pos, vel, mass) if tree else direct)
        assert np.array_equal(acc, direct) != tree


def test_clustered_pairs_grow_the_tree_arena(pkg, ph, oracle_lib):
    """Tight pairs make long single-child chains: far more cells than the first allocation holds."""
    rng = np.random.default_rng(8)
    n = 8192
    base = (rng.uniform(-1, 1, (n // 2, 2)) * 3.0e8).astype(np.float32).astype(np.float64)
    twin = base * (1.0 + 2.0 ** -19)
    pos = np.concatenate([base, twin]).astype(np.float32).astype(np.float64)
    vel = np.zeros_like(pos)
    mass = np.full(n, 7.342e22, np.float32)
    acc = gravity(ph, pos, vel, mass)
    ref, st = oracle_lib.bh_gravity(pos, vel, mass, with_stats=True)
    assert st["nodes"] > 6 * n + 4096, "input no longer forces the arena to grow"
    assert np.array_equal(acc, ref) and ph.solver_info()["cells"] == st["nodes"]
    # and the grown arena is reused by the next pass
    assert np.array_equal(gravity(ph, pos, vel, mass), ref)


def test_coincident_float_positions_fail_loudly(pkg, ph):
    pos, vel, mass = pkg.scenarios.uniform_disk(200, seed=3)
    pos[17] = pos[3] + 1e-3
    ph.upload(pos, vel, mass)
    with pytest.raises(pkg.NbodyError, match="float coordinates"):
        ph.compute_gravity()
    # the context stays usable
    pos[17] = pos[3] + 1.0e5
    assert np.isfinite(gravity(ph, pos, vel, mass)).all()


def test_three_d_upload_is_rejected(pkg, ph):
    pos, vel, mass = pkg.scenarios.plummer(300, seed=2, dim=3)
    ph.upload(pos, vel, mass)
    with pytest.raises(pkg.NbodyError, match="2-D"):
        ph.compute_gravity()


def test_full_size_2e20_bitexact(pkg, ph, oracle_lib):
    """Config 4's body count through the reference's default solver: every one of the 2^20 accelerations."""
    n = 1 << 20
    pos, vel, mass = pkg.scenarios.plummer(n, seed=1234, dim=2)
    ph.set_config(softening=3.844e8 / np.sqrt(n))
    acc = gravity(ph, pos, vel, mass)
    ref = oracle_lib.bh_gravity(pos, vel, mass, softening=np.float32(3.844e8 / np.sqrt(n)))
    assert np.array_equal(acc, ref)
    # 50 Verlet frames stay finite and conserve momentum to float-sum accuracy (the tree is not symmetric)
    ph.upload(pos, vel, mass)
    ph.step(50.0, 5)
    p, v, a = ph.download()
    assert np.isfinite(p).all() and np.isfinite(v).all() and ph.solver_info()["passes"] == 1 + 6


@pytest.mark.parametrize("n", [1, 2, 3, 5])
def test_tiny_worlds_on_the_tree_side(pkg, ph, oracle_lib, n):
    """bh_threshold = 0 sends even a single body through the tree (root leaf, no partner)."""
    rng = np.random.default_rng(n)
    pos = (rng.standard_normal((n, 2)) * 1e8).astype(np.float32).astype(np.float64)
    vel = np.zeros((n, 2))
    mass = rng.uniform(1e22, 1e24, n).astype(np.float32)
    ph.set_solver(pkg.SOLVER_REFERENCE, 0, 0.5)
    ph.set_config(softening=0.0)  # epsilon = 0 is allowed (Physics.hpp:186)
    acc = gravity(ph, pos, vel, mass)
    assert np.array_equal(acc, oracle_lib.bh_gravity(pos, vel, mass, softening=0.0))
    assert ph.solver_info()["bodies"] == n and (n > 1 or not acc.any())


def test_all_pinned_and_all_filtered(pkg, ph, oracle_lib):
    pos, vel, mass = pkg.scenarios.uniform_disk(300, seed=12)
    pinned = np.ones(300, np.uint8)
    acc = gravity(ph, pos, vel, mass, pinned)
    assert not acc.any() and ph.solver_info()["passes"] == 1  # the tree is built, nobody walks it (Physics.hpp:225)
    # nothing passes the gather filter: compute_gravity returns before touching anything (Physics.hpp:208)
    ph.upload(pos, vel, np.zeros(300, np.float32))
    ph.compute_gravity()
    assert not ph.download()[2].any()


def test_direct_solver_is_the_default_and_unchanged_by_the_threshold(pkg, oracle_lib):
    pos, vel, mass = pkg.scenarios.uniform_disk(400, seed=5)
    p = pkg.Physics(0)
    assert p.get_solver()[0] == pkg.SOLVER_DIRECT
    p.upload(pos, vel, mass)
    p.compute_gravity()
    assert np.array_equal(p.download()[2], oracle_lib.compute_gravity(pos, vel, mass))  # 400 <= 512: FP64-exact direct sum
    p.close()


def test_random_small_worlds_bitexact(pkg, ph, oracle_lib):
    """Many small clustered worlds (ties, pins, filtered bodies, theta from 0 to 2.5, epsilon 0): where walk order
    and float accumulation matter most. Same generator as tests/test_oracle_random.py, which pins the oracle to
    the live reference on these worlds."""
    from test_oracle_random import world
    rng = np.random.default_rng(20261104)
    done = 0
    for trial in range(120):
        n = int(rng.integers(1, 80))
        pos, vel, mass, pinned = world(rng, n)
        if len({(np.float32(p[0]), np.float32(p[1])) for p in pos}) < n:
            continue
        theta = float(rng.choice([0.0, 0.3, 0.5, 0.5, 1.0, 2.5]))
        eps = float(rng.choice([0.0, 4.0, 1e6]))
        ph.set_config(softening=eps)
        ph.set_solver(pkg.SOLVER_REFERENCE, 0, theta)
        got = gravity(ph, pos, vel, mass, pinned)
        want = oracle_lib.bh_gravity(pos, vel, mass, pinned, softening=eps, theta=theta)
        live = np.isfinite(pos).all(axis=1) & np.isfinite(vel).all(axis=1) & (mass > 0)
        assert np.array_equal(got[live], want[live]), f"trial {trial}: n={n} theta={theta} eps={eps}"
        done += 1
    assert done > 100
