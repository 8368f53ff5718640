"""Collision::resolve in merge mode (src/systems/Collision.hpp:49-132).

CPU part: oracle/nbody_oracle.c::orc_resolve_collisions against the UNMODIFIED reference (live library, build
container) and against tests/golden/collision_merge.npz (fixtures the reference generated, everywhere).
GPU part (-m gpu): nb_resolve_collisions through the C-ABI against the oracle, bit for bit: which bodies die
and every survivor's mass, velocity, position, pin and radius. The cases are chosen to hit the reference's
sequential quirks: chains (a survivor that grew reaches bodies it did not touch before), survivors that
are the LATER body of the pair (the row keeps running with a dead body's snapshot and later partners
merge into it), equal masses (>=), pinned pairs (no merge), pinned absorbers, filtered bodies."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

CASES = {  # name: (n, seed, spread, radius scale, pinned fraction, per-body densities, fraction without a Radius component)
    "sparse": (400, 1, 1e8, 3e6, 0.05, False, 0.0),
    "dense_density": (400, 2, 2e7, 3e6, 0.05, True, 0.0),
    "many_pinned": (1000, 3, 1e7, 2e6, 0.3, False, 0.0),
    "derived_radii": (1000, 4, 5e6, 2e6, 0.05, True, 1.0),
    "no_overlap": (2000, 5, 1e9, 1e6, 0.05, False, 0.0),
    "everything_overlaps": (300, 6, 1e6, 5e6, 0.05, False, 0.0),
}


def make_case(oracle_lib, ob, name):
    n, seed, spread, rad_scale, frac_pinned, with_density, no_radius_frac = CASES[name]
    rng = np.random.default_rng(seed)
    pos = (rng.standard_normal((n, 2)) * spread).astype(np.float32).astype(np.float64)
    vel = rng.standard_normal((n, 2)) * 1e3
    mass = rng.uniform(1e22, 1e24, n).astype(np.float32)
    mass[::7] = np.float32(5e23)  # equal masses: the >= of Collision.hpp:94
    pinned = (rng.random(n) < frac_pinned).astype(np.uint8)
    mass[3] = 0
    pos[5, 0] = np.nan
    vel[8, 1] = np.inf  # collisions do not filter on velocity (Collision.hpp:63-65)
    density = rng.uniform(1000, 8000, n) if with_density else np.full(n, ob.BODY_DENSITY)
    radius = rng.uniform(0.2, 1.0, n) * rad_scale
    norad = rng.random(n) < no_radius_frac
    radius_in = np.where(norad, -1.0, radius)  # for the reference shim: < 0 = no Radius component
    radius_eff = np.where(norad, oracle_lib.radius_from_mass(mass, density), radius)
    return pos, vel, mass, pinned, density, radius_in, radius_eff


def same_outcome(a, b):
    (ka, alive_a, ch_a, p_a, v_a, m_a, pin_a, r_a), (kb, alive_b, ch_b, p_b, v_b, m_b, pin_b, r_b) = a, b
    assert ka == kb and np.array_equal(alive_a, alive_b) and np.array_equal(ch_a, ch_b)
    s = (ch_a == 1)
    assert np.array_equal(p_a[s], p_b[s], equal_nan=True) and np.array_equal(v_a[s], v_b[s], equal_nan=True)
    assert np.array_equal(m_a[s], m_b[s]) and np.array_equal(pin_a[s], pin_b[s]) and np.array_equal(r_a[s], r_b[s])


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_matches_reference_fixture(oracle_lib, ob, name):
    g = np.load(os.path.join(GOLDEN, "collision_merge.npz"))
    pos, vel, mass, pinned, density, _, radius_eff = make_case(oracle_lib, ob, name)
    k, alive, changed, p, v, m, pin, r = oracle_lib.resolve_collisions(pos, vel, mass, pinned, radius_eff, density)
    assert np.array_equal(alive, g[f"{name}_alive"])
    live = alive == 1
    assert np.array_equal(p[live], g[f"{name}_pos"][live], equal_nan=True)
    assert np.array_equal(v[live], g[f"{name}_vel"][live], equal_nan=True)
    assert np.array_equal(m[live], g[f"{name}_mass"][live]) and np.array_equal(pin[live], g[f"{name}_pinned"][live])
    assert np.array_equal(r[live], g[f"{name}_radius"][live])
    assert (k == 0) == (name == "no_overlap")
    # survivors restart from rest in the reference (Acceleration = PrevAcceleration = 0, :118-119)
    assert not g[f"{name}_acc"][live & (changed == 1)].any()


@pytest.mark.parametrize("name", ["dense_density", "derived_radii"])
def test_oracle_against_live_reference_mixed_tables(oracle_lib, ob, name):
    """Bodies with and without a Radius component live in different flecs tables, so the pass visits them table
    by table: the restatement is run on the arrays in the reference's visiting order."""
    if not ob.have_reference():
        pytest.skip("oracle/_ref/libnbref.so not present (no /root/reference on this machine)")
    pos, vel, mass, pinned, density, radius_in, radius_eff = make_case(oracle_lib, ob, name)
    rng = np.random.default_rng(99)
    norad = rng.random(len(mass)) < 0.5
    radius_in = np.where(norad, -1.0, radius_eff)
    radius_eff = np.where(norad, oracle_lib.radius_from_mass(mass, density), radius_eff)
    r = ob.Reference()
    r.load_bodies(pos, vel, mass, pinned)
    alive, P, V, A, M, PIN, R, order = r.collision_resolve_full(radius_in, density)
    r.close()
    assert not np.array_equal(order, np.arange(len(order)))
    inv = np.argsort(order)
    k, al, ch, p, v, m, pin, rr = oracle_lib.resolve_collisions(pos[order], vel[order], mass[order], pinned[order],
                                                                radius_eff[order], density[order])
    al, p, v, m, pin, rr = al[inv], p[inv], v[inv], m[inv], pin[inv], rr[inv]
    live = alive == 1
    assert np.array_equal(alive, al) and k > 100
    assert np.array_equal(P[live], p[live], equal_nan=True) and np.array_equal(V[live], v[live], equal_nan=True)
    assert np.array_equal(M[live], m[live]) and np.array_equal(PIN[live], pin[live]) and np.array_equal(R[live], rr[live])


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(CASES))
def test_device_merge_pass_matches_oracle(pkg, oracle_lib, ob, name):
    pos, vel, mass, pinned, density, _, radius_eff = make_case(oracle_lib, ob, name)
    ph = pkg.Physics(0)
    ph.upload(pos, vel, mass, pinned)
    got = ph.resolve_collisions(radius_eff, density, with_log=True)
    want = oracle_lib.resolve_collisions(pos, vel, mass, pinned, radius_eff, density, with_log=True)
    same_outcome(got[:8], want[:8])
    # and the merges happen in the reference's order (an ECS caller replays its deferred destructs in it)
    assert np.array_equal(got[8], want[8])
    # the device state is untouched: the caller applies the outcome and uploads the new world
    p0, v0, _ = ph.download()
    assert np.array_equal(p0, pos, equal_nan=True) and np.array_equal(v0, vel, equal_nan=True)
    ph.close()


@pytest.mark.gpu
@pytest.mark.parametrize("n,spread", [(20000, 4e8), (60000, 1.2e9)])
def test_device_merge_pass_grid_broad_phase(pkg, oracle_lib, ob, n, spread):
    """Above 2048 bodies the initial detection is the uniform-grid broad phase, and the scan after every merge
    walks the same grid: merged bodies outgrow its cells and move off their initial positions, which the scan
    covers by widening its reach (largest current radius + largest displacement so far)."""
    rng = np.random.default_rng(n)
    pos = (rng.standard_normal((n, 2)) * spread).astype(np.float32).astype(np.float64)
    vel = rng.standard_normal((n, 2)) * 1e3
    mass = rng.uniform(1e22, 1e24, n).astype(np.float32)
    pinned = (rng.random(n) < 0.01).astype(np.uint8)
    density = np.full(n, ob.BODY_DENSITY)
    radius = oracle_lib.radius_from_mass(mass, density) * 1.5
    ph = pkg.Physics(0)
    ph.upload(pos, vel, mass, pinned)
    got = ph.resolve_collisions(radius, density)
    assert got[0] > 20
    same_outcome(got, oracle_lib.resolve_collisions(pos, vel, mass, pinned, radius, density))
    ph.close()


@pytest.mark.gpu
def test_device_merge_pass_random_small_worlds(pkg, oracle_lib, ob):
    from test_oracle_random import world
    rng = np.random.default_rng(4711)
    ph = pkg.Physics(0)
    for trial in range(120):
        n = int(rng.integers(2, 60))
        pos, vel, mass, pinned = world(rng, n)
        density = np.full(n, ob.BODY_DENSITY) if rng.random() < 0.5 else rng.uniform(0.5, 6000.0, n)
        radius = oracle_lib.radius_from_mass(mass, density) * float(rng.choice([0.3, 1.0, 3.0]))
        ph.upload(pos, vel, mass, pinned)
        got = ph.resolve_collisions(radius, density, with_log=True)
        want = oracle_lib.resolve_collisions(pos, vel, mass, pinned, radius, density, with_log=True)
        same_outcome(got[:8], want[:8])
        assert np.array_equal(got[8], want[8]), f"trial {trial}: merge order"
    ph.close()


@pytest.mark.gpu
def test_device_cbrt_is_glibc_bit_for_bit(pkg, oracle_lib):
    """update_radius_from_mass (Collision.hpp:37-46) calls std::cbrt; glibc's is not correctly rounded, so the
    device replay restates its algorithm operation by operation (merge.cuh::cbrt_glibc). Dense check against
    this machine's libm: the merge code's argument range (3 m / (4 pi rho)) and the whole exponent range."""
    rng = np.random.default_rng(99)
    x = np.concatenate([
        rng.uniform(1e-3, 1e30, 400000),
        np.exp(rng.uniform(np.log(1e-300), np.log(1e300), 400000)),
        3.0 * np.maximum(1.0, rng.uniform(1e20, 1e26, 200000).astype(np.float32).astype(np.float64)) / (4.0 * np.pi * 5510.0),
        np.array([0.0, 1.0, 8.0, 27.0, 5e-324, 2.2250738585072014e-308, 1.7976931348623157e308, np.inf]),
    ])
    ph = pkg.Physics(0)
    got = ph.debug_cbrt(x)
    ph.close()
    assert np.array_equal(got, oracle_lib.cbrt(x))   # (numpy's own cbrt is a different algorithm)
