"""Synthetic scenarios of BASELINE.json's configs (host-side numpy generators).

Every generator returns (pos, vel, mass) with positions and masses rounded to
float-representable values, so the FP64 reference and the FP32 j-stream of the CUDA path
see identical inputs. Units are SI like the reference's seed
(src/core/Constants.hpp:47-51).
"""
import numpy as np

G_DEFAULT = 6.67430e-11          # Constants.hpp:87
MOON_MASS = 7.342e22             # Constants.hpp:47
EARTH_MOON = 3.844e8             # Constants.hpp:51


def _f32(a):
    return np.asarray(a, np.float64).astype(np.float32).astype(np.float64)


def uniform_disk(n, seed=1234, radius=EARTH_MOON, mass=MOON_MASS, circular=False, g=G_DEFAULT):
    """Config 2: uniform surface density disk, equal masses, z = 0, 2-D arrays."""
    rng = np.random.default_rng(seed)
    r = radius * np.sqrt(rng.uniform(size=n))
    th = 2.0 * np.pi * rng.uniform(size=n)
    pos = _f32(np.stack([r * np.cos(th), r * np.sin(th)], axis=1))
    m = np.full(n, mass, np.float32)
    vel = np.zeros((n, 2))
    if circular:
        m_in = mass * n * (r / radius) ** 2
        vc = np.sqrt(g * m_in / np.maximum(r, 1e-9 * radius))
        vel = _f32(np.stack([-vc * np.sin(th), vc * np.cos(th)], axis=1))
    return pos, vel, m


def plummer(n, seed=1234, scale=EARTH_MOON, total_mass=None, g=G_DEFAULT, dim=3, rmax=10.0):
    """Configs 3/4: equal-mass Plummer sphere (Aarseth-Henon-Wielen sampling), truncated at
    rmax scale radii, virial velocities. dim=2 projects onto z = 0 for runs through the 2-D
    reference API."""
    rng = np.random.default_rng(seed)
    if total_mass is None:
        total_mass = MOON_MASS * n
    r = np.empty(n)
    filled = 0
    while filled < n:
        u = rng.uniform(size=n - filled)
        rr = 1.0 / np.sqrt(np.maximum(u, 1e-12) ** (-2.0 / 3.0) - 1.0)
        rr = rr[rr < rmax]
        r[filled:filled + rr.size] = rr
        filled += rr.size

    def iso(k):
        z = rng.uniform(-1.0, 1.0, k)
        ph = rng.uniform(0.0, 2.0 * np.pi, k)
        s = np.sqrt(1.0 - z * z)
        return np.stack([s * np.cos(ph), s * np.sin(ph), z], axis=1)

    pos = r[:, None] * iso(n) * scale
    # speed: q in [0,1] with density q^2 (1-q^2)^3.5 by rejection
    q = np.empty(n)
    filled = 0
    while filled < n:
        x = rng.uniform(size=2 * (n - filled))
        y = rng.uniform(0.0, 0.1, size=x.size)
        acc = x[y < x * x * (1.0 - x * x) ** 3.5][: n - filled]
        q[filled:filled + acc.size] = acc
        filled += acc.size
    vesc = np.sqrt(2.0 * g * total_mass / scale) * (1.0 + r * r) ** -0.25
    vel = (q * vesc)[:, None] * iso(n)
    mass = np.full(n, total_mass / n, np.float32)
    if dim == 2:
        pos, vel = pos[:, :2], vel[:, :2]
    pos = _f32(pos)
    vel = _f32(vel)
    return np.ascontiguousarray(pos), np.ascontiguousarray(vel), mass


def two_galaxies(n, seed=1234, scale=EARTH_MOON, g=G_DEFAULT, dim=3):
    """Config 5: two Plummer spheres on a bound collision orbit."""
    n1 = n // 2
    p1, v1, m1 = plummer(n1, seed, scale, g=g, dim=3)
    p2, v2, m2 = plummer(n - n1, seed + 1, scale, g=g, dim=3)
    sep = 8.0 * scale
    mt = float(m1.sum(dtype=np.float64) + m2.sum(dtype=np.float64))
    vrel = 0.5 * np.sqrt(2.0 * g * mt / sep)
    p1 = p1 + np.array([-sep / 2, -scale, 0.0])
    p2 = p2 + np.array([sep / 2, scale, 0.0])
    v1 = v1 + np.array([vrel / 2, 0.0, 0.0])
    v2 = v2 - np.array([vrel / 2, 0.0, 0.0])
    pos = _f32(np.concatenate([p1, p2]))
    vel = _f32(np.concatenate([v1, v2]))
    mass = np.concatenate([m1, m2])
    if dim == 2:
        pos, vel = np.ascontiguousarray(pos[:, :2]), np.ascontiguousarray(vel[:, :2])
    return pos, vel, mass


def crossing_time(n, scale=EARTH_MOON, mass=MOON_MASS, g=G_DEFAULT):
    """sqrt(a^3 / (G M)) of a Plummer sphere of n bodies: the natural time unit for dt."""
    return float(np.sqrt(scale ** 3 / (g * mass * n)))


# ---- scenario snapshots --------------------------------------------------------------------------
# The reference keeps scenarios in memory only (src/core/Scenario.hpp:14-50: bodies + a Config
# subset, nothing on disk). Same content here, as one .npz, so a device run can be checkpointed
# (Physics.download_state) and resumed (Physics.upload_state) bit for bit.
SNAPSHOT_CONFIG_KEYS = ("g", "softening", "max_speed", "integrator", "max_substep", "max_substeps_per_frame")


def save_snapshot(path, pos, vel, mass, pinned=None, acc=None, acc_prev=None, config=None, frame=0, stream_order=None):
    """Bodies (Scenario::BodyData, Scenario.hpp:14-21) plus the physics Config fields. stream_order:
    Physics.stream_order() of the run, for a resume that must be bit-identical (it fixes the summation order).
    A path ending in ".nbs" is written in the C-ABI's own scenario format (nb_scenario_write), the one the
    reference-side binding reads and writes too; anything else is a numpy .npz with the same content."""
    if str(path).endswith(".nbs"):
        from . import physics
        for k in (config or {}):
            if k not in SNAPSHOT_CONFIG_KEYS:
                raise KeyError(f"not a physics Config field: {k}")
        physics.scenario_write(path, pos, vel, mass, pinned, acc, acc_prev, config, stream_order, frame)
        return
    n = len(mass)
    data = {"pos": np.asarray(pos, np.float64), "vel": np.asarray(vel, np.float64),
            "mass": np.asarray(mass, np.float32),
            "pinned": np.zeros(n, np.uint8) if pinned is None else np.asarray(pinned, np.uint8),
            "frame": np.int64(frame)}
    if acc is not None:
        data["acc"] = np.asarray(acc, np.float64)
    if acc_prev is not None:
        data["acc_prev"] = np.asarray(acc_prev, np.float64)
    if stream_order is not None:
        data["order"] = np.asarray(stream_order[0], np.int32)
        data["order_age"] = np.int64(stream_order[1])
    for k, v in (config or {}).items():
        if k not in SNAPSHOT_CONFIG_KEYS:
            raise KeyError(f"not a physics Config field: {k}")
        data["cfg_" + k] = np.float64(v)
    np.savez_compressed(path, **data)


def load_snapshot(path):
    """-> dict(pos, vel, mass, pinned, acc|None, acc_prev|None, stream_order|None, config, frame); floats round-trip exactly."""
    if str(path).endswith(".nbs"):
        from . import physics
        d = physics.scenario_read(path)
        d["config"] = {k: (int(v) if k in ("integrator", "max_substeps_per_frame") else float(v))
                       for k, v in d["config"].items() if k in SNAPSHOT_CONFIG_KEYS}
        return d
    z = np.load(path)
    cfg = {k: (int(z["cfg_" + k]) if k in ("integrator", "max_substeps_per_frame") else float(z["cfg_" + k]))
           for k in SNAPSHOT_CONFIG_KEYS if "cfg_" + k in z.files}
    return {"pos": z["pos"], "vel": z["vel"], "mass": z["mass"], "pinned": z["pinned"],
            "acc": z["acc"] if "acc" in z.files else None, "acc_prev": z["acc_prev"] if "acc_prev" in z.files else None,
            "stream_order": (z["order"], int(z["order_age"])) if "order" in z.files else None,
            "config": cfg, "frame": int(z["frame"])}
