"""Developer check on a B200: per-body acceleration error of the FP32 path on EVOLVED states.

The 1e-5 bound of north_star is easy on float-representable inputs (every generator rounds its
positions to float). A drop-in sees FP64 positions that have moved: this script advances a world
K Verlet frames in the oracle (FP64, the reference's arithmetic), uploads the oracle's state,
and compares nb_compute_gravity per body with orc_compute_gravity at the reference's default
softening (4 m, Constants.hpp:89). Prints max / p99.9 / median / fraction over 1e-5, raw and
divided by the body's condition number where that is large (tests/conftest.py::conditioning).

    python tests/tools/evolved_accuracy.py [disk4096] [plummer65536] [disk3d] ...
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np

import __graft_entry__ as ge
from conftest import conditioning
from oracle import bindings as ob

pkg = ge.load_package()
sc = pkg.scenarios
o = ob.Oracle()


def report(tag, acc, ref, pos, mass, eps):
    e = np.linalg.norm(acc - ref, axis=1) / np.linalg.norm(ref, axis=1)
    line = (f"{tag}: max {e.max():.3e} p99.9 {np.quantile(e, 0.999):.3e} median {np.median(e):.3e} "
            f"frac>1e-5 {np.mean(e > 1e-5):.4f}")
    bad = np.nonzero(e > 5e-6)[0]
    if len(bad) and len(pos) <= 70000:
        sub = bad[:512]
        m64 = mass.astype(np.float64)
        cn = np.empty(len(sub))
        for q, i in enumerate(sub):
            d = pos - pos[i]
            r2 = (d * d).sum(axis=1) + eps * eps
            t = (m64 / (r2 * np.sqrt(r2)))[:, None] * d
            cn[q] = np.abs(t).sum() / max(np.linalg.norm(t.sum(axis=0)), 1e-300)
        line += f" | of {len(bad)} bodies over 5e-6: max err/cond {np.max(e[sub] / np.maximum(cn, 1.0)):.3e}, max cond {cn.max():.1f}"
    print(line, flush=True)
    return e


def evolve_in_oracle(pos, vel, mass, eps, dt, frames):
    n, dim = pos.shape
    cfg = ob.make_config(softening=eps)
    p, v = pos.copy(), vel.copy()
    a, a0 = np.zeros_like(p), np.zeros_like(p)
    o.physics_frames(p, v, a, a0, mass, np.zeros(n, np.uint8), cfg, dt, frames)
    return p, v


def case(tag, pos, vel, mass, eps, dt, frames):
    ph = pkg.Physics(0)
    ph.set_precision(pkg.PRECISION_FP32)
    ph.set_config(softening=eps)
    ph.upload(pos, vel, mass)
    ph.compute_gravity()
    acc0 = ph.download()[2]
    ref0 = o.compute_gravity(pos, vel, mass, softening=eps)
    report(f"{tag} seed state (float-representable) eps={eps:g}", acc0, ref0, pos, mass, eps)
    p, v = evolve_in_oracle(pos, vel, mass, eps, dt, frames)
    ph.upload(p, v, mass)
    ph.compute_gravity()
    acc = ph.download()[2]
    ref = o.compute_gravity(p, v, mass, softening=eps)
    e = report(f"{tag} after {frames} oracle Verlet frames of dt={dt:g} s, eps={eps:g}", acc, ref, p, mass, eps)
    ph.close()
    return e


if __name__ == "__main__":
    what = sys.argv[1:] or ["disk4096", "plummer16384"]
    for w in what:
        if w.startswith("disk3d"):
            n = int(w[6:] or 4096)
            p2, v2, m = sc.uniform_disk(n, seed=1234, circular=True)
            rng = np.random.default_rng(5)
            pos = np.concatenate([p2, (1e6 * rng.standard_normal((n, 1))).astype(np.float32).astype(np.float64)], axis=1)
            vel = np.concatenate([v2, np.zeros((n, 1))], axis=1)
            case(w, pos, vel, m, 4.0, 1000.0, 20)
        elif w.startswith("disk"):
            n = int(w[4:] or 4096)
            pos, vel, m = sc.uniform_disk(n, seed=1234, circular=True)
            case(w, pos, vel, m, 4.0, 1000.0, 20)
        elif w.startswith("plummer"):
            n = int(w[7:] or 16384)
            pos, vel, m = sc.plummer(n, seed=1234, dim=2)
            dt = 1e-3 * sc.crossing_time(n)
            case(w, pos, vel, m, 4.0, dt, 5)
