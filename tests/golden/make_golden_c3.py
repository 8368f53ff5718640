"""Config 3 golden record from the UNMODIFIED reference (oracle/_ref/libnbref.so): the 65 536-body Plummer
sphere (projected to the reference's 2-D), momentum zeroed, three Velocity Verlet frames
(Physics.hpp:45-57: gravity system + integrate) with Physics::compute_diagnostics (Physics.hpp:117-177)
before the first and after every frame. Committed: the four diagnostics rows, dt, softening and every 64th
body's final position/velocity  ->  tests/golden/c3_plummer2d_65536_3frames.npz
About 1-2 minutes on one core.   Run:  python tests/golden/make_golden_c3.py"""
import os, sys, time
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge
from oracle import bindings as ob


def main():
    pkg = ge.load_package()
    n, frames = 65536, 3
    pos, vel, mass = pkg.scenarios.plummer(n, seed=1234, dim=2)
    eps = float(np.float32(3.844e8 / np.sqrt(n)))
    dt = float(np.float32(1e-3 * pkg.scenarios.crossing_time(n)))
    r = ob.Reference()
    r.set_config(softening=eps, integrator=1, fixed_dt=dt, time_scale=1.0)
    r.load_bodies(pos, vel, mass)
    r.zero_net_momentum()
    assert abs(r.dt_eff() - dt) == 0.0
    t0 = time.time()
    diags = [r.compute_diagnostics()[0]]
    for _ in range(frames):
        r.physics_frame()
        d, ok = r.compute_diagnostics()
        assert ok
        diags.append(d)
    p, v, a = r.read_bodies()[:3]
    r.close()
    idx = np.arange(0, n, 64)
    diags = np.array(diags)
    np.savez_compressed(os.path.join(HERE, "c3_plummer2d_65536_3frames.npz"), diag=diags, dt=dt, softening=eps, n=n,
                        frames=frames, idx=idx, pos=p[idx], vel=v[idx], acc=a[idx], seconds=time.time() - t0)
    e = diags[:, 2]
    print(f"reference: E = {e}, drift per frame {(e[1:] - e[0]) / abs(e[0])}, {time.time() - t0:.1f} s")


if __name__ == "__main__":
    main()
