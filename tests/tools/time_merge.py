"""Wall time of nb_resolve_collisions (Collision::resolve, merge mode) against the reference / the oracle.
usage: python tests/tools/time_merge.py   (GPU box)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge
from oracle import bindings as ob

def main():
    pkg = ge.load_package()
    o = ob.Oracle()
    # (merge-heavy frames: a Plummer core of Earth-Moon-sized bodies is far denser than their radii. At 2^20 such a
    #  frame merges 90 % of the bodies and takes tens of seconds - pass "big" to include it)
    cases = [(20000, 1.5, True), (65536, 1.5, True)] + ([(1 << 20, 1.0, False)] if "big" in sys.argv else [])
    for n, scale, cpu in cases:
        pos, vel, mass = pkg.scenarios.plummer(n, seed=1234, dim=2)
        rng = np.random.default_rng(n)
        mass = (mass * rng.uniform(0.5, 2.0, n)).astype(np.float32)
        density = np.full(n, ob.BODY_DENSITY)
        radius = o.radius_from_mass(mass, density) * scale
        ph = pkg.Physics(0)
        ph.upload(pos, vel, mass)
        ph.resolve_collisions(radius, density)  # warm-up (allocations)
        t0 = time.perf_counter()
        k, alive, changed, *_ = ph.resolve_collisions(radius, density)
        t_gpu = time.perf_counter() - t0
        info = ph.merge_info()
        line = f"n={n} radius x{scale}: {info['initial_pairs']} pairs overlap initially, {info['events']} events re-tested, {info['scanned']} bodies scanned; {k} merges, {int((alive == 0).sum())} bodies destructed; device pass {t_gpu*1e3:.2f} ms"
        if cpu:
            t0 = time.perf_counter()
            if ob.have_reference():
                r = ob.Reference(); r.load_bodies(pos, vel, mass)
                t0 = time.perf_counter()
                r.collision_resolve(radius)
                t_cpu = time.perf_counter() - t0
                r.close()
                line += f"; reference {t_cpu:.2f} s (x{t_cpu/t_gpu:.0f})"
            else:
                o.resolve_collisions(pos, vel, mass, None, radius, density)
                line += f"; oracle {time.perf_counter()-t0:.2f} s"
        print(line, flush=True)
        ph.close()

if __name__ == "__main__":
    main()
