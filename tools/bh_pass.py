"""A few passes of the reference-solver (Barnes-Hut) compute_gravity, for ncu and for timing.
usage: python tools/bh_pass.py [n=1048576] [reps=5] [scenario=plummer|disk]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge

def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    kind = sys.argv[3] if len(sys.argv) > 3 else "plummer"
    pkg = ge.load_package()
    ph, sc = pkg.physics, pkg.scenarios
    pos, vel, mass = sc.plummer(n, seed=1234, dim=2) if kind == "plummer" else sc.uniform_disk(n, seed=1234)
    p = ph.Physics(0)
    p.set_config(softening=3.844e8 / np.sqrt(n))
    p.set_solver(ph.SOLVER_REFERENCE, 100, 0.5)
    p.solver_profile(True)
    p.upload(pos, vel, mass)
    p.compute_gravity()
    info = p.solver_info()
    p.solver_profile(False)
    p.compute_gravity(); p.compute_gravity()
    p.synchronize()
    p.timer_begin()
    for _ in range(reps):
        p.compute_gravity()
    ms = p.timer_end() / reps
    print(f"{kind} n={n}: {ms:.3f} ms/pass; cells {info['cells']} depth {info['depth']} visited/body "
          f"{info['visited']/n:.1f} terms/body {info['terms']/n:.1f} -> {info['visited']/ms*1e3:.3e} cells/s, "
          f"{info['terms']/ms*1e3:.3e} terms/s")
    p.close()

if __name__ == "__main__":
    main()
