"""Barnes-Hut branch on the GPU vs the unmodified reference / the oracle: bitwise, plus timings.
usage: python tests/tools/bh_check.py [n ...]   (run on a GPU box; the reference library is used when present)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge
from oracle import bindings as ob

def main():
    pkg = ge.load_package()
    ph, sc = pkg.physics, pkg.scenarios
    sizes = [int(a) for a in sys.argv[1:]] or [101, 4096, 65536, 1 << 20]
    orc = ob.Oracle()
    rng = np.random.default_rng(11)
    for n in sizes:
        pos, vel, mass = sc.plummer(n, seed=5, dim=2)
        mass = (mass * rng.uniform(0.5, 2.0, n)).astype(np.float32)
        pinned = (rng.random(n) < 0.01).astype(np.uint8)
        eps = 4.0
        t0 = time.time()
        a_orc, st = orc.bh_gravity(pos, vel, mass, pinned, softening=eps, with_stats=True)
        t_orc = time.time() - t0
        t_ref = None
        if ob.have_reference() and n <= (1 << 18):
            r = ob.Reference(); r.set_config(softening=eps); r.set_solver(100, 0.5)
            r.load_bodies(pos, vel, mass, pinned)
            t0 = time.time(); r.compute_gravity(); t_ref = time.time() - t0
            a_ref = r.read_bodies()[2]; r.close()
            assert np.array_equal(a_ref, a_orc), "oracle != reference"
        p = ph.Physics(0)
        cfg = ph.default_config(); cfg.softening = eps; p.set_config(cfg)
        p.set_solver(ph.SOLVER_REFERENCE, 100, 0.5)
        p.solver_profile(True)
        p.upload(pos, vel, mass, pinned)
        p.compute_gravity()
        a_gpu = p.download()[2]
        info = p.solver_info()
        same = np.array_equal(a_gpu, a_orc)
        bad = int((a_gpu != a_orc).any(axis=1).sum())
        p.solver_profile(False)
        for _ in range(3):
            p.compute_gravity()
        p.synchronize()
        reps = 10
        p.timer_begin()
        for _ in range(reps):
            p.compute_gravity()
        ms = p.timer_end() / reps
        print(f"n={n}: bit-identical={same} (rows differing {bad}) tree cells {info['cells']} (oracle {st['nodes']}) depth "
              f"{info['depth']} ({st['depth']}) visited {info['visited']} ({st['popped']}) terms {info['terms']} ({st['terms']}) "
              f"exact tests {info['exact_tests']} | GPU {ms:.3f} ms/pass, oracle {t_orc*1e3:.1f} ms"
              + (f", reference {t_ref*1e3:.1f} ms" if t_ref is not None else ""), flush=True)
        p.close()

if __name__ == "__main__":
    main()
