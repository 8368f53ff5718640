"""The reference's own CPU physics (oracle/_ref/libnbref.so, unmodified Physics.hpp) timed on this
box's host: force pass alone, diagnostics alone, and the full world.progress() frame
(collision + diagnostics + gravity + integrate), at the sizes BASELINE.md §3 names. One core: the
reference is single-threaded. Prints a markdown table; a reported baseline, not the target."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge
from oracle import bindings as ob
pkg = ge.load_package()
sizes = [int(a) for a in sys.argv[1:]] or [4096, 16384, 65536]
assert ob.have_reference(), "oracle/_ref/libnbref.so missing"
print(f"host cores available: {os.cpu_count()} (1 used)")
print("| N | compute_gravity (s) | pairs/s | interactions/s N(N-1) | compute_diagnostics (s) | progress() Euler (s) | progress() Verlet (s) |")
print("|---|---|---|---|---|---|---|")
for n in sizes:
    pos, vel, mass = pkg.scenarios.plummer(1 << 20, seed=1234, dim=2)
    pos, vel, mass = pos[:n].copy(), vel[:n].copy(), mass[:n].copy()
    eps = 3.844e8 / np.sqrt(1 << 20)
    r = ob.Reference(); r.set_config(softening=eps, integrator=0); r.load_bodies(pos, vel, mass)
    t0 = time.perf_counter(); r.compute_gravity(); tg = time.perf_counter() - t0
    t0 = time.perf_counter(); r.compute_diagnostics(); td = time.perf_counter() - t0
    frames = {}
    for integ, name in ((0, "euler"), (1, "verlet")):
        if n > 16384 and integ == 1:
            frames[name] = float("nan"); continue
        r.set_config(softening=eps, integrator=integ)
        t0 = time.perf_counter(); r.progress(); frames[name] = time.perf_counter() - t0
    r.close()
    print(f"| {n} | {tg:.3f} | {n*(n-1)/2/tg:.3e} | {n*(n-1)/tg:.3e} | {td:.3f} | {frames['euler']:.3f} | {frames['verlet']:.3f} |", flush=True)
