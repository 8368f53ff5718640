"""C-ABI surface checks that need no GPU: the library loads, exports every symbol the header
declares, mirrors the reference's config surface, and fails loudly (never falls back) without
a CUDA device."""
import ctypes as C
import json
import os
import re

import numpy as np
import pytest

from conftest import GOLDEN, ROOT, _have_gpu


def header_functions():
    src = open(os.path.join(ROOT, "include", "nbody_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(pkg):
    lib = pkg.load_library()
    names = header_functions()
    assert len(names) >= 30
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/nbody_b200.h but not exported"
    assert sorted(pkg.EXPORTS) == names
    assert lib.nb_abi_version() == 1


def test_default_config_matches_reference_defaults(pkg):
    with open(os.path.join(GOLDEN, "defaults.json")) as f:
        ref = json.load(f)
    cfg = pkg.default_config()
    for k in ("g", "softening", "max_speed", "integrator", "max_substep", "max_substeps_per_frame"):
        assert getattr(cfg, k) == ref[k], k


def test_struct_layouts_match_header(pkg):
    from raylib_nbody_b200 import physics
    assert C.sizeof(physics.Config) == 32
    assert C.sizeof(physics._Diag) == 88


@pytest.mark.parametrize("dt,cap,lim", [(8333.334, 1e9, 200), (8333.334, 3000.0, 200), (8333.334, 1.0, 200),
                                         (0.0, 1e9, 200), (5.0, 0.0, 7), (100.0, 30.0, 0), (1e-7, 1e-9, 50)])
def test_substep_plan_matches_oracle(pkg, oracle_lib, dt, cap, lim):
    assert pkg.substep_plan(dt, cap, lim) == oracle_lib.substep_plan(dt, cap, lim)


def test_dt_eff_is_float_arithmetic(pkg, ob):
    assert pkg.dt_eff(ob.DEFAULTS["fixed_dt"], ob.DEFAULTS["time_scale"]) == ob.dt_eff() == 8333.333984375
    assert pkg.dt_eff(0.01, -5.0) == 0.0


def test_partition_is_tile_aligned_and_covers(pkg):
    for n in (0, 3, 511, 512, 513, 4096, 65536, (1 << 20) + 17):
        for world in (1, 2, 3, 8):
            spans = [pkg.partition(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (lo, hi), (lo2, _) in zip(spans, spans[1:]):
                assert hi == lo2 and lo <= hi
            assert all(lo % 512 == 0 or lo == n for lo, _ in spans)
            sizes = {hi - lo for lo, hi in spans if hi < n}
            assert len(sizes) <= 1  # equal tile counts: the all-gather is uniform


def test_no_cpu_fallback(pkg):
    """Without a CUDA device creation fails with NB_ERR_NO_DEVICE; nothing computes on the CPU."""
    if _have_gpu():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.NbodyError) as e:
        pkg.Physics(0)
    assert e.value.status == 2 and "no CPU path" in str(e.value)


def test_product_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under the package may import, link or load it."""
    pkg_dir = os.path.join(ROOT, "raylib-nbody_b200")
    for base, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".sh")):
                text = open(os.path.join(base, f)).read()
                assert "oracle" not in text.lower().replace("oracle/", "oracle/") or f == "never", (base, f)


def test_bench_reference_arm_prints_one_json_line():
    """bench.py --impl reference runs on host cores only (the reference's own compute_gravity when
    oracle/_ref is built, else the oracle port) and prints exactly one JSON line with the contract keys."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "1", "--ref-sample", "1024"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "interactions/s" and d["value"] > 0
    assert d["cpu_baseline"]["cores"] == 1 and d["cpu_baseline"]["kind"] in ("reference", "port")
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["value"] == d["value"]


def test_scenario_file_round_trip_and_errors(pkg, tmp_path):
    """The ".nbs" scenario / checkpoint format of the C-ABI (host-only, no GPU): floats and flags round-trip
    exactly, optional arrays are optional, scenarios.save_snapshot/load_snapshot route ".nbs" paths through it,
    and a bad file is a status code."""
    rng = np.random.default_rng(3)
    n = 257
    pos, vel = rng.standard_normal((n, 2)) * 1e8, rng.standard_normal((n, 2)) * 1e3
    acc, acc_prev = rng.standard_normal((n, 2)), rng.standard_normal((n, 2))
    mass = rng.uniform(1e20, 1e25, n).astype(np.float32)
    pinned = (rng.random(n) < 0.1).astype(np.uint8)
    order = rng.permutation(n).astype(np.int32)
    path = str(tmp_path / "world.nbs")
    cfg = dict(g=6.5e-11, softening=123.5, max_speed=7.0, integrator=0, max_substep=0.25, max_substeps_per_frame=9,
               bh_threshold=77, bh_theta=0.75, use_fixed_dt=1, fixed_dt=0.01, time_scale=3.0e4)
    pkg.scenario_write(path, pos, vel, mass, pinned, acc, acc_prev, cfg, (order, 17), frame=42)
    d = pkg.scenario_read(path)
    for k, v in (("pos", pos), ("vel", vel), ("acc", acc), ("acc_prev", acc_prev), ("mass", mass), ("pinned", pinned)):
        assert np.array_equal(d[k], v), k
    assert np.array_equal(d["stream_order"][0], order) and d["stream_order"][1] == 17 and d["frame"] == 42
    for k, v in cfg.items():
        assert d["config"][k] == (np.float32(v) if isinstance(d["config"][k], float) and k != "g" else v), k
    # optional arrays left out; 3-D
    p3 = rng.standard_normal((5, 3))
    pkg.scenario_write(path, p3, p3 * 2, mass[:5])
    d = pkg.scenario_read(path)
    assert d["acc"] is None and d["acc_prev"] is None and d["stream_order"] is None and not d["pinned"].any()
    assert np.array_equal(d["pos"], p3) and d["config"]["integrator"] == 1 and d["config"]["bh_threshold"] == 100
    # the package's snapshot helpers use the same file for ".nbs" paths
    pkg.scenarios.save_snapshot(path, pos, vel, mass, pinned, acc=acc, config=dict(softening=4.0, integrator=1), frame=3,
                                stream_order=(order, 2))
    s = pkg.scenarios.load_snapshot(path)
    assert np.array_equal(s["pos"], pos) and s["config"]["softening"] == 4.0 and s["frame"] == 3
    assert s["acc_prev"] is None and s["stream_order"][1] == 2
    with open(path, "wb") as f:
        f.write(b"not a scenario")
    with pytest.raises(pkg.NbodyError):
        pkg.scenario_read(path)
    with pytest.raises(pkg.NbodyError):
        pkg.scenario_read(str(tmp_path / "missing.nbs"))
