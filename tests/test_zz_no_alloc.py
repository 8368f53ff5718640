"""nb_diagnostics_noalloc: the allocation-free entry point (static device scratch), against nb_compute_diagnostics.

Collected last on purpose (file name): a failure here hides nothing.
The entry point has no counterpart in the reference (include/nbody_b200.h says why it exists); what it must
equal is this library's own nb_compute_diagnostics on a context in NB_PRECISION_FP64_EXACT — the kernel it is
a variant of — on the same bodies. Same FP64 operations in the same order, so the tolerance is the one
tests/test_gpu_parity.py uses for the exact diagnostics path (1e-12 relative), and in practice the values are
identical. Two calls in a row on different worlds: the second must not see the first one's scratch contents.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-12


def _existing(pkg, pos, vel, mass, g, eps):
    ph = pkg.Physics(0)
    ph.set_precision(pkg.PRECISION_FP64_EXACT)
    ph.set_config(g=g, softening=eps)
    ph.upload(pos, vel, mass)
    d = ph.compute_diagnostics()
    ph.close()
    return d


def _noalloc(pkg, torch, pos, vel, mass, g, eps, out, check_no_allocation=False):
    """Inputs and output are torch CUDA tensors the TEST allocates; the entry point gets raw device addresses."""
    n, dim = pos.shape
    d_pos = torch.from_numpy(np.ascontiguousarray(pos)).cuda()
    d_vel = torch.from_numpy(np.ascontiguousarray(vel)).cuda()
    d_mass = torch.from_numpy(np.ascontiguousarray(mass, np.float32)).cuda()
    out.fill_(0xAB)
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    free0 = torch.cuda.mem_get_info()[0]
    pkg.diagnostics_noalloc(0, n, dim, d_pos.data_ptr(), d_vel.data_ptr(), d_mass.data_ptr(), g, eps, out.data_ptr(), stream)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    if check_no_allocation:  # after the first call (module and symbol already resolved) device memory must not move
        assert free1 == free0
    return pkg.diag_from_bytes(out.cpu().numpy().tobytes())


def _same(a, b):
    assert a.ok and b.ok
    scale_p = max(np.abs(b.momentum).max(), 1e-300)
    assert abs(a.kinetic - b.kinetic) <= RTOL * abs(b.kinetic)
    assert abs(a.potential - b.potential) <= RTOL * abs(b.potential)
    assert abs(a.energy - b.energy) <= RTOL * max(abs(b.kinetic), abs(b.potential))
    assert np.abs(a.momentum - b.momentum).max() <= RTOL * max(scale_p, np.sqrt(2 * b.kinetic * b.total_mass))
    assert np.abs(a.com - b.com).max() <= RTOL * max(np.abs(b.com).max(), 3.844e8)
    assert abs(a.total_mass - b.total_mass) <= RTOL * b.total_mass


def test_noalloc_diagnostics_equal_compute_diagnostics_twice_in_a_row(pkg):
    import torch
    sc = pkg.scenarios
    out = torch.empty(pkg.DIAG_STRUCT_BYTES, dtype=torch.uint8, device="cuda")
    # call 1: 3-D Plummer sphere, 5000 bodies (20 partial rows)
    p1, v1, m1 = sc.plummer(5000, seed=3, dim=3)
    eps1 = 3.844e8 / np.sqrt(5000)
    got1 = _noalloc(pkg, torch, p1, v1, m1, 6.6743e-11, eps1, out)
    # call 2 right behind it: smaller 2-D disk with mixed masses (3 rows: the stale rows 3..19 must not be read)
    p2, v2, m2 = sc.uniform_disk(700, seed=9, circular=True)
    m2 = (m2 * np.random.default_rng(4).uniform(0.5, 2.0, len(m2))).astype(np.float32)
    got2 = _noalloc(pkg, torch, p2, v2, m2, 6.6743e-11, 4.0, out, check_no_allocation=True)
    _same(got1, _existing(pkg, p1, v1, m1, 6.6743e-11, eps1))
    _same(got2, _existing(pkg, p2, v2, m2, 6.6743e-11, 4.0))
    # and the first world again: equal to its first answer bit for bit (fixed summation order, shared scratch)
    again = _noalloc(pkg, torch, p1, v1, m1, 6.6743e-11, eps1, out, check_no_allocation=True)
    assert again.energy == got1.energy and np.array_equal(again.momentum, got1.momentum)


def test_noalloc_rejects_sizes_beyond_the_static_scratch_and_handles_the_empty_world(pkg):
    import torch
    out = torch.empty(pkg.DIAG_STRUCT_BYTES, dtype=torch.uint8, device="cuda")
    one = torch.zeros(8, dtype=torch.float64, device="cuda")
    with pytest.raises(pkg.NbodyError) as err:
        pkg.diagnostics_noalloc(0, pkg.NOALLOC_MAX_BODIES + 1, 3, one.data_ptr(), one.data_ptr(), one.data_ptr(), 1.0, 1.0,
                                out.data_ptr(), 0)
    assert "NB_NOALLOC_MAX_BODIES" in str(err.value)
    with pytest.raises(pkg.NbodyError):
        pkg.diagnostics_noalloc(0, 10, 4, one.data_ptr(), one.data_ptr(), one.data_ptr(), 1.0, 1.0, out.data_ptr(), 0)
    # Physics.hpp:124-127: no bodies -> zeroed diagnostics with ok = true
    out.fill_(0xAB)
    pkg.diagnostics_noalloc(0, 0, 2, 0, 0, 0, 1.0, 1.0, out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    d = pkg.diag_from_bytes(out.cpu().numpy().tobytes())
    assert d.ok and d.energy == 0.0 and d.total_mass == 0.0
