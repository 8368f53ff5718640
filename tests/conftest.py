import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def pkg():
    import __graft_entry__ as ge
    p = ge.load_package()
    if not os.path.exists(p.LIB_PATH):
        ge.build()
    return p


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle import bindings as ob
    if not os.path.exists(ob.ORACLE_SO):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=True)
    return ob.Oracle()


@pytest.fixture(scope="session")
def ob():
    from oracle import bindings
    return bindings


def golden(name):
    return np.load(os.path.join(GOLDEN, name))


def rel_err(a, ref):
    return np.linalg.norm(a - ref, axis=1) / np.linalg.norm(ref, axis=1)


def conditioning(pos, mass, eps=4.0):
    """Per-body condition number of the direct sum: sum_j |term_ij| / |sum_j term_ij| (FP64, numpy).
    A body whose pulls nearly cancel has a large value; any FP32 evaluation of the terms carries a
    relative error of about (per-term rounding ~1e-7) x this number."""
    n = len(pos)
    out = np.empty(n)
    m = mass.astype(np.float64)
    for i in range(n):
        d = pos - pos[i]
        r2 = (d * d).sum(axis=1) + eps * eps
        t = (m / (r2 * np.sqrt(r2)))[:, None] * d
        out[i] = np.abs(t).sum() / max(np.linalg.norm(t.sum(axis=0)), 1e-300)
    return out
