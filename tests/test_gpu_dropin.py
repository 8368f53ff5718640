"""The flecs drop-in: oracle/_ref/dropin_check runs the UNMODIFIED reference
systems and include/reference_side/PhysicsB200.hpp (-> libnbody_b200.so) on twin flecs worlds and
compares the components. The binary is built here by oracle/Makefile (it needs the reference's
headers, which do not exist on the GPU box) and travels as a built artefact."""
import os
import subprocess

import pytest

from conftest import ROOT

BIN = os.path.join(ROOT, "oracle", "_ref", "dropin_check")


@pytest.mark.gpu
def test_flecs_dropin_matches_reference_systems():
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/dropin_check not built (needs /root/reference at build time)")
    r = subprocess.run([BIN], capture_output=True, text=True, timeout=600)
    print(r.stdout, r.stderr)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "DROPIN CHECK PASSED" in r.stdout and "FAIL" not in r.stdout


@pytest.mark.gpu
def test_flecs_dropin_on_several_gpus_in_one_process():
    """Scenario F of the same binary: PhysicsB200::register_systems(w, devices) -> nb_create_multi. The reference's
    systems are one process and one thread; so is this (no torchrun, no NCCL: peer-memory exchange)."""
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/dropin_check not built (needs /root/reference at build time)")
    env = dict(os.environ, NB_DROPIN_DEVICES=",".join(str(i) for i in range(min(n, 8))), NB_DROPIN_ONLY="F")
    r = subprocess.run([BIN], capture_output=True, text=True, timeout=900, env=env)
    print(r.stdout, r.stderr)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "GPUs in one process" in r.stdout and "FAIL" not in r.stdout


def test_reference_side_binding_builds_and_links():
    """CPU: the binding compiles against the reference's headers and resolves every nb_* symbol
    from the product library (no compute call)."""
    if not os.path.isdir("/root/reference/src"):
        if not os.path.exists(BIN):
            pytest.skip("no reference tree and no prebuilt binary")
    else:
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=True)
    assert os.path.exists(BIN)
    out = subprocess.run(["ldd", BIN], capture_output=True, text=True).stdout
    assert "libnbody_b200.so" in out and "not found" not in out.split("libnbody_b200.so")[1].split("\n")[0]
    syms = subprocess.run(["nm", "-D", "--undefined-only", BIN], capture_output=True, text=True).stdout
    for s in ("nb_create", "nb_upload_state", "nb_step", "nb_download_state", "nb_compute_diagnostics"):
        assert s in syms
