"""raylib-nbody_b200 — B200-native physics step of tr0yspradling/raylib-nbody.

Only the hot path lives here: the CUDA kernels + C-ABI (csrc/, libnbody_b200.so) and the
host-side mirror of the reference's Physics interface (physics.py). Import through
``__graft_entry__.load_package()`` (the directory name is not a Python identifier).
"""
from .physics import (Config, Diagnostics, NbodyError, Physics, comm_unique_id, default_config, partition, pinned_array, dt_eff,  # noqa: F401
                      load_library, substep_plan, scenario_read, scenario_write, ScenarioConfig, EXPORTS, LIB_PATH, PRECISION_AUTO, PRECISION_FP32,
                      PRECISION_FP64_EXACT, SOLVER_DIRECT, SOLVER_REFERENCE, REF_BH_THRESHOLD, REF_BH_THETA,
                      diagnostics_noalloc, diag_from_bytes, NOALLOC_MAX_BODIES, DIAG_STRUCT_BYTES)
from . import scenarios, sharded  # noqa: F401
