"""Host-side mirror of the reference's ``nbody::Physics`` interface over the C-ABI.

Same names, argument meaning and error behaviour as the statics of
``src/systems/Physics.hpp`` (reference file:line in each docstring); every call goes
through ``libnbody_b200.so`` (include/nbody_b200.h). There is no CPU fallback: if the
CUDA library or a B200 is missing this module raises.
"""
import ctypes as C
import os

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NB_LIB_PATH") or os.path.join(_PKG, "libnbody_b200.so")  # NB_LIB_PATH: developer A/B builds

_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)
_bp = C.POINTER(C.c_uint8)


class NbodyError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"nbody_b200 status {status}: {message}")
        self.status = status


class Config(C.Structure):
    """Physics fields of the reference's Config singleton (src/core/Config.hpp:13-29)."""
    _fields_ = [("g", C.c_double), ("softening", C.c_float), ("max_speed", C.c_float),
                ("integrator", C.c_int32), ("max_substep", C.c_float), ("max_substeps_per_frame", C.c_int32)]


class ScenarioConfig(C.Structure):
    """The Config subset a Scenario carries (Scenario.hpp:26-38), physics fields only (nb_scenario_config)."""
    _fields_ = [("physics", Config), ("bh_threshold", C.c_int32), ("bh_theta", C.c_float), ("use_fixed_dt", C.c_int32),
                ("fixed_dt", C.c_float), ("time_scale", C.c_float)]


class _Diag(C.Structure):
    _fields_ = [("kinetic", C.c_double), ("potential", C.c_double), ("energy", C.c_double),
                ("momentum", C.c_double * 3), ("com", C.c_double * 3), ("total_mass", C.c_double),
                ("ok", C.c_int32), ("_pad", C.c_int32)]


class Diagnostics:
    """Physics::Diagnostics (Physics.hpp:22-30)."""

    def __init__(self, d):
        self.kinetic = d.kinetic
        self.potential = d.potential
        self.energy = d.energy
        self.momentum = np.array(d.momentum[:])
        self.com = np.array(d.com[:])
        self.total_mass = d.total_mass
        self.ok = bool(d.ok)

    def as_reference_array(self):
        """[KE, PE, E, Px, Py, COMx, COMy, M] — the 2-D field order of Physics.hpp:22-30."""
        return np.array([self.kinetic, self.potential, self.energy, self.momentum[0], self.momentum[1],
                         self.com[0], self.com[1], self.total_mass])


EXPORTS = ["nb_create", "nb_create_multi", "nb_num_devices", "nb_comm_unique_id", "nb_create_sharded", "nb_destroy", "nb_last_error", "nb_abi_version",
           "nb_default_config", "nb_set_config", "nb_get_config", "nb_set_precision", "nb_set_solver", "nb_get_solver", "nb_solver_info", "nb_solver_profile", "nb_upload", "nb_upload3", "nb_upload_state", "nb_download",
           "nb_download3", "nb_download_state", "nb_stream_order", "nb_set_stream_order", "nb_num_bodies", "nb_local_range", "nb_partition", "nb_compute_gravity", "nb_integrate", "nb_step",
           "nb_step_with_diagnostics", "nb_step_diagnostics", "nb_compute_diagnostics", "nb_zero_net_momentum", "nb_count_overlaps", "nb_resolve_collisions", "nb_debug_cbrt", "nb_merge_info",
           "nb_substep_plan", "nb_reset_scenario", "nb_scenario_default_config", "nb_scenario_write", "nb_scenario_read_header", "nb_scenario_read", "nb_synchronize", "nb_timer_begin", "nb_timer_end", "nb_counters",
           "nb_profile_forces", "nb_profile_read", "nb_near_stats", "nb_overlap_profile", "nb_device_info", "nb_flush_l2", "nb_alloc_pinned", "nb_free_pinned",
           "nb_diagnostics_noalloc", "nb_partition_info"]

PRECISION_AUTO, PRECISION_FP32, PRECISION_FP64_EXACT = 0, 1, 2
SOLVER_DIRECT, SOLVER_REFERENCE = 0, 1
BODY_DENSITY = 5510.0  # Constants.hpp:34
REF_BH_THRESHOLD, REF_BH_THETA = 100, 0.5  # Constants.hpp:90-91

_lib = None


def load_library(path=LIB_PATH):
    """dlopen the C-ABI library; fails loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(path):
        raise NbodyError(-1, f"{path} not found: build it with raylib-nbody_b200/csrc/build.sh "
                             "(or __graft_entry__.build()); there is no CPU fallback")
    L = C.CDLL(path)
    L.nb_last_error.restype = C.c_char_p
    L.nb_num_bodies.restype = C.c_int64
    L.nb_num_bodies.argtypes = [C.c_void_p]
    L.nb_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    if hasattr(L, "nb_create_multi"):
        L.nb_create_multi.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_void_p)]
        L.nb_num_devices.argtypes = [C.c_void_p]
    L.nb_create_sharded.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]
    L.nb_destroy.argtypes = [C.c_void_p]
    L.nb_destroy.restype = None
    L.nb_upload.argtypes = [C.c_void_p, C.c_int64, _dp, _dp, _fp, _bp]
    L.nb_upload3.argtypes = [C.c_void_p, C.c_int64, _dp, _dp, _fp, _bp]
    L.nb_upload_state.argtypes = [C.c_void_p, C.c_int64, C.c_int, _dp, _dp, _dp, _fp, _bp]
    L.nb_flush_l2.argtypes = [C.c_void_p]
    L.nb_device_info.argtypes = [C.c_void_p] + [C.POINTER(C.c_int)] * 4
    L.nb_alloc_pinned.argtypes = [C.c_size_t, C.POINTER(C.c_void_p)]
    L.nb_free_pinned.argtypes = [C.c_void_p]
    L.nb_download.argtypes = [C.c_void_p, _dp, _dp, _dp]
    L.nb_download3.argtypes = [C.c_void_p, _dp, _dp, _dp]
    L.nb_download_state.argtypes = [C.c_void_p, C.c_int, _dp, _dp, _dp, _dp]
    if hasattr(L, "nb_stream_order"):
        L.nb_stream_order.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int64)]
        L.nb_set_stream_order.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.c_int64]
    L.nb_count_overlaps.argtypes = [C.c_void_p, _dp, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    L.nb_resolve_collisions.argtypes = [C.c_void_p, _dp, _dp, C.POINTER(C.c_int64), C.POINTER(C.c_uint8), C.POINTER(C.c_uint8),
                                        _dp, _dp, C.POINTER(C.c_float), C.POINTER(C.c_uint8), _dp, C.POINTER(C.c_int64), C.c_int64]
    if hasattr(L, "nb_debug_cbrt"):
        L.nb_debug_cbrt.argtypes = [C.c_void_p, _dp, _dp, C.c_int64]
    if hasattr(L, "nb_merge_info"):
        L.nb_merge_info.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
    L.nb_set_config.argtypes = [C.c_void_p, C.POINTER(Config)]
    L.nb_get_config.argtypes = [C.c_void_p, C.POINTER(Config)]
    L.nb_default_config.argtypes = [C.POINTER(Config)]
    L.nb_set_precision.argtypes = [C.c_void_p, C.c_int]
    L.nb_set_solver.argtypes = [C.c_void_p, C.c_int, C.c_int32, C.c_float]
    L.nb_get_solver.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int32), C.POINTER(C.c_float)]
    L.nb_solver_info.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
    L.nb_solver_profile.argtypes = [C.c_void_p, C.c_int]
    L.nb_local_range.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    L.nb_partition.argtypes = [C.c_int64, C.c_int, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    L.nb_compute_gravity.argtypes = [C.c_void_p]
    L.nb_integrate.argtypes = [C.c_void_p, C.c_float]
    L.nb_step.argtypes = [C.c_void_p, C.c_float, C.c_int64]
    L.nb_step_with_diagnostics.argtypes = [C.c_void_p, C.c_float, C.c_int64]
    L.nb_step_diagnostics.argtypes = [C.c_void_p, C.POINTER(_Diag)]
    L.nb_compute_diagnostics.argtypes = [C.c_void_p, C.POINTER(_Diag)]
    L.nb_zero_net_momentum.argtypes = [C.c_void_p]
    if hasattr(L, "nb_diagnostics_noalloc"):
        L.nb_diagnostics_noalloc.argtypes = [C.c_int, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double,
                                             C.c_float, C.c_void_p, C.c_void_p]
    L.nb_substep_plan.argtypes = [C.c_float, C.c_float, C.c_int32, _fp]
    L.nb_reset_scenario.argtypes = [C.c_void_p]
    if hasattr(L, "nb_scenario_write"):
        L.nb_scenario_default_config.argtypes = [C.POINTER(ScenarioConfig)]
        L.nb_scenario_write.argtypes = [C.c_char_p, C.c_int64, C.c_int, _dp, _dp, _dp, _dp, _fp, _bp, C.POINTER(ScenarioConfig),
                                        C.POINTER(C.c_int32), C.c_int64, C.c_int64]
        L.nb_scenario_read_header.argtypes = [C.c_char_p, C.POINTER(C.c_int64), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                              C.POINTER(ScenarioConfig), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.nb_scenario_read.argtypes = [C.c_char_p, _dp, _dp, _dp, _dp, _fp, _bp, C.POINTER(C.c_int32)]
    L.nb_synchronize.argtypes = [C.c_void_p]
    L.nb_timer_begin.argtypes = [C.c_void_p]
    L.nb_timer_end.argtypes = [C.c_void_p, _fp]
    L.nb_counters.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), _dp]
    L.nb_profile_forces.argtypes = [C.c_void_p, C.c_int]
    L.nb_profile_read.argtypes = [C.c_void_p, C.POINTER(C.c_int64), _dp]
    L.nb_comm_unique_id.argtypes = [C.c_void_p]
    if hasattr(L, "nb_near_stats"):  # absent from developer A/B builds of older sources (NB_LIB_PATH)
        L.nb_overlap_profile.argtypes = [C.c_void_p, C.c_int, _fp]
        L.nb_near_stats.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    _lib = L
    return L


def default_config():
    """Defaults of src/core/Constants.hpp:86-100 and Config.hpp:25."""
    cfg = Config()
    _check(load_library().nb_default_config(C.byref(cfg)))
    return cfg


def _check(status):
    if status != 0:
        raise NbodyError(status, (load_library().nb_last_error() or b"").decode())


def scenario_write(path, pos, vel, mass, pinned=None, acc=None, acc_prev=None, config=None, stream_order=None, frame=0):
    """One ".nbs" scenario / checkpoint file (nb_scenario_write): readable by every binding of the C-ABI."""
    L = load_library()
    pos = np.ascontiguousarray(pos, np.float64)
    vel = np.ascontiguousarray(vel, np.float64)
    mass = np.ascontiguousarray(mass, np.float32)
    n, dim = pos.shape
    cfg = ScenarioConfig()
    _check(L.nb_scenario_default_config(C.byref(cfg)))
    for k, v in (config or {}).items():
        target = cfg.physics if hasattr(cfg.physics, k) else cfg
        if not hasattr(target, k):
            raise KeyError(f"not a scenario config field: {k}")
        setattr(target, k, type(getattr(target, k))(v))
    opt = lambda a, t: None if a is None else np.ascontiguousarray(a, t)
    acc, acc_prev, pinned = opt(acc, np.float64), opt(acc_prev, np.float64), opt(pinned, np.uint8)
    order = None if stream_order is None else np.ascontiguousarray(stream_order[0], np.int32)
    ptr = lambda a, t: None if a is None else a.ctypes.data_as(t)
    _check(L.nb_scenario_write(os.fsencode(path), n, dim, ptr(pos, _dp), ptr(vel, _dp), ptr(acc, _dp), ptr(acc_prev, _dp),
                               ptr(mass, _fp), ptr(pinned, _bp), C.byref(cfg), ptr(order, C.POINTER(C.c_int32)),
                               0 if stream_order is None else int(stream_order[1]), int(frame)))


def scenario_read(path):
    """-> dict(pos, vel, mass, pinned, acc|None, acc_prev|None, stream_order|None, config, frame) of a ".nbs" file."""
    L = load_library()
    n, dim, flags, age, frame = C.c_int64(), C.c_int(), C.c_int(), C.c_int64(), C.c_int64()
    cfg = ScenarioConfig()
    _check(L.nb_scenario_read_header(os.fsencode(path), C.byref(n), C.byref(dim), C.byref(flags), C.byref(cfg), C.byref(age),
                                     C.byref(frame)))
    n, dim, flags = n.value, dim.value, flags.value
    pos, vel = np.zeros((n, dim)), np.zeros((n, dim))
    mass, pinned = np.zeros(n, np.float32), np.zeros(n, np.uint8)
    acc = np.zeros((n, dim)) if flags & 1 else None
    acc_prev = np.zeros((n, dim)) if flags & 2 else None
    order = np.zeros(n, np.int32) if flags & 4 else None
    ptr = lambda a, t: None if a is None else a.ctypes.data_as(t)
    _check(L.nb_scenario_read(os.fsencode(path), ptr(pos, _dp), ptr(vel, _dp), ptr(acc, _dp), ptr(acc_prev, _dp), ptr(mass, _fp),
                              ptr(pinned, _bp), ptr(order, C.POINTER(C.c_int32))))
    config = {k: getattr(cfg.physics, k) for k, _ in Config._fields_}
    config.update({k: getattr(cfg, k) for k in ("bh_threshold", "bh_theta", "use_fixed_dt", "fixed_dt", "time_scale")})
    return {"pos": pos, "vel": vel, "mass": mass, "pinned": pinned, "acc": acc, "acc_prev": acc_prev,
            "stream_order": None if order is None else (order, age.value), "config": config, "frame": frame.value}


def substep_plan(dt, max_substep, max_substeps_per_frame):
    """(nSteps, dtSub) of Physics.hpp:268-271."""
    out = C.c_float()
    n = load_library().nb_substep_plan(C.c_float(dt), C.c_float(max_substep), C.c_int32(max_substeps_per_frame),
                                       C.byref(out))
    return n, out.value


def dt_eff(fixed_dt, time_scale):
    """dtEff = baseDt * max(0, time_scale) in float, Physics.hpp:54-55 with use_fixed_dt."""
    return float(np.float32(fixed_dt) * max(np.float32(0.0), np.float32(time_scale)))


NOALLOC_MAX_BODIES = 262144  # NB_NOALLOC_MAX_BODIES
DIAG_STRUCT_BYTES = C.sizeof(_Diag)


def diagnostics_noalloc(device, n, dim, d_pos, d_vel, d_mass, g, softening, d_out, stream=0):
    """nb_diagnostics_noalloc (no counterpart in the reference): compute_diagnostics over DEVICE arrays given as
    integer addresses, result as one nb_diag at device address d_out; enqueues on `stream` and does not
    synchronise, allocates nothing. diag_from_bytes() decodes the struct once it is back on the host."""
    _check(load_library().nb_diagnostics_noalloc(device, n, dim, C.c_void_p(d_pos), C.c_void_p(d_vel), C.c_void_p(d_mass),
                                                 C.c_double(g), C.c_float(softening), C.c_void_p(d_out),
                                                 C.c_void_p(stream)))


def diag_from_bytes(raw):
    """Diagnostics from the DIAG_STRUCT_BYTES bytes of an nb_diag."""
    return Diagnostics(_Diag.from_buffer_copy(bytes(raw)))


def partition(n, rank, world):
    """Rows [lo,hi) rank `rank` of `world` owns (host-only; mirrors nb_create_sharded's split)."""
    lo, hi = C.c_int64(), C.c_int64()
    _check(load_library().nb_partition(n, rank, world, C.byref(lo), C.byref(hi)))
    return lo.value, hi.value


def pinned_array(shape, dtype):
    """numpy array over page-locked host memory (nb_alloc_pinned); freed with the array."""
    dtype = np.dtype(dtype)
    nbytes = int(np.prod(shape)) * dtype.itemsize
    ptr = C.c_void_p()
    _check(load_library().nb_alloc_pinned(nbytes, C.byref(ptr)))
    buf = (C.c_char * max(nbytes, 1)).from_address(ptr.value)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    import weakref
    weakref.finalize(buf, load_library().nb_free_pinned, ptr)
    return arr


def comm_unique_id():
    buf = C.create_string_buffer(128)
    _check(load_library().nb_comm_unique_id(buf))
    return buf.raw


class Physics:
    """Device-resident world state with the reference's Physics entry points."""

    def __init__(self, device=0, rank=0, world=1, unique_id=None, devices=None):
        """devices=[0, 1, ...]: one single-process context over several GPUs (nb_create_multi)."""
        L = load_library()
        self._L = L
        self._h = C.c_void_p()
        if devices is not None:
            arr = (C.c_int * len(devices))(*devices)
            _check(L.nb_create_multi(arr, len(devices), C.byref(self._h)))
        elif world > 1:
            _check(L.nb_create_sharded(device, rank, world, unique_id, C.byref(self._h)))
        else:
            _check(L.nb_create(device, C.byref(self._h)))
        self.rank, self.world = rank, world
        self.dim = 2

    def close(self):
        if self._h:
            self._L.nb_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- Config singleton ------------------------------------------------------------
    def set_config(self, cfg=None, **fields):
        if cfg is None:
            cfg = self.get_config()
        for k, v in fields.items():
            setattr(cfg, k, v)
        _check(self._L.nb_set_config(self._h, C.byref(cfg)))
        return cfg

    def get_config(self):
        cfg = Config()
        _check(self._L.nb_get_config(self._h, C.byref(cfg)))
        return cfg

    def set_precision(self, mode):
        """PRECISION_AUTO / PRECISION_FP32 / PRECISION_FP64_EXACT (include/nbody_b200.h)."""
        _check(self._L.nb_set_precision(self._h, int(mode)))

    def set_solver(self, solver=SOLVER_REFERENCE, bh_threshold=REF_BH_THRESHOLD, bh_theta=REF_BH_THETA):
        """Config::bh_threshold / bh_theta (Config.hpp:17-18). SOLVER_REFERENCE follows the reference's
        switch (Physics.hpp:212): its float Barnes-Hut quadtree above bh_threshold bodies, bit-identical;
        SOLVER_DIRECT (the default) always evaluates the all-pairs sum."""
        _check(self._L.nb_set_solver(self._h, int(solver), int(bh_threshold), float(bh_theta)))

    def get_solver(self):
        s, t, th = C.c_int(), C.c_int32(), C.c_float()
        _check(self._L.nb_get_solver(self._h, C.byref(s), C.byref(t), C.byref(th)))
        return s.value, t.value, th.value

    def solver_profile(self, enable=True):
        _check(self._L.nb_solver_profile(self._h, 1 if enable else 0))

    def solver_info(self):
        """Tree of the last Barnes-Hut pass (and, after solver_profile(True), what its walks did)."""
        v = (C.c_int64 * 8)()
        _check(self._L.nb_solver_info(self._h, v))
        return dict(passes=v[0], cells=v[1], depth=v[2], bodies=v[3], visited=v[4], terms=v[5], exact_tests=v[6])

    # ---- bodies ------------------------------------------------------------------------
    def upload(self, pos, vel, mass, pinned=None):
        """Replace all bodies; pos/vel are (n,2) [reference layout] or (n,3) float64."""
        pos = np.ascontiguousarray(pos, np.float64)
        vel = np.ascontiguousarray(vel, np.float64)
        mass = np.ascontiguousarray(mass, np.float32)
        n, dim = (pos.shape[0], pos.shape[1]) if pos.ndim == 2 else (0, 2)
        if pos.ndim != 2 or dim not in (2, 3) or vel.shape != pos.shape or mass.shape != (n,):
            raise NbodyError(1, "upload: pos/vel must be (n,2) or (n,3), mass (n,)")
        pin = None
        if pinned is not None:
            pin = np.ascontiguousarray(pinned, np.uint8)
            if pin.shape != (n,):
                raise NbodyError(1, "upload: pinned must be (n,)")
        fn = self._L.nb_upload if dim == 2 else self._L.nb_upload3
        _check(fn(self._h, n, pos.ctypes.data_as(_dp), vel.ctypes.data_as(_dp), mass.ctypes.data_as(_fp),
                  None if pin is None else pin.ctypes.data_as(_bp)))
        self.dim = dim

    def upload_state(self, pos, vel, acc, mass, pinned=None):
        """upload() that also carries the caller's Acceleration components (asserted == a(x))."""
        pos = np.ascontiguousarray(pos, np.float64)
        vel = np.ascontiguousarray(vel, np.float64)
        mass = np.ascontiguousarray(mass, np.float32)
        if pos.ndim != 2 or pos.shape[1] not in (2, 3) or vel.shape != pos.shape or mass.shape != (pos.shape[0],):
            raise NbodyError(1, "upload_state: pos/vel must be (n,2) or (n,3), mass (n,)")
        n, dim = pos.shape
        if acc is not None:
            acc = np.ascontiguousarray(acc, np.float64)
            if acc.shape != pos.shape:
                raise NbodyError(1, "upload_state: acc must have the shape of pos")
        if pinned is not None:
            pinned = np.ascontiguousarray(pinned, np.uint8)
            if pinned.shape != (n,):
                raise NbodyError(1, "upload_state: pinned must be (n,)")
        _check(self._L.nb_upload_state(self._h, n, dim, pos.ctypes.data_as(_dp), vel.ctypes.data_as(_dp),
                                       None if acc is None else acc.ctypes.data_as(_dp), mass.ctypes.data_as(_fp),
                                       None if pinned is None else pinned.ctypes.data_as(_bp)))
        self.dim = dim

    def download_into(self, pos, vel, acc):
        """download() into caller-owned (n,dim) float64 arrays (e.g. pinned ones)."""
        n = self.num_bodies()
        for a in (pos, vel, acc):  # written in place: must already be what the C side expects
            if a is not None and not (isinstance(a, np.ndarray) and a.dtype == np.float64 and a.flags.c_contiguous
                                      and a.flags.writeable and a.shape == (n, self.dim)):
                raise NbodyError(1, f"download_into: arrays must be writable C-contiguous float64 of shape ({n}, {self.dim})")
        fn = self._L.nb_download if self.dim == 2 else self._L.nb_download3
        _check(fn(self._h, *[None if a is None else a.ctypes.data_as(_dp) for a in (pos, vel, acc)]))

    def device_info(self):
        v = [C.c_int() for _ in range(4)]
        _check(self._L.nb_device_info(self._h, *[C.byref(x) for x in v]))
        return dict(sm_count=v[0].value, cc=(v[1].value, v[2].value), sm_clock_khz=v[3].value)

    def flush_l2(self):
        _check(self._L.nb_flush_l2(self._h))

    def download(self, pos=True, vel=True, acc=True):
        """(pos, vel, acc) as (n,dim) float64; a sharded context fills its own rows only."""
        n = self.num_bodies()
        out = [np.zeros((n, self.dim)) if want else None for want in (pos, vel, acc)]
        fn = self._L.nb_download if self.dim == 2 else self._L.nb_download3
        _check(fn(self._h, *[None if a is None else a.ctypes.data_as(_dp) for a in out]))
        return tuple(out)

    def count_overlaps(self, radius):
        """Detection phase of Collision::resolve, Collision.hpp:63-84: (n_pairs, first_pair or None)."""
        r = np.ascontiguousarray(radius, dtype=np.float64)
        if r.shape != (self.num_bodies(),):
            raise ValueError("radius must have one entry per body")
        cnt, first = C.c_int64(), (C.c_int64 * 2)()
        _check(self._L.nb_count_overlaps(self._h, r.ctypes.data_as(_dp), C.byref(cnt), first))
        return cnt.value, ((first[0], first[1]) if cnt.value else None)

    def resolve_collisions(self, radius, density=None, with_log=False):
        """Collision::resolve, merge mode (Collision.hpp:49-132). Returns (n_merges, alive, changed, pos, vel, mass,
        pinned, radius): alive[i] = 0 where the reference destructs the body; for changed[i] = 1 the survivor's
        new components (other rows of the arrays are zero). The device state is left as it is."""
        n = self.num_bodies()
        r = np.ascontiguousarray(radius, dtype=np.float64)
        rho = np.full(n, BODY_DENSITY) if density is None else np.ascontiguousarray(density, dtype=np.float64)
        if r.shape != (n,) or rho.shape != (n,):
            raise ValueError("radius and density must have one entry per body")
        alive, changed = np.ones(n, np.uint8), np.zeros(n, np.uint8)
        pos, vel = np.zeros((n, 2)), np.zeros((n, 2))
        mass, pinned, rout = np.zeros(n, np.float32), np.zeros(n, np.uint8), np.zeros(n)
        k = C.c_int64()
        u8 = C.POINTER(C.c_uint8)
        _check(self._L.nb_resolve_collisions(self._h, r.ctypes.data_as(_dp), rho.ctypes.data_as(_dp), C.byref(k),
                                             alive.ctypes.data_as(u8), changed.ctypes.data_as(u8), pos.ctypes.data_as(_dp),
                                             vel.ctypes.data_as(_dp), mass.ctypes.data_as(C.POINTER(C.c_float)),
                                             pinned.ctypes.data_as(u8), rout.ctypes.data_as(_dp), None, 0))
        if with_log:  # the merges in the reference's order, {survivor, absorbed}
            log = np.zeros((max(k.value, 1), 2), np.int64)
            k2 = C.c_int64()
            _check(self._L.nb_resolve_collisions(self._h, r.ctypes.data_as(_dp), rho.ctypes.data_as(_dp), C.byref(k2),
                                                 alive.ctypes.data_as(u8), changed.ctypes.data_as(u8), pos.ctypes.data_as(_dp),
                                                 vel.ctypes.data_as(_dp), mass.ctypes.data_as(C.POINTER(C.c_float)),
                                                 pinned.ctypes.data_as(u8), rout.ctypes.data_as(_dp),
                                                 log.ctypes.data_as(C.POINTER(C.c_int64)), log.shape[0]))
            return k.value, alive, changed, pos, vel, mass, pinned, rout, log[:k.value]
        return k.value, alive, changed, pos, vel, mass, pinned, rout

    def merge_info(self):
        v = (C.c_int64 * 4)()
        _check(self._L.nb_merge_info(self._h, v))
        return dict(merges=v[0], events=v[1], scanned=v[2], initial_pairs=v[3])

    def debug_cbrt(self, x):
        """Device restatement of glibc's cbrt (merge.cuh) on an array (test hook)."""
        x = np.ascontiguousarray(x, np.float64)
        out = np.zeros_like(x)
        _check(self._L.nb_debug_cbrt(self._h, x.ctypes.data_as(_dp), out.ctypes.data_as(_dp), x.size))
        return out

    def download_state(self):
        """(pos, vel, acc, acc_prev): the four DVec2 components of Components.hpp:10-21."""
        n = self.num_bodies()
        out = [np.zeros((n, self.dim)) for _ in range(4)]
        _check(self._L.nb_download_state(self._h, self.dim, *[a.ctypes.data_as(_dp) for a in out]))
        return tuple(out)

    def stream_order(self):
        """(order, generations_since_sort): the summation order of the FP32 path (nb_stream_order)."""
        order = np.zeros(self.num_bodies(), np.int32)
        age = C.c_int64()
        _check(self._L.nb_stream_order(self._h, order.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(age)))
        return order, age.value

    def partition_info(self):
        """nb_partition_info: dict(curve, partitions, uploads_since, sorts) of a (multi-GPU) context."""
        v = (C.c_int64 * 4)()
        self._L.nb_partition_info.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
        _check(self._L.nb_partition_info(self._h, v))
        return {"curve": bool(v[0]), "partitions": int(v[1]), "uploads_since": int(v[2]), "sorts": int(v[3])}

    def set_stream_order(self, order, generations_since_sort=0):
        order = np.ascontiguousarray(order, np.int32)
        if order.shape != (self.num_bodies(),):
            raise NbodyError(1, "set_stream_order: one entry per body")
        _check(self._L.nb_set_stream_order(self._h, order.ctypes.data_as(C.POINTER(C.c_int32)), int(generations_since_sort)))

    def num_bodies(self):
        return int(self._L.nb_num_bodies(self._h))

    def local_range(self):
        lo, hi = C.c_int64(), C.c_int64()
        _check(self._L.nb_local_range(self._h, C.byref(lo), C.byref(hi)))
        return lo.value, hi.value

    # ---- reference statics ---------------------------------------------------------------
    def compute_gravity(self):
        """Physics::compute_gravity, Physics.hpp:183-261 (direct-sum branch)."""
        _check(self._L.nb_compute_gravity(self._h))

    def integrate(self, dt):
        """Physics::integrate, Physics.hpp:263-327."""
        _check(self._L.nb_integrate(self._h, C.c_float(dt)))

    def compute_diagnostics(self):
        """Physics::compute_diagnostics, Physics.hpp:117-177."""
        d = _Diag()
        _check(self._L.nb_compute_diagnostics(self._h, C.byref(d)))
        return Diagnostics(d)

    def zero_net_momentum(self):
        """Physics::zero_net_momentum, Physics.hpp:66-85."""
        _check(self._L.nb_zero_net_momentum(self._h))

    def reset_scenario(self):
        """Physics::reset_scenario, Physics.hpp:87-115 (three-body seed)."""
        _check(self._L.nb_reset_scenario(self._h))
        self.dim = 2

    def step(self, dt, n_frames=1, diagnostics=False):
        """n_frames x (gravity system + integrate system), Physics.hpp:45-57 — fused path."""
        fn = self._L.nb_step_with_diagnostics if diagnostics else self._L.nb_step
        _check(fn(self._h, C.c_float(dt), n_frames))

    def step_diagnostics(self):
        d = _Diag()
        _check(self._L.nb_step_diagnostics(self._h, C.byref(d)))
        return Diagnostics(d)

    # ---- measurement -------------------------------------------------------------------
    def synchronize(self):
        _check(self._L.nb_synchronize(self._h))

    def timer_begin(self):
        _check(self._L.nb_timer_begin(self._h))

    def timer_end(self):
        ms = C.c_float()
        _check(self._L.nb_timer_end(self._h, C.byref(ms)))
        return ms.value

    def counters(self):
        k, f, x = C.c_int64(), C.c_int64(), C.c_double()
        _check(self._L.nb_counters(self._h, C.byref(k), C.byref(f), C.byref(x)))
        return dict(kernel_launches=k.value, force_evals=f.value, interactions=x.value)

    def overlap_profile(self, enable=True):
        """{own_ms, comm_start_ms, comm_end_ms, comm_ms} of the last sharded Verlet step (nb_overlap_profile)."""
        v = (C.c_float * 4)()
        _check(self._L.nb_overlap_profile(self._h, int(enable), v))
        return dict(own_launch_ms=v[0], exchange_start_ms=v[1], exchange_end_ms=v[2], exchange_ms=v[3])

    def near_stats(self, enable=True):
        """(warp x group visits, near-path visits) of the force launches since the last call."""
        a, b = C.c_int64(), C.c_int64()
        _check(self._L.nb_near_stats(self._h, int(enable), C.byref(a), C.byref(b)))
        return a.value, b.value

    def profile_forces(self, enable=True):
        _check(self._L.nb_profile_forces(self._h, int(enable)))

    def profile_read(self):
        n, ms = C.c_int64(), C.c_double()
        _check(self._L.nb_profile_read(self._h, C.byref(n), C.byref(ms)))
        return n.value, ms.value
