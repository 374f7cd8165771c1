"""epp_b200 — Python host side of the B200-native validity-check path.

Thin ctypes binding over the C ABI of ``libepp_b200.so`` (``include/epp_c_api.h``) plus a
``World`` class that mirrors the reference ``World`` (``include/World.h:16-127`` of
Autonomous-Drone-Racing-Lab/Efficient-Path-Planner): same method names, argument meaning and
error behaviour, with batched entry points added.  PyTorch is used only for device memory,
streams and ``torch.distributed``; all checks run in the hand-written sm_100a kernels.  There
is no CPU fallback: if the library is missing or no GPU is present the calls raise.

The directory name contains a hyphen, so import it through the repo-root shim:
``import epp_b200``.
"""
import ctypes as C
import json
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("EPP_B200_LIB") or os.path.join(_HERE, "libepp_b200.so")  # override: profiling builds
DEFAULT_CONFIG = os.path.join(_HERE, "config", "config.json")

BP_AUTO, BP_BRUTE, BP_BITS, BP_GRID = 0, 1, 2, 3
BP_NAMES = {BP_AUTO: "auto", BP_BRUTE: "brute", BP_BITS: "bits", BP_GRID: "grid"}


class EppError(RuntimeError):
    """Raised for every non-zero epp_status (message = epp_last_error())."""

    def __init__(self, status, message):
        super().__init__(message)
        self.status = status


class ObbDesc(C.Structure):
    _fields_ = [("position", C.c_double * 3), ("size", C.c_double * 3), ("filling", C.c_int)]


class Obb(C.Structure):
    _fields_ = [("center", C.c_double * 3), ("half", C.c_double * 3), ("rot", C.c_double * 9),
                ("aabb_min", C.c_double * 3), ("aabb_max", C.c_double * 3), ("inflate", C.c_double),
                ("filling", C.c_int), ("is_gate", C.c_int), ("object_id", C.c_int),
                ("obb_index", C.c_int)]


_dp = C.POINTER(C.c_double)
_up = C.POINTER(C.c_uint32)
_vp = C.c_void_p

# name -> (restype, argtypes); every symbol include/epp_c_api.h declares.
SIGNATURES = {
    "epp_last_error": (C.c_char_p, []),
    "epp_version": (C.c_char_p, []),
    "epp_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "epp_world_create": (C.c_int, [C.c_double, C.c_double, C.POINTER(_vp)]),
    "epp_world_destroy": (C.c_int, [_vp]),
    "epp_world_set_gate_geometry": (C.c_int, [_vp, C.c_int, C.POINTER(ObbDesc), C.c_int]),
    "epp_world_set_obstacle_geometry": (C.c_int, [_vp, C.POINTER(ObbDesc), C.c_int]),
    "epp_world_reset": (C.c_int, [_vp]),
    "epp_world_add_gate": (C.c_int, [_vp, C.c_int, _dp]),
    "epp_world_update_gate": (C.c_int, [_vp, C.c_int, _dp]),
    "epp_world_add_obstacle": (C.c_int, [_vp, C.c_int, _dp]),
    "epp_world_parse_gates_and_obstacles": (C.c_int, [_vp, _dp, C.c_int, _dp, C.c_int]),
    "epp_world_size": (C.c_int, [_vp, C.POINTER(C.c_int)]),
    "epp_world_get_obb": (C.c_int, [_vp, C.c_int, C.POINTER(Obb)]),
    "epp_world_set_broadphase": (C.c_int, [_vp, C.c_int]),
    "epp_world_commit": (C.c_int, [_vp]),
    "epp_world_epoch": (C.c_int, [_vp, C.POINTER(C.c_uint64)]),
    "epp_world_active_broadphase": (C.c_int, [_vp, C.POINTER(C.c_int)]),
    "epp_check_states": (C.c_int, [_vp, _vp, C.c_size_t, C.c_int, _vp]),
    "epp_check_states_mindist": (C.c_int, [_vp, _vp, C.c_size_t, C.c_double, _vp]),
    "epp_check_segments": (C.c_int, [_vp, _vp, _vp, C.c_size_t, C.c_int, _vp]),
    "epp_is_valid": (C.c_int, [_vp, _dp, C.c_int, C.POINTER(C.c_int)]),
    "epp_check_motion": (C.c_int, [_vp, _dp, _dp, C.c_int, C.POINTER(C.c_int)]),
    "epp_check_trajectories": (C.c_int, [_vp, _vp, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t,
                                         C.c_int, C.c_double, _vp, _vp, _vp]),
    "epp_check_poly_trajectories": (C.c_int, [_vp, _vp, _vp, _vp, C.c_size_t, C.c_int, C.c_int,
                                              C.c_double, C.c_size_t, _vp, _vp, _vp, _vp]),
    "epp_check_states_dev": (C.c_int, [_vp, _vp, C.c_size_t, C.c_int, _vp, _vp]),
    "epp_check_states_mindist_dev": (C.c_int, [_vp, _vp, C.c_size_t, C.c_double, _vp, _vp]),
    "epp_check_segments_dev": (C.c_int, [_vp, _vp, _vp, C.c_size_t, C.c_int, _vp, _vp]),
    "epp_check_trajectories_dev": (C.c_int, [_vp, _vp, C.c_size_t, C.c_size_t, C.c_size_t,
                                             C.c_size_t, C.c_int, C.c_double, _vp, _vp, _vp, _vp]),
    "epp_check_poly_trajectories_dev": (C.c_int, [_vp, _vp, _vp, _vp, C.c_size_t, C.c_int, C.c_int,
                                                  C.c_double, C.c_size_t, _vp, _vp, _vp, _vp, _vp]),
    "epp_check_poly_trajectories_static_dev": (C.c_int, [_vp, _vp, _vp, _vp, C.c_size_t, C.c_int, C.c_int,
                                                         C.c_double, C.c_size_t, _vp, _vp, _vp, _vp, _vp]),
    "epp_shard_range": (C.c_int, [C.c_size_t, C.c_int, C.c_int, C.c_size_t,
                                  C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "epp_mask_summary_dev": (C.c_int, [_vp, C.c_size_t, C.c_uint64, _vp, _vp]),
    "epp_measure_fp64_nofma": (C.c_int, [C.c_int, C.c_int, _vp, _dp, _dp]),
    "epp_check_states_multi_dev": (C.c_int, [_vp, _vp, _vp, _vp, C.c_int, C.c_int, _vp]),
    "epp_check_segments_multi_dev": (C.c_int, [_vp, _vp, _vp, _vp, _vp, C.c_int, C.c_int, _vp]),
    "epp_gather_set_deferred": (C.c_int, [_vp, C.c_int]),
    "epp_gather_flush": (C.c_int, [_vp, _vp]),
    "epp_measure_single_latency": (C.c_int, [_vp, C.c_int, _dp]),
    "epp_count_candidates_dev": (C.c_int, [_vp, _vp, _vp, C.c_size_t, _vp, _vp]),
    "epp_gen_states_dev": (C.c_int, [C.c_uint64, C.c_uint64, C.c_size_t, _dp, _dp, _vp, _vp]),
    "epp_gen_segments_dev": (C.c_int, [C.c_uint64, C.c_uint64, C.c_size_t, _dp, _dp, C.c_double,
                                       _vp, _vp, _vp]),
    "epp_gather_create": (C.c_int, [C.c_int, C.c_int, C.c_size_t, C.c_int, C.POINTER(_vp)]),
    "epp_gather_destroy": (C.c_int, [_vp]),
    "epp_gather_export": (C.c_int, [_vp, _vp]),
    "epp_gather_connect": (C.c_int, [_vp, _vp]),
    "epp_gather_connect_local": (C.c_int, [C.POINTER(_vp), C.c_int]),
    "epp_gather_buffer": (C.c_int, [_vp, C.c_int, C.POINTER(_vp)]),
    "epp_gather_set_timeout_ms": (C.c_int, [_vp, C.c_uint]),
    "epp_gather_status": (C.c_int, [_vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "epp_check_states_gather_dev": (C.c_int, [_vp, _vp, C.c_size_t, C.c_int, _vp, C.c_int, _vp]),
    "epp_check_states_mindist_gather_dev": (C.c_int, [_vp, _vp, C.c_size_t, C.c_double, _vp, C.c_int, _vp]),
    "epp_check_segments_gather_dev": (C.c_int, [_vp, _vp, _vp, C.c_size_t, C.c_int, _vp, C.c_int, _vp]),
    "epp_segment_times": (C.c_int, [_vp, C.c_int, C.c_double, C.c_double, _vp]),
    "epp_minsnap": (C.c_int, [_vp, _vp, _vp, C.c_size_t, C.c_int, _vp]),
    "epp_minsnap_dev": (C.c_int, [_vp, _vp, _vp, C.c_size_t, C.c_int, _vp, _vp]),
    "epp_sample_trajectories": (C.c_int, [_vp, _vp, _vp, _vp, C.c_size_t, C.c_int, C.c_int, C.c_size_t, _vp, _vp]),
    "epp_sample_trajectories_dev": (C.c_int, [_vp, _vp, _vp, _vp, C.c_size_t, C.c_int, C.c_int, C.c_size_t, _vp,
                                              _vp, _vp]),
    "epp_kernel_launch_count": (C.c_uint64, []),
    "epp_set_tuning": (C.c_int, [C.c_char_p, C.c_int]),
    "epp_get_tuning": (C.c_int, [C.c_char_p, C.POINTER(C.c_int)]),
}

_lib = None


def lib():
    """Load libepp_b200.so (built in-tree by __graft_entry__.build()).  No fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise EppError(-1, "libepp_b200.so is not built (run `python __graft_entry__.py build` "
                               "or `make -C efficient-path-planner_b200/csrc`); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
        # EPP_TUNE="knob=value,knob=value": tuning knobs for a whole process (A/B runs of the test suite and of bench.py)
        for item in filter(None, os.environ.get("EPP_TUNE", "").split(",")):
            key, _, value = item.partition("=")
            if L.epp_set_tuning(key.strip().encode(), int(value)) != 0:
                raise EppError(-1, "EPP_TUNE: unknown knob %r" % key)
    return _lib


def _chk(status):
    if status != 0:
        raise EppError(status, lib().epp_last_error().decode())


def device_count():
    n = C.c_int(0)
    _chk(lib().epp_device_count(C.byref(n)))
    return n.value


def kernel_launch_count():
    return int(lib().epp_kernel_launch_count())


def set_tuning(key, value):
    _chk(lib().epp_set_tuning(key.encode(), int(value)))


def get_tuning(key):
    v = C.c_int(0)
    _chk(lib().epp_get_tuning(key.encode(), C.byref(v)))
    return v.value


def shard_range(n, rank, world_size, align=1024):
    """Contiguous unit range of `rank`: boundaries on whole mask words."""
    b, e = C.c_size_t(0), C.c_size_t(0)
    _chk(lib().epp_shard_range(n, rank, world_size, align, C.byref(b), C.byref(e)))
    return b.value, e.value


def _f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None:
        a = a.reshape(shape)
    return a


def _ptr(a):
    return a.ctypes.data if a is not None else None


def unpack_mask(mask, n):
    """uint32 mask words -> bool[n] (bit i&31 of word i>>5)."""
    bits = np.unpackbits(np.ascontiguousarray(mask, dtype="<u4").view(np.uint8), bitorder="little")
    return bits[:n].astype(bool)


def mask_words(n):
    return (n + 31) // 32


def descs_from_component(component):
    """config component dict (file order, src/ConfigParserYAML.cpp:54-73) -> ObbDesc array."""
    arr = (ObbDesc * len(component))()
    for i, (_, o) in enumerate(component.items()):
        arr[i].position[:] = [float(v) for v in o["position"]]
        arr[i].size[:] = [float(v) for v in o["size"]]
        arr[i].filling = 1 if o["type"] == "filling" else 0
    return arr


def load_config(path=None):
    with open(path or DEFAULT_CONFIG) as f:
        return json.load(f)


class World:
    """GPU-resident World.  Mirrors include/World.h of the reference.

    ``World(config)`` takes the parsed config dict (or a path) where the reference takes a
    ``ConfigParser``.  Mutators mark the world dirty; the next check commits a new device
    epoch automatically (the reference mutates its R-tree in place).
    """

    def __init__(self, config=None, broadphase=BP_AUTO):
        if config is None or isinstance(config, str):
            config = load_config(config)
        self.config = config
        L = lib()
        infl = config["world_properties"]["inflate_radius"]  # include/World.h:22-23
        h = _vp()
        _chk(L.epp_world_create(float(infl["gate"]), float(infl["obstacle"]), C.byref(h)))
        self._h = h
        for tid, name in config["gate_id_to_name_mapping"].items():
            d = descs_from_component(config["component_geometry"][name])
            _chk(L.epp_world_set_gate_geometry(self._h, int(tid), d, len(d)))
        d = descs_from_component(config["component_geometry"]["obstacle"])
        _chk(L.epp_world_set_obstacle_geometry(self._h, d, len(d)))
        if broadphase != BP_AUTO:
            self.set_broadphase(broadphase)
        self._dirty = True

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                lib().epp_world_destroy(h)
            except Exception:
                pass
            self._h = None

    # -- reference World API (include/World.h:29-92) -------------------------------------
    def resetWorld(self):
        _chk(lib().epp_world_reset(self._h))
        self._dirty = True

    def addGate(self, gateId, coordinates):
        c = _f64(coordinates, 7)
        _chk(lib().epp_world_add_gate(self._h, int(gateId), c.ctypes.data_as(_dp)))
        self._dirty = True

    def updateGatePosition(self, gateId, coordinates):
        c = _f64(coordinates, 7)
        _chk(lib().epp_world_update_gate(self._h, int(gateId), c.ctypes.data_as(_dp)))
        self._dirty = True

    def addObstacle(self, obstacleId, coordinates):
        c = _f64(coordinates, 6)
        _chk(lib().epp_world_add_obstacle(self._h, int(obstacleId), c.ctypes.data_as(_dp)))
        self._dirty = True

    def checkPointValidity(self, point, arg=False):
        """Both reference overloads: bool -> canPassGate, float -> minDistance
        (src/World.cpp:80-128)."""
        p = _f64(point, (1, 3))
        if isinstance(arg, (bool, np.bool_)):
            self._sync()
            v = C.c_int(0)
            _chk(lib().epp_is_valid(self._h, p.ctypes.data_as(_dp), int(arg), C.byref(v)))
            return bool(v.value)
        return bool(unpack_mask(self.check_states_mindist(p, float(arg)), 1)[0])

    def checkRayValid(self, start, end, canPassGate=False):
        self._sync()
        a, b = _f64(start, 3), _f64(end, 3)
        v = C.c_int(0)
        _chk(lib().epp_check_motion(self._h, a.ctypes.data_as(_dp), b.ctypes.data_as(_dp),
                                    int(canPassGate), C.byref(v)))
        return bool(v.value)

    # -- PathPlanner::parseGatesAndObstacles (src/PathPlanner.cpp:60-78) -----------------
    def parse_gates_and_obstacles(self, gates, obstacles):
        g = _f64(gates, (-1, 7)) if len(gates) else np.zeros((0, 7))
        o = _f64(obstacles, (-1, 6)) if len(obstacles) else np.zeros((0, 6))
        _chk(lib().epp_world_parse_gates_and_obstacles(self._h, g.ctypes.data_as(_dp), g.shape[0],
                                                       o.ctypes.data_as(_dp), o.shape[0]))
        self._dirty = False

    # -- housekeeping ---------------------------------------------------------------------
    def set_broadphase(self, mode):
        _chk(lib().epp_world_set_broadphase(self._h, int(mode)))
        self._dirty = True

    def commit(self):
        _chk(lib().epp_world_commit(self._h))
        self._dirty = False

    def _sync(self):
        if self._dirty:
            self.commit()

    @property
    def handle(self):
        self._sync()
        return self._h

    def size(self):
        n = C.c_int(0)
        _chk(lib().epp_world_size(self._h, C.byref(n)))
        return n.value

    def epoch(self):
        e = C.c_uint64(0)
        _chk(lib().epp_world_epoch(self._h, C.byref(e)))
        return e.value

    def active_broadphase(self):
        self._sync()
        m = C.c_int(0)
        _chk(lib().epp_world_active_broadphase(self._h, C.byref(m)))
        return m.value

    def obbs(self):
        out = []
        for i in range(self.size()):
            o = Obb()
            _chk(lib().epp_world_get_obb(self._h, i, C.byref(o)))
            out.append(dict(center=list(o.center), half=list(o.half), rot=list(o.rot),
                            aabb_min=list(o.aabb_min), aabb_max=list(o.aabb_max), inflate=o.inflate,
                            filling=o.filling, is_gate=o.is_gate, object_id=o.object_id,
                            obb_index=o.obb_index))
        return out

    # -- batched checks on host (numpy) buffers --------------------------------------------
    def check_states(self, xyz, can_pass_gate=False):
        """(n,3) float64 -> uint32 mask words (bit = valid)."""
        self._sync()
        xyz = _f64(xyz, (-1, 3))
        n = xyz.shape[0]
        mask = np.zeros(mask_words(n), dtype=np.uint32)
        _chk(lib().epp_check_states(self._h, _ptr(xyz), n, int(can_pass_gate), _ptr(mask)))
        return mask

    def check_states_mindist(self, xyz, min_distance):
        self._sync()
        xyz = _f64(xyz, (-1, 3))
        n = xyz.shape[0]
        mask = np.zeros(mask_words(n), dtype=np.uint32)
        _chk(lib().epp_check_states_mindist(self._h, _ptr(xyz), n, float(min_distance), _ptr(mask)))
        return mask

    def check_segments(self, a, b, can_pass_gate=False):
        self._sync()
        a, b = _f64(a, (-1, 3)), _f64(b, (-1, 3))
        if a.shape != b.shape:
            raise ValueError("start/end batches differ in shape")
        n = a.shape[0]
        mask = np.zeros(mask_words(n), dtype=np.uint32)
        _chk(lib().epp_check_segments(self._h, _ptr(a), _ptr(b), n, int(can_pass_gate), _ptr(mask)))
        return mask

    def check_trajectories(self, trajs, min_distance, want_samples=True):
        """trajs: (n_traj, rows, cols) array (each slice one trajectory matrix, evaluated as the
        reference's column-major Eigen matrix) -> (traj_valid words, sample masks, first_invalid)."""
        self._sync()
        t = np.asarray(trajs, dtype=np.float64)
        if t.ndim == 2:
            t = t[None]
        n_traj, rows, cols = t.shape
        cm = np.ascontiguousarray(np.transpose(t, (0, 2, 1)))  # per trajectory: cols x rows = column-major
        valid = np.zeros(mask_words(n_traj), dtype=np.uint32)
        smask = np.zeros((n_traj, mask_words(rows)), dtype=np.uint32) if want_samples else None
        first = np.zeros(n_traj, dtype=np.int64)
        _chk(lib().epp_check_trajectories(self._h, _ptr(cm), n_traj, rows, rows, rows * cols, cols,
                                          float(min_distance), _ptr(valid), _ptr(smask), _ptr(first)))
        return valid, smask, first

    def checkTrajectoryValidity(self, trajectory, minDistance):
        """PathPlanner::checkTrajectoryValidity (src/PathPlanner.cpp:267-280)."""
        valid, _, _ = self.check_trajectories(np.asarray(trajectory)[None], minDistance, False)
        return bool(valid[0] & 1)

    def check_poly_trajectories(self, coeffs, durations, dt, min_distance, max_samples=None,
                                want_samples=True):
        """coeffs (n_traj, K, 3, N), durations (n_traj, K), dt scalar or (n_traj,).  max_samples (rows reserved per
        trajectory) defaults to what the longest trajectory needs; a trajectory that would produce more rows than
        max_samples is reported invalid with first row = max_samples, never silently valid."""
        self._sync()
        c = _f64(coeffs)
        d = _f64(durations)
        if c.ndim == 3:
            c, d = c[None], d[None]
        n_traj, K, _, N = c.shape
        dts = np.ascontiguousarray(np.broadcast_to(np.asarray(dt, dtype=np.float64), (n_traj,)))
        if max_samples is None:
            max_samples = poly_max_samples(d, dts)
        max_samples = (int(max_samples) + 31) // 32 * 32
        valid = np.zeros(mask_words(n_traj), dtype=np.uint32)
        smask = np.zeros((n_traj, max_samples // 32), dtype=np.uint32) if want_samples else None
        ns = np.zeros(n_traj, dtype=np.uint32)
        first = np.zeros(n_traj, dtype=np.int64)
        _chk(lib().epp_check_poly_trajectories(self._h, _ptr(c), _ptr(d), _ptr(dts), n_traj, K, N,
                                               float(min_distance), max_samples, _ptr(valid),
                                               _ptr(smask), _ptr(ns), _ptr(first)))
        return valid, smask, ns, first

    # -- batched checks on device (torch) buffers; asynchronous on the current stream -------
    @staticmethod
    def _stream():
        import torch
        return torch.cuda.current_stream().cuda_stream

    def check_states_dev(self, xyz, mask=None, can_pass_gate=False, min_distance=None):
        import torch
        self._sync()
        n = xyz.numel() // 3
        if mask is None:
            mask = torch.empty(mask_words(n), dtype=torch.int32, device=xyz.device)
        if min_distance is None:
            _chk(lib().epp_check_states_dev(self._h, xyz.data_ptr(), n, int(can_pass_gate),
                                            mask.data_ptr(), self._stream()))
        else:
            _chk(lib().epp_check_states_mindist_dev(self._h, xyz.data_ptr(), n, float(min_distance),
                                                    mask.data_ptr(), self._stream()))
        return mask

    def check_states_multi_dev(self, batches, masks, can_pass_gate=False):
        """Independent state batches (device tensors) -> their mask tensors, one library call (epp_check_states_multi_dev)."""
        self._sync()
        k = len(batches)
        ptrs = (C.c_void_p * k)(*[b.data_ptr() for b in batches])
        ns = (C.c_size_t * k)(*[b.numel() // 3 for b in batches])
        ms = (C.c_void_p * k)(*[m.data_ptr() for m in masks])
        _chk(lib().epp_check_states_multi_dev(self._h, ptrs, ns, ms, k, int(can_pass_gate), self._stream()))
        return masks

    def check_segments_dev(self, a, b, mask=None, can_pass_gate=False):
        import torch
        self._sync()
        n = a.numel() // 3
        if mask is None:
            mask = torch.empty(mask_words(n), dtype=torch.int32, device=a.device)
        _chk(lib().epp_check_segments_dev(self._h, a.data_ptr(), b.data_ptr(), n, int(can_pass_gate),
                                          mask.data_ptr(), self._stream()))
        return mask

    def check_segments_multi_dev(self, batches, masks, can_pass_gate=False):
        """Independent segment batches [(a, b), ...] (device tensors) -> their mask tensors, one library call
        (epp_check_segments_multi_dev)."""
        self._sync()
        k = len(batches)
        pa = (C.c_void_p * k)(*[a.data_ptr() for a, _ in batches])
        pb = (C.c_void_p * k)(*[b.data_ptr() for _, b in batches])
        ns = (C.c_size_t * k)(*[a.numel() // 3 for a, _ in batches])
        ms = (C.c_void_p * k)(*[m.data_ptr() for m in masks])
        _chk(lib().epp_check_segments_multi_dev(self._h, pa, pb, ns, ms, k, int(can_pass_gate), self._stream()))
        return masks

    def single_latency_us(self, n_calls=2000):
        """Microseconds per epp_is_valid call, timed inside the library (no ctypes overhead)."""
        self._sync()
        us = C.c_double(0.0)
        _chk(lib().epp_measure_single_latency(self._h, int(n_calls), C.byref(us)))
        return us.value

    def count_candidates_dev(self, a, b=None):
        """(mean broad-phase candidates visited, mean candidates passing the exact stored-AABB test) per unit of a
        device batch of points (b None) or segments."""
        import torch
        self._sync()
        n = a.numel() // 3
        out = torch.zeros(2, dtype=torch.int64, device=a.device)
        _chk(lib().epp_count_candidates_dev(self._h, a.data_ptr(), b.data_ptr() if b is not None else None, n,
                                            out.data_ptr(), self._stream()))
        v, h = out.cpu().numpy().view(np.uint64)
        return float(v) / max(n, 1), float(h) / max(n, 1)

    def check_trajectories_dev(self, mats, n_traj, rows, ld, traj_stride, cols, min_distance,
                               want_samples=True):
        import torch
        self._sync()
        dev = mats.device
        valid = torch.empty(mask_words(n_traj), dtype=torch.int32, device=dev)
        smask = torch.empty((n_traj, mask_words(rows)), dtype=torch.int32, device=dev) if want_samples else None
        first = torch.empty(n_traj, dtype=torch.int64, device=dev)
        _chk(lib().epp_check_trajectories_dev(self._h, mats.data_ptr(), n_traj, rows, ld, traj_stride,
                                              cols, float(min_distance), valid.data_ptr(),
                                              smask.data_ptr() if want_samples else None,
                                              first.data_ptr(), self._stream()))
        return valid, smask, first

    def check_poly_trajectories_dev(self, coeffs, durations, dt, min_distance, max_samples=None,
                                    want_samples=True):
        import torch
        self._sync()
        n_traj, K, _, N = coeffs.shape
        dev = coeffs.device
        if max_samples is None:  # one device -> host read; pass max_samples to stay asynchronous
            max_samples = poly_max_samples(durations, dt)
        max_samples = (int(max_samples) + 31) // 32 * 32
        valid = torch.empty(mask_words(n_traj), dtype=torch.int32, device=dev)
        smask = torch.empty((n_traj, max_samples // 32), dtype=torch.int32, device=dev) if want_samples else None
        ns = torch.empty(n_traj, dtype=torch.int32, device=dev)
        first = torch.empty(n_traj, dtype=torch.int64, device=dev)
        _chk(lib().epp_check_poly_trajectories_dev(self._h, coeffs.data_ptr(), durations.data_ptr(),
                                                   dt.data_ptr(), n_traj, K, N, float(min_distance),
                                                   max_samples, valid.data_ptr(),
                                                   smask.data_ptr() if want_samples else None,
                                                   ns.data_ptr(), first.data_ptr(), self._stream()))
        return valid, smask, ns, first

    def check_poly_trajectories_static_dev(self, coeffs, durations, dt, min_distance, max_samples, valid, smask=None,
                                           ns=None, first=None):
        """Allocation-free form (include/epp_c_api.h epp_check_poly_trajectories_static_dev; no counterpart in the
        reference): every buffer is the caller's, the kernels' per-trajectory scratch is a static device array of
        STATIC_MAX_TRAJECTORIES words inside the library.  `max_samples` must be a multiple of 32; `valid` is
        mask_words(n_traj) int32, `smask` (n_traj, max_samples / 32) int32, `ns` n_traj int32, `first` n_traj int64
        (the last three optional).  Calls that overlap on the GPU race on the scratch array."""
        if self._dirty:
            raise EppError(7, "commit() the world first: this entry point allocates nothing, a commit does")
        n_traj, K, _, N = coeffs.shape
        _chk(lib().epp_check_poly_trajectories_static_dev(
            self._h, coeffs.data_ptr(), durations.data_ptr(), dt.data_ptr(), n_traj, K, N, float(min_distance),
            int(max_samples), valid.data_ptr(), None if smask is None else smask.data_ptr(),
            None if ns is None else ns.data_ptr(), None if first is None else first.data_ptr(), self._stream()))


STATIC_MAX_TRAJECTORIES = 65536  # EPP_STATIC_MAX_TRAJECTORIES


class _DevWords:
    """__cuda_array_interface__ view of library-owned device words (lets torch wrap them without a copy)."""

    def __init__(self, ptr, shape, owner):
        self.owner = owner
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<i4", "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class Gather:
    """Mask exchange fused into the check kernels (include/epp_c_api.h "Mask exchange").

    One per rank.  ``connect()`` exchanges the CUDA IPC handles over ``torch.distributed`` (one process per
    GPU); ``Gather.connect_local(list)`` wires several handles of ONE process together.  After
    ``check_states(world, xyz, buffer)`` has completed in stream order, ``tensor(buffer)`` — a
    (world_size, slot_words) int32 view of library memory — holds the masks of all ranks."""

    def __init__(self, rank, world_size, slot_words, n_buffers=2):
        self.rank, self.world_size, self.slot_words, self.n_buffers = rank, world_size, slot_words, n_buffers
        h = _vp()
        _chk(lib().epp_gather_create(rank, world_size, slot_words, n_buffers, C.byref(h)))
        self._h = h

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                lib().epp_gather_destroy(h)
            except Exception:
                pass

    def export(self):
        buf = (C.c_ubyte * 64)()
        _chk(lib().epp_gather_export(self._h, buf))
        return bytes(buf)

    def connect(self, group=None):
        """Collective over torch.distributed: all-gather the 64-byte handles, open the peers'."""
        import torch.distributed as dist
        if self.world_size == 1:
            return self
        handles = [None] * self.world_size
        dist.all_gather_object(handles, self.export(), group=group)
        blob = b"".join(handles)
        _chk(lib().epp_gather_connect(self._h, C.c_char_p(blob)))
        dist.barrier(group=group)
        return self

    @staticmethod
    def connect_local(gathers):
        arr = (_vp * len(gathers))(*[g._h for g in gathers])
        _chk(lib().epp_gather_connect_local(arr, len(gathers)))

    def set_timeout_ms(self, ms):
        _chk(lib().epp_gather_set_timeout_ms(self._h, int(ms)))

    def set_deferred(self, on=True):
        """Deferred form: a call unpacks the peers' words of the previous call; tensor(b) of call s is complete after
        call s + 1 or flush() (include/epp_c_api.h)."""
        _chk(lib().epp_gather_set_deferred(self._h, int(bool(on))))
        return self

    def flush(self):
        import torch
        _chk(lib().epp_gather_flush(self._h, torch.cuda.current_stream().cuda_stream))

    def status(self):
        calls, timeouts = C.c_uint64(0), C.c_uint64(0)
        _chk(lib().epp_gather_status(self._h, C.byref(calls), C.byref(timeouts)))
        return calls.value, timeouts.value

    def tensor(self, buffer, device=None):
        import torch
        p = _vp()
        _chk(lib().epp_gather_buffer(self._h, buffer, C.byref(p)))
        dev = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        return torch.as_tensor(_DevWords(p.value, (self.world_size, self.slot_words), self), device=dev)

    def check_states(self, world, xyz, buffer, can_pass_gate=False, min_distance=None):
        world._sync()
        n = xyz.numel() // 3
        if min_distance is None:
            _chk(lib().epp_check_states_gather_dev(world._h, xyz.data_ptr(), n, int(can_pass_gate), self._h,
                                                   buffer, World._stream()))
        else:
            _chk(lib().epp_check_states_mindist_gather_dev(world._h, xyz.data_ptr(), n, float(min_distance),
                                                           self._h, buffer, World._stream()))

    def check_segments(self, world, a, b, buffer, can_pass_gate=False):
        world._sync()
        _chk(lib().epp_check_segments_gather_dev(world._h, a.data_ptr(), b.data_ptr(), a.numel() // 3,
                                                 int(can_pass_gate), self._h, buffer, World._stream()))


def segment_times(waypoints, v_max, a_max):
    """estimateSegmentTimesNfabian (external/poly_traj/src/vertex.cpp:272-289): (..., V, 3) -> (..., V - 1)."""
    w = _f64(waypoints)
    lead, V = w.shape[:-2], w.shape[-2]
    w2 = w.reshape(-1, V, 3)
    out = np.zeros((w2.shape[0], V - 1))
    for i in range(w2.shape[0]):
        _chk(lib().epp_segment_times(_ptr(w2[i]), V, float(v_max), float(a_max), _ptr(out[i])))
    return out.reshape(lead + (V - 1,))


def minsnap(waypoints, durations, start_vel_acc=None):
    """Batched linear min-snap solve on the GPU: waypoints (n, K+1, 3), durations (n, K), optional start
    velocity / acceleration (n, 6) -> coefficients (n, K, 3, 10) in check_poly_trajectories' layout."""
    w, d = _f64(waypoints), _f64(durations)
    n, K = d.shape
    va = _f64(start_vel_acc, (n, 6)) if start_vel_acc is not None else None
    out = np.zeros((n, K, 3, 10))
    _chk(lib().epp_minsnap(_ptr(w), _ptr(d), _ptr(va) if va is not None else None, n, K, _ptr(out)))
    return out


def minsnap_dev(waypoints, durations, start_vel_acc=None, out=None):
    import torch
    n, K = durations.shape
    if out is None:
        out = torch.empty((n, K, 3, 10), dtype=torch.float64, device=durations.device)
    _chk(lib().epp_minsnap_dev(waypoints.data_ptr(), durations.data_ptr(),
                               start_vel_acc.data_ptr() if start_vel_acc is not None else None, n, K,
                               out.data_ptr(), torch.cuda.current_stream().cuda_stream))
    return out


def sample_trajectories(coeffs, durations, dt, time_offsets=None, max_rows=1024):
    """The reference's trajectory matrices (poly_traj::generateTrajectory's sampling, trajectory_generator.cpp:60-97):
    coeffs (n, K, 3, N), durations (n, K), dt (n,) -> list of (rows_i, 10) arrays [px vx ax py vy ay pz vz az t]."""
    c, d = _f64(coeffs), _f64(durations)
    n, K, _, N = c.shape
    dts = np.ascontiguousarray(np.broadcast_to(np.asarray(dt, dtype=np.float64), (n,)))
    offs = _f64(np.broadcast_to(np.asarray(time_offsets, dtype=np.float64), (n,))) if time_offsets is not None else None
    out = np.zeros((n, 10, max_rows))
    rows = np.zeros(n, dtype=np.uint32)
    _chk(lib().epp_sample_trajectories(_ptr(c), _ptr(d), _ptr(dts), _ptr(offs), n, K, N, max_rows, _ptr(out), _ptr(rows)))
    if (rows > max_rows).any():
        raise EppError(6, "sample_trajectories: a trajectory has %d rows, max_rows is %d" % (rows.max(), max_rows))
    return [np.ascontiguousarray(out[i, :, : rows[i]].T) for i in range(n)]


def sample_trajectories_dev(coeffs, durations, dt, ld=1024, time_offsets=None):
    """Device form of sample_trajectories: torch tensors in, (n, 10, ld) column-major matrices + row counts out."""
    import torch
    n, K, _, N = coeffs.shape
    out = torch.zeros((n, 10, ld), dtype=torch.float64, device=coeffs.device)
    rows = torch.zeros(n, dtype=torch.int32, device=coeffs.device)
    _chk(lib().epp_sample_trajectories_dev(coeffs.data_ptr(), durations.data_ptr(), dt.data_ptr(),
                                           time_offsets.data_ptr() if time_offsets is not None else None, n, K, N, ld,
                                           out.data_ptr(), rows.data_ptr(), torch.cuda.current_stream().cuda_stream))
    return out, rows


def gen_states_dev(seed, first, n, lo, hi, out=None, device="cuda"):
    import torch
    if out is None:
        out = torch.empty((n, 3), dtype=torch.float64, device=device)
    lo, hi = _f64(lo, 3), _f64(hi, 3)
    _chk(lib().epp_gen_states_dev(seed, first, n, lo.ctypes.data_as(_dp), hi.ctypes.data_as(_dp),
                                  out.data_ptr(), torch.cuda.current_stream().cuda_stream))
    return out


def gen_segments_dev(seed, first, n, lo, hi, max_len, device="cuda"):
    import torch
    a = torch.empty((n, 3), dtype=torch.float64, device=device)
    b = torch.empty((n, 3), dtype=torch.float64, device=device)
    lo, hi = _f64(lo, 3), _f64(hi, 3)
    _chk(lib().epp_gen_segments_dev(seed, first, n, lo.ctypes.data_as(_dp), hi.ctypes.data_as(_dp),
                                    float(max_len), a.data_ptr(), b.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream))
    return a, b


def mask_summary_dev(mask, first_word=0):
    """(popcount, xor checksum) of a device mask; checksum is position dependent."""
    import torch
    out = torch.zeros(2, dtype=torch.int64, device=mask.device)
    _chk(lib().epp_mask_summary_dev(mask.data_ptr(), mask.numel(), first_word, out.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream))
    pop, x = out.cpu().numpy().view(np.uint64)
    return int(pop), int(x)


def poly_max_samples(durations, dt):
    """Rows to reserve per trajectory for evaluateRange(0, sum(durations), dt): the running sums of the sampler
    (trajectory.cpp:121-140) can emit a few rows more than sum / dt, hence the margin."""
    worst = float((durations.sum(1) / dt).max()) if durations.shape[0] else 0.0
    if not np.isfinite(worst):
        raise ValueError("dt must be positive and durations finite")
    return int(worst * (1.0 + 1e-9)) + 8


def measure_fp64_nofma(iters=4096, reps=5):
    """(thread-level non-fused FP64 instructions per second, ms per launch) of the current device."""
    import torch
    rate, ms = C.c_double(0.0), C.c_double(0.0)
    _chk(lib().epp_measure_fp64_nofma(int(iters), int(reps), torch.cuda.current_stream().cuda_stream,
                                      C.byref(rate), C.byref(ms)))
    return rate.value, ms.value


def _mix64(seed, counter):
    """numpy restatement of the device mix64 (for host-side checksum comparison)."""
    with np.errstate(over="ignore"):
        z = np.uint64(seed) + np.uint64(0x9E3779B97F4A7C15) * (counter.astype(np.uint64) + np.uint64(1))
        z ^= z >> np.uint64(30)
        z *= np.uint64(0xBF58476D1CE4E5B9)
        z ^= z >> np.uint64(27)
        z *= np.uint64(0x94D049BB133111EB)
        z ^= z >> np.uint64(31)
    return z


def mask_summary_host(mask, first_word=0):
    m = np.ascontiguousarray(mask, dtype=np.uint32)
    pop = int(np.unpackbits(m.view(np.uint8)).sum())
    idx = np.arange(first_word, first_word + m.size, dtype=np.uint64)
    h = _mix64(0x5EED, (idx << np.uint64(32)) | m.astype(np.uint64))
    x = int(np.bitwise_xor.reduce(h)) if m.size else 0
    return pop, x


from . import worlds  # noqa: E402,F401  (synthetic world / input builders)
