"""ctypes binding of the CPU oracle (oracle/libepp_oracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package never
imports this module.
"""
import ctypes as C
import json
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libepp_oracle.so")


class ObbDesc(C.Structure):
    _fields_ = [("position", C.c_double * 3), ("size", C.c_double * 3), ("filling", C.c_int)]


class Obb(C.Structure):
    _fields_ = [("center", C.c_double * 3), ("half", C.c_double * 3), ("rot", C.c_double * 9),
                ("aabb_min", C.c_double * 3), ("aabb_max", C.c_double * 3), ("inflate", C.c_double),
                ("filling", C.c_int), ("is_gate", C.c_int), ("object_id", C.c_int),
                ("obb_index", C.c_int)]


def build():
    subprocess.check_call(["make", "-s", "-C", HERE, "libepp_oracle.so"])


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        dp = C.POINTER(C.c_double)
        up = C.POINTER(C.c_uint32)
        L.orc_world_create.restype = C.c_void_p
        L.orc_world_create.argtypes = [C.c_double, C.c_double]
        L.orc_world_destroy.argtypes = [C.c_void_p]
        L.orc_world_set_gate_geometry.argtypes = [C.c_void_p, C.c_int, C.POINTER(ObbDesc), C.c_int]
        L.orc_world_set_obstacle_geometry.argtypes = [C.c_void_p, C.POINTER(ObbDesc), C.c_int]
        L.orc_world_reset.argtypes = [C.c_void_p]
        L.orc_world_add_gate.argtypes = [C.c_void_p, C.c_int, dp]
        L.orc_world_update_gate.argtypes = [C.c_void_p, C.c_int, dp]
        L.orc_world_add_obstacle.argtypes = [C.c_void_p, C.c_int, dp]
        L.orc_world_size.argtypes = [C.c_void_p]
        L.orc_world_get.argtypes = [C.c_void_p, C.c_int, C.POINTER(Obb)]
        L.orc_obb_point_hit.argtypes = [C.POINTER(Obb), dp, C.c_double]
        L.orc_obb_ray_hit.argtypes = [C.POINTER(Obb), dp, dp, C.c_double]
        L.orc_check_point.argtypes = [C.c_void_p, dp, C.c_int]
        L.orc_check_point_mindist.argtypes = [C.c_void_p, dp, C.c_double]
        L.orc_check_ray.argtypes = [C.c_void_p, dp, dp, C.c_int]
        L.orc_check_states.argtypes = [C.c_void_p, dp, C.c_size_t, C.c_int, up, C.c_int]
        L.orc_check_states_mindist.argtypes = [C.c_void_p, dp, C.c_size_t, C.c_double, up, C.c_int]
        L.orc_check_segments.argtypes = [C.c_void_p, dp, dp, C.c_size_t, C.c_int, up, C.c_int]
        L.orc_check_trajectory.argtypes = [C.c_void_p, dp, C.c_size_t, C.c_size_t, C.c_double, up,
                                           C.POINTER(C.c_long)]
        L.orc_poly_eval.restype = C.c_double
        L.orc_poly_eval.argtypes = [dp, C.c_int, C.c_double]
        L.orc_traj_evaluate_range.restype = C.c_size_t
        L.orc_traj_evaluate_range.argtypes = [C.c_int, C.c_int, dp, dp, C.c_double, C.c_double,
                                              C.c_double, dp, dp, C.c_size_t]
        L.orc_traj_max_time.restype = C.c_double
        L.orc_traj_max_time.argtypes = [C.c_int, dp]
        L.orc_check_poly_traj.argtypes = [C.c_void_p, C.c_int, C.c_int, dp, dp, C.c_double,
                                          C.c_double, up, C.c_size_t, C.POINTER(C.c_size_t),
                                          C.POINTER(C.c_long)]
        L.orc_mix64.restype = C.c_uint64
        L.orc_mix64.argtypes = [C.c_uint64, C.c_uint64]
        L.orc_gen_states.argtypes = [C.c_uint64, C.c_uint64, C.c_size_t, dp, dp, dp]
        L.orc_gen_segments.argtypes = [C.c_uint64, C.c_uint64, C.c_size_t, dp, dp, C.c_double, dp, dp]
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _up(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint32))


def _f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None:
        a = a.reshape(shape)
    return a


ERRORS = {1: "Rotation around x axis is not supported", 2: "Rotation around y axis is not supported",
          3: "Center z position must be zero", 4: "no geometry for component", 5: "no such obb"}


def descs_from_component(component):
    """component = dict name -> {position,size,type} in file order (ConfigParserYAML.cpp:54-73)."""
    arr = (ObbDesc * len(component))()
    for i, (_, o) in enumerate(component.items()):
        arr[i].position[:] = [float(v) for v in o["position"]]
        arr[i].size[:] = [float(v) for v in o["size"]]
        arr[i].filling = 1 if o["type"] == "filling" else 0
    return arr


class OracleWorld:
    """Mirror of the reference World (include/World.h) on the oracle."""

    def __init__(self, config):
        if isinstance(config, str):
            with open(config) as f:
                config = json.load(f)
        self.config = config
        L = lib()
        infl = config["world_properties"]["inflate_radius"]
        self.h = L.orc_world_create(float(infl["gate"]), float(infl["obstacle"]))
        for tid, name in config["gate_id_to_name_mapping"].items():
            d = descs_from_component(config["component_geometry"][name])
            L.orc_world_set_gate_geometry(self.h, int(tid), d, len(d))
        d = descs_from_component(config["component_geometry"]["obstacle"])
        L.orc_world_set_obstacle_geometry(self.h, d, len(d))

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_world_destroy(self.h)
            self.h = None

    @staticmethod
    def _chk(rc):
        if rc:
            raise RuntimeError(ERRORS.get(rc, "oracle error %d" % rc))

    def reset(self):
        lib().orc_world_reset(self.h)

    def add_gate(self, gate_id, pose7):
        self._chk(lib().orc_world_add_gate(self.h, gate_id, _dp(_f64(pose7, 7))))

    def update_gate(self, gate_id, pose7):
        self._chk(lib().orc_world_update_gate(self.h, gate_id, _dp(_f64(pose7, 7))))

    def add_obstacle(self, obstacle_id, pose6):
        self._chk(lib().orc_world_add_obstacle(self.h, obstacle_id, _dp(_f64(pose6, 6))))

    def parse_gates_and_obstacles(self, gates, obstacles):
        self.reset()
        for i, g in enumerate(gates):
            g = list(g)
            g[2] = 0.0
            self.add_gate(i, g)
        for i, o in enumerate(obstacles):
            self.add_obstacle(i, o)

    def size(self):
        return lib().orc_world_size(self.h)

    def obbs(self):
        out = []
        for i in range(self.size()):
            o = Obb()
            lib().orc_world_get(self.h, i, C.byref(o))
            out.append(dict(center=list(o.center), half=list(o.half), rot=list(o.rot),
                            aabb_min=list(o.aabb_min), aabb_max=list(o.aabb_max), inflate=o.inflate,
                            filling=o.filling, is_gate=o.is_gate, object_id=o.object_id,
                            obb_index=o.obb_index))
        return out

    def check_point(self, p, can_pass_gate=False):
        return bool(lib().orc_check_point(self.h, _dp(_f64(p, 3)), int(can_pass_gate)))

    def check_point_mindist(self, p, min_distance):
        return bool(lib().orc_check_point_mindist(self.h, _dp(_f64(p, 3)), float(min_distance)))

    def check_ray(self, a, b, can_pass_gate=False):
        return bool(lib().orc_check_ray(self.h, _dp(_f64(a, 3)), _dp(_f64(b, 3)), int(can_pass_gate)))

    def check_states(self, xyz, can_pass_gate=False, threads=1):
        xyz = _f64(xyz, (-1, 3))
        n = xyz.shape[0]
        mask = np.zeros((n + 31) // 32, dtype=np.uint32)
        lib().orc_check_states(self.h, _dp(xyz), n, int(can_pass_gate), _up(mask), threads)
        return mask

    def check_states_mindist(self, xyz, min_distance, threads=1):
        xyz = _f64(xyz, (-1, 3))
        n = xyz.shape[0]
        mask = np.zeros((n + 31) // 32, dtype=np.uint32)
        lib().orc_check_states_mindist(self.h, _dp(xyz), n, float(min_distance), _up(mask), threads)
        return mask

    def check_segments(self, a, b, can_pass_gate=False, threads=1):
        a = _f64(a, (-1, 3))
        b = _f64(b, (-1, 3))
        n = a.shape[0]
        mask = np.zeros((n + 31) // 32, dtype=np.uint32)
        lib().orc_check_segments(self.h, _dp(a), _dp(b), n, int(can_pass_gate), _up(mask), threads)
        return mask

    def check_trajectory(self, traj, min_distance):
        """traj: (rows, cols) array; evaluated as the column-major Eigen matrix."""
        t = np.asfortranarray(np.asarray(traj, dtype=np.float64))
        rows = t.shape[0]
        mask = np.zeros((rows + 31) // 32, dtype=np.uint32)
        first = C.c_long(-1)
        ok = lib().orc_check_trajectory(self.h, _dp(t), rows, rows, float(min_distance), _up(mask),
                                        C.byref(first))
        return bool(ok), mask, first.value

    def check_poly_traj(self, coeffs, durations, dt, min_distance, max_samples=4096):
        coeffs = _f64(coeffs)
        durations = _f64(durations)
        k = durations.shape[0]
        ncoef = coeffs.size // (3 * k)
        words = (max_samples + 31) // 32
        mask = np.zeros(words, dtype=np.uint32)
        n = C.c_size_t(0)
        first = C.c_long(-1)
        ok = lib().orc_check_poly_traj(self.h, k, ncoef, _dp(coeffs), _dp(durations), float(dt),
                                       float(min_distance), _up(mask), words, C.byref(n),
                                       C.byref(first))
        return bool(ok), mask, n.value, first.value


def poly_eval(coeffs, t):
    c = _f64(coeffs)
    return lib().orc_poly_eval(_dp(c), c.size, float(t))


def traj_evaluate_range(coeffs, durations, t_start, t_end, dt, max_out=1 << 16):
    coeffs = _f64(coeffs)
    durations = _f64(durations)
    k = durations.shape[0]
    ncoef = coeffs.size // (3 * k)
    xyz = np.zeros((max_out, 3))
    times = np.zeros(max_out)
    n = lib().orc_traj_evaluate_range(k, ncoef, _dp(coeffs), _dp(durations), t_start, t_end, dt,
                                      _dp(xyz), _dp(times), max_out)
    m = min(n, max_out)
    return xyz[:m].copy(), times[:m].copy(), n


def traj_max_time(durations):
    d = _f64(durations)
    return lib().orc_traj_max_time(d.shape[0], _dp(d))


def gen_states(seed, first, n, lo, hi):
    xyz = np.empty((n, 3))
    lib().orc_gen_states(seed, first, n, _dp(_f64(lo, 3)), _dp(_f64(hi, 3)), _dp(xyz))
    return xyz


def gen_segments(seed, first, n, lo, hi, max_len):
    a = np.empty((n, 3))
    b = np.empty((n, 3))
    lib().orc_gen_segments(seed, first, n, _dp(_f64(lo, 3)), _dp(_f64(hi, 3)), float(max_len), _dp(a),
                           _dp(b))
    return a, b
