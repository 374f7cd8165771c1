"""ref_py — ctypes binding of oracle/_ref/libepp_ref.so: the reference's OWN src/OBB.cpp, src/Object.cpp,
src/World.cpp, src/StateValidator.cpp, src/MotionValidator.cpp compiled unmodified against the stand-in
headers in oracle/shim/ (TEST INFRASTRUCTURE; built by `make -C oracle ref` where /root/reference exists,
the prebuilt .so travels to the GPU box).  Same surface as oracle_py.OracleWorld for the calls the tests
and bench.py use."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_ref", "libepp_ref.so")


class RefDesc(C.Structure):
    _fields_ = [("position", C.c_double * 3), ("size", C.c_double * 3), ("filling", C.c_int)]


_lib = None
_vp = C.c_void_p


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH)
        L.ref_world_create.restype = _vp
        L.ref_world_create.argtypes = [C.c_double, C.c_double]
        L.ref_world_destroy.argtypes = [_vp]
        L.ref_last_error.restype = C.c_char_p
        L.ref_last_error.argtypes = [_vp]
        L.ref_world_set_gate_geometry.argtypes = [_vp, C.c_int, C.POINTER(RefDesc), C.c_int]
        L.ref_world_set_obstacle_geometry.argtypes = [_vp, C.POINTER(RefDesc), C.c_int]
        L.ref_world_reset.argtypes = [_vp]
        for f in (L.ref_world_add_gate, L.ref_world_update_gate, L.ref_world_add_obstacle):
            f.argtypes = [_vp, C.c_int, _vp]
        L.ref_object_records.argtypes = [_vp, C.c_int, C.c_int, _vp, _vp, C.c_double, _vp, C.c_int]
        L.ref_check_states.argtypes = [_vp, _vp, C.c_size_t, C.c_int, _vp, C.c_int, C.c_int]
        L.ref_check_states_mindist.argtypes = [_vp, _vp, C.c_size_t, C.c_double, _vp, C.c_int]
        L.ref_check_segments.argtypes = [_vp, _vp, _vp, C.c_size_t, C.c_int, _vp, C.c_int, C.c_int]
        _lib = L
    return _lib


def _descs(component):
    arr = (RefDesc * len(component))()
    for i, (_, o) in enumerate(component.items()):
        arr[i].position[:] = [float(v) for v in o["position"]]
        arr[i].size[:] = [float(v) for v in o["size"]]
        arr[i].filling = 1 if o["type"] == "filling" else 0
    return arr


def _f64(a, shape):
    return np.ascontiguousarray(a, dtype=np.float64).reshape(shape)


class RefWorld:
    """The reference World (include/World.h) behind a C shim.  via_validators=True routes the batched
    loops through StateValidator::isValid / MotionValidator::checkMotion on OMPL-shaped states."""

    def __init__(self, config, via_validators=False):
        L = lib()
        infl = config["world_properties"]["inflate_radius"]
        self._h = L.ref_world_create(float(infl["gate"]), float(infl["obstacle"]))
        self.via = 1 if via_validators else 0
        for tid, name in config["gate_id_to_name_mapping"].items():
            d = _descs(config["component_geometry"][name])
            L.ref_world_set_gate_geometry(self._h, int(tid), d, len(d))
        d = _descs(config["component_geometry"]["obstacle"])
        L.ref_world_set_obstacle_geometry(self._h, d, len(d))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().ref_world_destroy(self._h)
            self._h = None

    def _chk(self, rc):
        if rc:
            raise RuntimeError(lib().ref_last_error(self._h).decode())

    def reset(self):
        lib().ref_world_reset(self._h)

    def add_gate(self, gate_id, pose7):
        p = _f64(pose7, 7)
        self._chk(lib().ref_world_add_gate(self._h, int(gate_id), p.ctypes.data))

    def update_gate(self, gate_id, pose7):
        p = _f64(pose7, 7)
        self._chk(lib().ref_world_update_gate(self._h, int(gate_id), p.ctypes.data))

    def add_obstacle(self, obstacle_id, pose6):
        p = _f64(pose6, 6)
        self._chk(lib().ref_world_add_obstacle(self._h, int(obstacle_id), p.ctypes.data))

    def parse_gates_and_obstacles(self, gates, obstacles):
        """PathPlanner::parseGatesAndObstacles (src/PathPlanner.cpp:60-78)."""
        self.reset()
        for i, g in enumerate(gates):
            g = np.array(g, dtype=np.float64)
            g[2] = 0.0
            self.add_gate(i, g)
        for i, o in enumerate(obstacles):
            self.add_obstacle(i, o)

    def object_records(self, is_gate, type_id, pos, rot, inflate, max_obbs=16):
        """Object::createFromDescription + getAABBs -> list of dicts (center, half, rot, aabb_min, aabb_max)."""
        out = np.zeros((max_obbs, 24))
        pos, rot = _f64(pos, 3), _f64(rot, 3)
        n = lib().ref_object_records(self._h, int(is_gate), int(type_id), pos.ctypes.data, rot.ctypes.data,
                                     float(inflate), out.ctypes.data, max_obbs)
        if n < 0:
            raise RuntimeError(lib().ref_last_error(self._h).decode())
        return [dict(center=r[0:3].tolist(), half=r[3:6].tolist(), rot=r[6:15].tolist(),
                     aabb_min=r[15:18].tolist(), aabb_max=r[18:21].tolist()) for r in out[:n]]

    def check_states(self, xyz, can_pass_gate=False, threads=1):
        xyz = _f64(xyz, (-1, 3))
        n = xyz.shape[0]
        mask = np.zeros((n + 31) // 32, dtype=np.uint32)
        lib().ref_check_states(self._h, xyz.ctypes.data, n, int(can_pass_gate), mask.ctypes.data, threads, self.via)
        return mask

    def check_states_mindist(self, xyz, min_distance, threads=1):
        xyz = _f64(xyz, (-1, 3))
        n = xyz.shape[0]
        mask = np.zeros((n + 31) // 32, dtype=np.uint32)
        lib().ref_check_states_mindist(self._h, xyz.ctypes.data, n, float(min_distance), mask.ctypes.data, threads)
        return mask

    def check_segments(self, a, b, can_pass_gate=False, threads=1):
        a, b = _f64(a, (-1, 3)), _f64(b, (-1, 3))
        n = a.shape[0]
        mask = np.zeros((n + 31) // 32, dtype=np.uint32)
        lib().ref_check_segments(self._h, a.ctypes.data, b.ctypes.data, n, int(can_pass_gate), mask.ctypes.data,
                                 threads, self.via)
        return mask


class RefPathWriter:
    """The reference's own PathWriter (src/PathWriter.cpp, unmodified; Eigen's stream format restated in the shim)."""

    def __init__(self, folder):
        L = lib()
        L.ref_writer_create.restype = _vp
        L.ref_writer_create.argtypes = [C.c_char_p]
        L.ref_writer_destroy.argtypes = [_vp]
        L.ref_writer_write_path.argtypes = [_vp, _vp, C.c_size_t]
        L.ref_writer_write_checkpoints.argtypes = [_vp, _vp, C.c_size_t]
        L.ref_writer_update_gate.argtypes = [_vp, C.c_int, _vp, C.c_int]
        L.ref_writer_update_obstacle.argtypes = [_vp, C.c_int, _vp, C.c_int]
        self._h = L.ref_writer_create(os.fsencode(folder))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().ref_writer_destroy(self._h)
            self._h = None

    def writePath(self, path):
        p = _f64(path, (-1, 3))
        lib().ref_writer_write_path(self._h, p.ctypes.data, p.shape[0])

    def writeCheckpoints(self, checkpoints):
        p = _f64(checkpoints, (-1, 3))
        lib().ref_writer_write_checkpoints(self._h, p.ctypes.data, p.shape[0])

    def updateGatePos(self, gateId, row):
        r = _f64(row, (-1,))
        lib().ref_writer_update_gate(self._h, int(gateId), r.ctypes.data, r.shape[0])

    def updateObstaclePos(self, obstacleId, row):
        r = _f64(row, (-1,))
        lib().ref_writer_update_obstacle(self._h, int(obstacleId), r.ctypes.data, r.shape[0])
