"""refpoly_py — ctypes binding of oracle/_ref/libepp_refpoly.so: the reference's OWN
external/poly_traj/src/{polynomial,segment,trajectory,vertex,motion_defines}.cpp compiled unmodified against
the stand-in headers in oracle/shim_poly/ (TEST INFRASTRUCTURE; built by `make -C oracle ref` where
/root/reference exists, the prebuilt .so travels to the GPU box)."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_ref", "libepp_refpoly.so")
_lib = None
_dp = C.POINTER(C.c_double)


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH)
        L.refpoly_horner.restype = C.c_double
        L.refpoly_horner.argtypes = [_dp, C.c_int, C.c_double, C.c_int]
        L.refpoly_base_coefficient.restype = C.c_double
        L.refpoly_base_coefficient.argtypes = [C.c_int, C.c_int]
        L.refpoly_max_time.restype = C.c_double
        L.refpoly_max_time.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_int]
        L.refpoly_evaluate_range.restype = C.c_long
        L.refpoly_evaluate_range.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, _dp, _dp,
                                             C.c_long]
        L.refpoly_evaluate.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, _dp]
        L.refpoly_segment_times.argtypes = [_dp, C.c_int, C.c_int, C.c_double, C.c_double, _dp]
        _lib = L
    return _lib


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _p(a):
    return a.ctypes.data_as(_dp)


def horner(coeffs, t, derivative=0):
    c = _f64(coeffs)
    return lib().refpoly_horner(_p(c), c.size, float(t), int(derivative))


def max_time(coeffs, durations):
    c, d = _f64(coeffs), _f64(durations)
    k = d.shape[0]
    return lib().refpoly_max_time(_p(c), _p(d), k, 3, c.size // (3 * k))


def evaluate_range(coeffs, durations, dt, derivative=0, max_out=1 << 16):
    """coeffs (K, 3, N), durations (K,) -> (samples (n, 3), times (n,), n) exactly as
    poly_traj::generateTrajectory samples its trajectory (trajectory_generator.cpp:60-69)."""
    c, d = _f64(coeffs), _f64(durations)
    k = d.shape[0]
    out = np.zeros((max_out, 3))
    times = np.zeros(max_out)
    n = lib().refpoly_evaluate_range(_p(c), _p(d), k, 3, c.size // (3 * k), float(dt), int(derivative), _p(out),
                                     _p(times), max_out)
    m = min(n, max_out)
    return out[:m].copy(), times[:m].copy(), n


def segment_times(waypoints, v_max, a_max):
    w = _f64(waypoints).reshape(-1, 3)
    out = np.zeros(w.shape[0] - 1)
    lib().refpoly_segment_times(_p(w), w.shape[0], 3, float(v_max), float(a_max), _p(out))
    return out


def has_minsnap():
    """False for a prebuilt oracle/_ref that predates the min-snap entry points."""
    return available() and hasattr(lib(), "refpoly_minsnap")


def minsnap(waypoints, times, v0=None, a0=None):
    """The reference's PolynomialOptimization<10> (setupFromVertices + solveLinear, unmodified) on the vertex pattern
    of poly_traj::generateTrajectory, for GIVEN segment times.  Returns coefficients [K][3][10]."""
    L = lib()
    L.refpoly_minsnap.argtypes = [_dp, C.c_int, _dp, _dp, _dp, _dp]
    wp = _f64(np.asarray(waypoints, float).reshape(-1, 3))
    t = _f64(times)
    v = _f64(np.zeros(3) if v0 is None else v0)
    a = _f64(np.zeros(3) if a0 is None else a0)
    out = np.zeros((wp.shape[0] - 1, 3, 10))
    rc = L.refpoly_minsnap(_p(wp), wp.shape[0], _p(t), _p(v), _p(a), _p(out))
    if rc:
        raise RuntimeError("refpoly_minsnap failed")
    return out


def generate_trajectory(waypoints, v_max, a_max, dt, offset=0.0, v0=None, a0=None, max_rows=1 << 16):
    """poly_traj::generateTrajectory (trajectory_generator.cpp:12-100), unmodified: N x 10 rows."""
    L = lib()
    L.refpoly_generate_trajectory.restype = C.c_long
    L.refpoly_generate_trajectory.argtypes = [_dp, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, _dp, _dp,
                                              _dp, C.c_long]
    wp = _f64(np.asarray(waypoints, float).reshape(-1, 3))
    v = _f64(np.zeros(3) if v0 is None else v0)
    a = _f64(np.zeros(3) if a0 is None else a0)
    rows = np.zeros((max_rows, 10))
    n = L.refpoly_generate_trajectory(_p(wp), wp.shape[0], v_max, a_max, dt, offset, _p(v), _p(a), _p(rows), max_rows)
    if n < 0:
        raise RuntimeError("refpoly_generate_trajectory failed")
    return rows[:n].copy()
