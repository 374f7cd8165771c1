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
