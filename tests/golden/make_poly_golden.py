"""Generates tests/golden/poly_golden.npz by RUNNING THE REFERENCE'S OWN CODE in this container:
oracle/_ref/libepp_refpoly.so = /root/reference/external/poly_traj/src/{polynomial,segment,trajectory,vertex,
motion_defines}.cpp compiled unmodified against oracle/shim_poly, and oracle/_ref/libepp_ref.so (the reference's
World) for the validity of the sampled rows (make -C oracle ref).

    python tests/golden/make_poly_golden.py

Contents: Horner known answers (Polynomial::evaluate, derivatives 0..2), evaluateRange samples / sampling times /
counts of seeded trajectories sampled the way poly_traj::generateTrajectory samples them, the Nfabian segment
times, and — end to end — PathPlanner::checkTrajectoryValidity's per-row verdicts on those samples against the
track world (reference sampler -> reference World::checkPointValidity(point, minDistance))."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]

import epp_b200  # noqa: E402
import ref_py  # noqa: E402
import refpoly_py  # noqa: E402
from epp_b200 import worlds  # noqa: E402

MIN_DISTANCE = 0.1
MAX_SAMPLES = 512


def main():
    assert ref_py.available() and refpoly_py.available(), "build oracle/_ref first: make -C oracle ref"
    rng = np.random.default_rng(0x5EED0004)
    out = {}

    # Horner known answers
    hc = rng.normal(size=(256, 10)) * 10.0 ** rng.integers(-3, 3, size=(256, 1))
    ht = rng.uniform(0.0, 6.0, size=256)
    out["horner_coeffs"], out["horner_t"] = hc, ht
    for der in (0, 1, 2):
        out["horner_d%d" % der] = np.array([refpoly_py.horner(hc[i], ht[i], der) for i in range(256)])

    # trajectories: rest-to-rest closed forms (config 4's generator) + arbitrary coefficients; dt values that
    # do and do not divide the segment times
    config = epp_b200.load_config()
    g, o, lo, hi = worlds.track_world(config)
    c1, T1, dt1 = worlds.rest_to_rest_trajectories(24, 4, lo, hi, n_samples=200)
    c2 = rng.normal(size=(8, 3, 3, 10)) * 0.05
    c2[..., 0] = rng.uniform(lo, hi, size=(8, 3, 3))
    T2 = np.array([[0.5, 0.25, 1.0]] * 4 + [[0.3, 0.7, 0.1]] * 4)
    dt2 = np.array([0.125, 0.25, 0.01, 0.05, 0.1, 0.0999, 0.3, 0.011])
    K = 4
    coeffs = np.zeros((32, K, 3, 10))
    T = np.full((32, K), 1e-3)  # a 1 ms tail segment pads the 3-segment trajectories to K = 4
    coeffs[:24], T[:24] = c1, T1
    coeffs[24:, :3], T[24:, :3] = c2, T2
    coeffs[24:, 3, :, 0] = c2[:, 2, :, 0]
    dt = np.concatenate([dt1, dt2])
    out["traj_coeffs"], out["traj_T"], out["traj_dt"] = coeffs, T, dt

    rw = ref_py.RefWorld(config)
    rw.parse_gates_and_obstacles(g, o)
    n_samples = np.zeros(32, dtype=np.int64)
    samples = np.zeros((32, MAX_SAMPLES, 3))
    vel = np.zeros((32, MAX_SAMPLES, 3))   # the velocity / acceleration passes of generateTrajectory
    acc = np.zeros((32, MAX_SAMPLES, 3))   # (trajectory_generator.cpp:71-79)
    times = np.zeros((32, MAX_SAMPLES))
    masks = np.zeros((32, MAX_SAMPLES // 32), dtype=np.uint32)
    first_invalid = np.full(32, -1, dtype=np.int64)
    max_time = np.zeros(32)
    for k in range(32):
        x, t, n = refpoly_py.evaluate_range(coeffs[k], T[k], dt[k])
        assert n <= MAX_SAMPLES, n
        n_samples[k] = n
        samples[k, :n], times[k, :n] = x, t
        v, tv, nv = refpoly_py.evaluate_range(coeffs[k], T[k], dt[k], derivative=1)
        a, ta, na = refpoly_py.evaluate_range(coeffs[k], T[k], dt[k], derivative=2)
        assert nv == n and na == n and np.array_equal(tv, t) and np.array_equal(ta, t)
        vel[k, :n], acc[k, :n] = v, a
        max_time[k] = refpoly_py.max_time(coeffs[k], T[k])
        m = rw.check_states_mindist(x, MIN_DISTANCE)
        masks[k, : m.size] = m
        bits = np.unpackbits(m.view(np.uint8), bitorder="little")[:n]
        bad = np.nonzero(bits == 0)[0]
        first_invalid[k] = bad[0] if bad.size else -1
    out.update(traj_vel=vel, traj_acc=acc, traj_n=n_samples, traj_samples=samples, traj_times=times, traj_masks=masks,
               traj_first_invalid=first_invalid, traj_max_time=max_time,
               min_distance=np.array(MIN_DISTANCE))

    wp = rng.uniform(lo, hi, size=(33, 3))
    out["nfabian_waypoints"] = wp
    out["nfabian_times"] = refpoly_py.segment_times(wp, 1.0, 2.0)

    path = os.path.join(HERE, "poly_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; samples per trajectory:", n_samples.tolist())


if __name__ == "__main__":
    main()
