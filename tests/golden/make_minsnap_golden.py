"""Generates tests/golden/minsnap_golden.npz by RUNNING THE REFERENCE'S OWN MIN-SNAP CODE in this container:
oracle/_ref/libepp_refpoly.so now also holds external/poly_traj/src/trajectory_generator.cpp and, through it,
PolynomialOptimization<10>::setupFromVertices / solveLinear (impl/polynomial_optimization_linear_impl.h), compiled
unmodified.  Only the linear algebra underneath (Eigen fixed-size / sparse matrices, SparseQR) is a stand-in
(oracle/shim_poly/Eigen/{Core,Sparse}), so these vectors pin the FORMULATION to rounding, not Eigen's bits.

    make -C oracle ref && python tests/golden/make_minsnap_golden.py

Contents: per problem p (K_p segments): waypoints, segment times, start velocity / acceleration, the reference's
coefficients [K][3][10]; and for a few waypoint lists the complete N x 10 matrix of poly_traj::generateTrajectory
(segment-time heuristic + solve + the three sampling passes), as OnlineTrajGenerator receives it."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]

import refpoly_py  # noqa: E402

KS = [1, 1, 2, 2, 3, 4, 5, 6, 8, 8, 12, 16, 24, 40, 64]
MAXK = max(KS)


def main():
    assert refpoly_py.has_minsnap(), "build oracle/_ref first: make -C oracle ref"
    rng = np.random.default_rng(0x5EED0006)
    P = len(KS)
    wp = np.zeros((P, MAXK + 1, 3))
    T = np.zeros((P, MAXK))
    va = np.zeros((P, 6))
    coeffs = np.zeros((P, MAXK, 3, 10))
    for p, K in enumerate(KS):
        wp[p, : K + 1] = np.cumsum(rng.normal(size=(K + 1, 3)) * [0.8, 0.8, 0.3], axis=0) + [0.0, 0.0, 1.0]
        T[p, :K] = rng.uniform(0.4, 3.0, size=K)
        if p % 3:
            va[p] = rng.normal(size=6) * 0.5
        coeffs[p, :K] = refpoly_py.minsnap(wp[p, : K + 1], T[p, :K], va[p, :3], va[p, 3:])
    out = dict(K=np.array(KS), waypoints=wp, times=T, start_vel_acc=va, coeffs=coeffs)

    # complete generateTrajectory outputs (v_max, a_max, dt and the takeoff offset of config.json's neighbourhood)
    gen_wp, gen_rows, gen_par, gen_va = [], [], [], []
    for g, (n_wp, v_max, a_max, dt, off) in enumerate([(2, 1.0, 2.0, 0.05, 0.0), (5, 2.0, 2.0, 0.05, 0.5),
                                                       (9, 3.0, 4.0, 0.02, 1.25), (14, 1.5, 1.0, 0.1, 0.0)]):
        w = np.cumsum(rng.normal(size=(n_wp, 3)) * [0.9, 0.9, 0.25], axis=0) + [0.0, 0.0, 1.0]
        v0a0 = rng.normal(size=6) * 0.4 if g % 2 else np.zeros(6)
        rows = refpoly_py.generate_trajectory(w, v_max, a_max, dt, off, v0a0[:3], v0a0[3:])
        gen_wp.append(w)
        gen_rows.append(rows)
        gen_par.append([v_max, a_max, dt, off])
        gen_va.append(v0a0)
    out["gen_count"] = np.array(len(gen_wp))
    for g in range(len(gen_wp)):
        out["gen%d_waypoints" % g], out["gen%d_rows" % g] = gen_wp[g], gen_rows[g]
    out["gen_params"], out["gen_start_vel_acc"] = np.array(gen_par), np.array(gen_va)
    path = os.path.join(HERE, "minsnap_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", [r.shape for r in gen_rows])


if __name__ == "__main__":
    main()
