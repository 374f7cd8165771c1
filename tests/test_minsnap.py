"""Min-snap trajectory generation: the batched CUDA solve against
  * tests/golden/minsnap_golden.npz — outputs of THE REFERENCE'S OWN PolynomialOptimization<10>::setupFromVertices /
    solveLinear and poly_traj::generateTrajectory, compiled unmodified into oracle/_ref/libepp_refpoly.so
    (tests/golden/make_minsnap_golden.py); only the linear algebra underneath that code (Eigen's fixed-size and sparse
    matrices, SparseQR) is a stand-in, so these pin the formulation to rounding, not Eigen's bits;
  * oracle/minsnap_np.py — a numpy restatement of the same formulation in its unscaled coordinates, itself checked
    against the golden vectors and the live reference here.

FLOATING-POINT PATH — tolerance, stated here: positions along the trajectory agree to 1e-9 m (the unscaled formulation
inverts matrices of powers of T in double and is itself only ~1e-10 accurate), velocities to 1e-8 m/s, accelerations to
1e-7 m/s^2; boundary values and continuity of the derivatives hold to 1e-9 in units scaled by the segment time.
What IS bit-exact: the segment-time heuristic and the sampling times against the reference's own vertex.cpp /
trajectory.cpp (golden), the run-to-run determinism of the kernel, and everything downstream of the coefficients
(sampler + World check)."""
import os

import numpy as np
import pytest

import epp_b200
from epp_b200 import worlds

HERE = os.path.dirname(os.path.abspath(__file__))
TOL = 1e-9


@pytest.fixture(scope="module")
def ms():
    import minsnap_np
    return minsnap_np


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "minsnap_golden.npz"))


def worst_position_gap(ms, a, b, T, taus=np.linspace(0.0, 1.0, 17)):
    worst = 0.0
    for k in range(len(T)):
        for d in range(3):
            for tau in taus:
                worst = max(worst, abs(ms.evaluate(a[k, d], tau * T[k]) - ms.evaluate(b[k, d], tau * T[k])))
    return worst


def random_problem(rng, n, K, lo, hi):
    wp = rng.uniform(lo, hi, size=(n, K + 1, 3))
    T = epp_b200.segment_times(wp, 1.0, 2.0)
    va = rng.normal(size=(n, 6)) * 0.3
    return wp, T, va


def test_segment_times_equal_reference_golden():
    """epp_segment_times (host side of the C ABI, no GPU needed) vs the reference's estimateSegmentTimesNfabian."""
    z = np.load(os.path.join(HERE, "golden", "poly_golden.npz"))
    got = epp_b200.segment_times(z["nfabian_waypoints"], 1.0, 2.0)
    assert np.array_equal(got, z["nfabian_times"])


def test_oracle_equals_reference_golden(ms, golden):
    """numpy restatement vs the reference's own solver, 15 problems of 1..64 segments."""
    for p, K in enumerate(golden["K"]):
        wp, T, va = golden["waypoints"][p, : K + 1], golden["times"][p, :K], golden["start_vel_acc"][p]
        mine = ms.solve(wp, T, va[:3], va[3:])
        assert worst_position_gap(ms, mine, golden["coeffs"][p, :K], T) < TOL, (p, K)


def test_live_reference_solver_equals_oracle(ms):
    refpoly = pytest.importorskip("refpoly_py")
    if not refpoly.has_minsnap():
        pytest.skip("oracle/_ref/libepp_refpoly.so without the min-snap entry points (needs /root/reference)")
    rng = np.random.default_rng(31)
    for K in (1, 2, 4, 7, 11, 20):
        wp = rng.uniform(-2, 2, size=(K + 1, 3))
        T = rng.uniform(0.5, 3.0, size=K)
        v0, a0 = rng.normal(size=3) * 0.4, rng.normal(size=3) * 0.4
        ref = refpoly.minsnap(wp, T, v0, a0)
        # violent random problems (3 m in 0.7 s: coefficients of 1e2) lose digits in BOTH unscaled solves: the gap is
        # bounded relative to the largest coefficient
        assert worst_position_gap(ms, ms.solve(wp, T, v0, a0), ref, T) < TOL * max(1.0, np.abs(ref).max()), K


def test_oracle_properties(ms):
    rng = np.random.default_rng(3)
    for K in (1, 2, 5, 8):
        wp = rng.uniform(-2, 2, size=(K + 1, 3))
        T = rng.uniform(0.7, 3.0, size=K)
        v0, a0 = rng.normal(size=3) * 0.3, rng.normal(size=3) * 0.3
        c = ms.solve(wp, T, v0, a0)
        for k in range(K):
            for d in range(3):
                assert abs(ms.evaluate(c[k, d], 0.0) - wp[k, d]) < 1e-8
                assert abs(ms.evaluate(c[k, d], T[k]) - wp[k + 1, d]) < 1e-8
                if k + 1 < K:
                    for r in range(1, 5):
                        s = T[k] ** r
                        assert abs(ms.evaluate(c[k, d], T[k], r) - ms.evaluate(c[k + 1, d], 0.0, r)) * s < 1e-7
        for d in range(3):
            assert abs(ms.evaluate(c[0, d], 0.0, 1) - v0[d]) < 1e-9 and abs(ms.evaluate(c[0, d], 0.0, 2) - a0[d]) < 1e-9
            for r in range(1, 5):
                assert abs(ms.evaluate(c[K - 1, d], T[K - 1], r)) * T[K - 1] ** r < 1e-7
    # one segment at rest at both ends = the closed form config 4 uses (vertex.cpp:146-170)
    c = ms.solve(np.array([[0.0, 0, 0], [1, 2, 3]]), [2.0])
    s = np.array([0, 0, 0, 0, 0, 126, -420, 540, -315, 70.0]) / 2.0 ** np.arange(10)
    assert np.abs(c[0, 0] - s).max() < 1e-10 and np.abs(c[0, 2] - 3 * s).max() < 1e-9


@pytest.mark.gpu
@pytest.mark.parametrize("K", [1, 2, 3, 8, 16, 40])
def test_gpu_minsnap_equals_oracle(ms, config, K):
    g, o, lo, hi = worlds.track_world(config)
    rng = np.random.default_rng(100 + K)
    n = 40
    wp, T, va = random_problem(rng, n, K, lo, hi)
    got = epp_b200.minsnap(wp, T, va)
    assert got.shape == (n, K, 3, 10)
    taus = np.linspace(0.0, 1.0, 17)
    worst = 0.0
    for t in range(n):
        want = ms.solve(wp[t], T[t], va[t, :3], va[t, 3:])
        for k in range(K):
            for d in range(3):
                for tau in taus:
                    x = tau * T[t, k]
                    worst = max(worst, abs(ms.evaluate(got[t, k, d], x) - ms.evaluate(want[k, d], x)))
    assert worst < TOL, worst
    # exact interpolation conditions hold for the kernel's own output (tighter than the restatement's)
    for t in range(n):
        for k in range(K):
            for d in range(3):
                assert abs(got[t, k, d, 0] - wp[t, k, d]) < 1e-12
                assert abs(ms.evaluate(got[t, k, d], T[t, k]) - wp[t, k + 1, d]) < 1e-10
    at_rest = epp_b200.minsnap(wp, T)  # start at rest
    for d in range(3):
        assert at_rest[0, 0, d, 1] == 0.0 and at_rest[0, 0, d, 2] == 0.0


@pytest.mark.gpu
def test_gpu_minsnap_equals_reference_golden(ms, golden):
    """The CUDA solve vs the reference's own PolynomialOptimization<10> on the committed problems (1..64 segments)."""
    for p, K in enumerate(golden["K"]):
        wp, T, va = golden["waypoints"][p, : K + 1], golden["times"][p, :K], golden["start_vel_acc"][p]
        got = epp_b200.minsnap(wp[None], T[None], va[None])[0]
        assert worst_position_gap(ms, got, golden["coeffs"][p, :K], T) < TOL, (p, K)


@pytest.mark.gpu
def test_gpu_generate_trajectory_equals_reference_golden(golden):
    """generate_trajectory (segment times -> k_minsnap -> k_sample_poly) vs poly_traj::generateTrajectory's N x 10
    matrix: same row count, time column bit-equal, positions / velocities / accelerations within the stated tolerances."""
    import online_traj_planner as otp
    for g in range(int(golden["gen_count"])):
        wp, want = golden["gen%d_waypoints" % g], golden["gen%d_rows" % g]
        v_max, a_max, dt, off = golden["gen_params"][g]
        va = golden["gen_start_vel_acc"][g]
        got = otp.generate_trajectory(wp, v_max, a_max, dt, off, va[:3], va[3:])
        assert got.shape == want.shape, g
        assert np.array_equal(got[:, 9], want[:, 9]), g
        assert np.abs(got[:, [0, 3, 6]] - want[:, [0, 3, 6]]).max() < TOL, g
        assert np.abs(got[:, [1, 4, 7]] - want[:, [1, 4, 7]]).max() < 1e-8, g
        assert np.abs(got[:, [2, 5, 8]] - want[:, [2, 5, 8]]).max() < 1e-7, g


@pytest.mark.gpu
def test_gpu_minsnap_is_deterministic_and_feeds_the_sweep(oracle, config):
    """Same bits every run; generate -> sample -> check entirely on the device equals the CPU oracle's sampler +
    World run on the SAME coefficients (bit-exact: the coefficients are the sweep's input)."""
    import torch
    g, o, lo, hi = worlds.track_world(config)
    rng = np.random.default_rng(77)
    n, K = 96, 6
    wp, T, va = random_problem(rng, n, K, lo, hi)
    d_wp, d_T, d_va = (torch.from_numpy(x).cuda() for x in (wp, T, va))
    c1 = epp_b200.minsnap_dev(d_wp, d_T, d_va)
    c2 = epp_b200.minsnap_dev(d_wp, d_T, d_va)
    torch.cuda.synchronize()
    assert torch.equal(c1, c2)
    assert np.array_equal(c1.cpu().numpy(), epp_b200.minsnap(wp, T, va))
    gw = epp_b200.World(config)
    gw.parse_gates_and_obstacles(g, o)
    ow = oracle.OracleWorld(config)
    ow.parse_gates_and_obstacles(g, o)
    dt = T.sum(axis=1) / 199.0
    valid, smask, ns, first = gw.check_poly_trajectories_dev(c1, d_T, torch.from_numpy(dt).cuda(), 0.1, max_samples=256)
    torch.cuda.synchronize()
    coeffs = c1.cpu().numpy()
    ns, first, smask = ns.cpu().numpy(), first.cpu().numpy(), smask.cpu().numpy().view(np.uint32)
    for t in range(n):
        ok, mk, cnt, f = ow.check_poly_traj(coeffs[t], T[t], dt[t], 0.1, max_samples=256)
        assert ns[t] == cnt and first[t] == f and np.array_equal(smask[t], mk[: smask.shape[1]])
    assert (first >= 0).any()
