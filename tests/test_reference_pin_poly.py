"""Pins the polynomial-sampler rows against THE REFERENCE'S OWN CODE.

Sources, both produced from /root/reference/external/poly_traj/src/{polynomial,segment,trajectory,vertex,
motion_defines}.cpp compiled unmodified against oracle/shim_poly (Eigen / glog stand-ins only):
  * tests/golden/poly_golden.npz        — committed outputs (tests/golden/make_poly_golden.py), incl. the
                                           reference World's verdicts on the reference's samples;
  * oracle/_ref/libepp_refpoly.so       — the same code, live, when it has been built.
CPU tests check the oracle's restatement; the GPU tests check the fused sweep (K3b) and the pre-sampled sweep
(K3a) through the C ABI against the reference outputs.  Bit-exact everywhere except the Nfabian time heuristic
in numpy (exp of numpy vs libm: 1 ulp allowed, said in the test)."""
import os

import numpy as np
import pytest

import epp_b200
from epp_b200 import worlds

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "poly_golden.npz"))


def test_oracle_horner_equals_reference_golden(oracle, golden):
    c, t = golden["horner_coeffs"], golden["horner_t"]
    got = np.array([oracle.poly_eval(c[i], t[i]) for i in range(c.shape[0])])
    assert np.array_equal(got, golden["horner_d0"])


def test_oracle_evaluate_range_equals_reference_golden(oracle, golden):
    coeffs, T, dt = golden["traj_coeffs"], golden["traj_T"], golden["traj_dt"]
    for k in range(coeffs.shape[0]):
        t_end = oracle.traj_max_time(T[k])
        assert t_end == golden["traj_max_time"][k]
        xyz, times, n = oracle.traj_evaluate_range(coeffs[k], T[k], 0.0, t_end, dt[k])
        assert n == golden["traj_n"][k], k
        assert np.array_equal(xyz, golden["traj_samples"][k, :n]), k
        assert np.array_equal(times, golden["traj_times"][k, :n]), k


def test_oracle_trajectory_verdicts_equal_reference_golden(oracle, config, golden):
    """oracle sampler + oracle World == reference sampler + reference World, row for row."""
    g, o, lo, hi = worlds.track_world(config)
    ow = oracle.OracleWorld(config)
    ow.parse_gates_and_obstacles(g, o)
    coeffs, T, dt = golden["traj_coeffs"], golden["traj_T"], golden["traj_dt"]
    md = float(golden["min_distance"])
    words = golden["traj_masks"].shape[1]
    for k in range(coeffs.shape[0]):
        ok, mask, n, first = ow.check_poly_traj(coeffs[k], T[k], dt[k], md, max_samples=words * 32)
        assert n == golden["traj_n"][k]
        assert np.array_equal(mask[:words], golden["traj_masks"][k]), k
        assert first == golden["traj_first_invalid"][k]
        assert ok == (golden["traj_first_invalid"][k] < 0)
    assert (golden["traj_first_invalid"] >= 0).any() and (golden["traj_first_invalid"] < 0).any()


def test_nfabian_times_equal_reference_golden(golden):
    wp = golden["nfabian_waypoints"]
    d = np.sqrt(((wp[1:] - wp[:-1]) ** 2).sum(axis=1))
    got = worlds.nfabian_time(d, 1.0, 2.0)
    want = golden["nfabian_times"]
    # numpy's exp and sqrt-of-sum vs libm's exp and Eigen's norm(): at most 2 ulp apart
    assert np.all(np.abs(got - want) <= 2 * np.spacing(np.abs(want)))


def test_live_reference_sampler_equals_oracle(oracle, config):
    refpoly = pytest.importorskip("refpoly_py")
    if not refpoly.available():
        pytest.skip("oracle/_ref/libepp_refpoly.so not built (needs /root/reference)")
    rng = np.random.default_rng(7)
    for _ in range(500):
        c = rng.normal(size=10) * 10.0 ** rng.integers(-4, 4)
        t = rng.uniform(0.0, 8.0)
        assert oracle.poly_eval(c, t) == refpoly.horner(c, t)
    g, o, lo, hi = worlds.track_world(config)
    for K, ns in ((1, 50), (3, 333), (8, 1000)):
        coeffs, T, dt = worlds.rest_to_rest_trajectories(12, K, lo, hi, n_samples=ns, seed=1234 + K)
        for k in range(12):
            xr, tr, nr = refpoly.evaluate_range(coeffs[k], T[k], dt[k])
            xo, to, no = oracle.traj_evaluate_range(coeffs[k], T[k], 0.0, oracle.traj_max_time(T[k]), dt[k])
            assert nr == no and np.array_equal(xr, xo) and np.array_equal(tr, to)
            assert refpoly.max_time(coeffs[k], T[k]) == oracle.traj_max_time(T[k])


# ---- GPU: K3b (fused) and K3a (pre-sampled) against the reference outputs ---------------------------------

@pytest.mark.gpu
@pytest.mark.parametrize("bp", [epp_b200.BP_BRUTE, epp_b200.BP_BITS, epp_b200.BP_GRID], ids=["brute", "bits", "grid"])
def test_gpu_fused_sweep_equals_reference_golden(config, golden, bp):
    g, o, lo, hi = worlds.track_world(config)
    gw = epp_b200.World(config, broadphase=bp)
    gw.parse_gates_and_obstacles(g, o)
    coeffs, T, dt = golden["traj_coeffs"], golden["traj_T"], golden["traj_dt"]
    words = golden["traj_masks"].shape[1]
    valid, smask, ns, first = gw.check_poly_trajectories(coeffs, T, dt, float(golden["min_distance"]),
                                                         max_samples=words * 32)
    assert np.array_equal(ns.astype(np.int64), golden["traj_n"])
    assert np.array_equal(smask, golden["traj_masks"])
    assert np.array_equal(first, golden["traj_first_invalid"])
    bits = np.unpackbits(valid.view(np.uint8), bitorder="little")[: coeffs.shape[0]]
    assert np.array_equal(bits.astype(bool), golden["traj_first_invalid"] < 0)


@pytest.mark.gpu
def test_gpu_presampled_sweep_equals_reference_golden(config, golden):
    """checkTrajectoryValidity on N x 10 matrices holding the reference's own samples in columns 0, 3, 6."""
    g, o, lo, hi = worlds.track_world(config)
    gw = epp_b200.World(config)
    gw.parse_gates_and_obstacles(g, o)
    md = float(golden["min_distance"])
    for k in (0, 5, 24, 26, 31):
        n = int(golden["traj_n"][k])
        traj = np.zeros((n, 10))
        traj[:, 0], traj[:, 3], traj[:, 6] = golden["traj_samples"][k, :n].T
        traj[:, 9] = golden["traj_times"][k, :n]
        assert gw.checkTrajectoryValidity(traj, md) == (golden["traj_first_invalid"][k] < 0)
        want = golden["traj_masks"][k][: (n + 31) // 32]
        got = gw.check_states_mindist(golden["traj_samples"][k, :n], md)
        assert np.array_equal(got, want)


@pytest.mark.gpu
def test_gpu_trajectory_matrices_equal_reference_golden(golden):
    """epp_sample_trajectories = the sampling half of poly_traj::generateTrajectory: rows [px vx ax py vy ay pz vz az
    t + offset], bit-equal to the reference's three evaluateRange passes."""
    coeffs, T, dt = golden["traj_coeffs"], golden["traj_T"], golden["traj_dt"]
    mats = epp_b200.sample_trajectories(coeffs, T, dt, time_offsets=1.25, max_rows=512)
    for k, m in enumerate(mats):
        n = int(golden["traj_n"][k])
        assert m.shape == (n, 10), k
        for d in range(3):
            assert np.array_equal(m[:, 3 * d], golden["traj_samples"][k, :n, d]), (k, d)
            assert np.array_equal(m[:, 3 * d + 1], golden["traj_vel"][k, :n, d]), (k, d)
            assert np.array_equal(m[:, 3 * d + 2], golden["traj_acc"][k, :n, d]), (k, d)
        assert np.array_equal(m[:, 9], golden["traj_times"][k, :n] + 1.25), k
    with pytest.raises(epp_b200.EppError):
        epp_b200.sample_trajectories(coeffs, T, dt, max_rows=64)  # the trajectories have ~200 rows
