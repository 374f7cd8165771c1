"""CPU tests of the oracle (oracle/epp_oracle.c): hand-derived known answers and the semantics
checklist of SURVEY.md §8c.  The reference ships no test that pins a validity result, so these
KATs (derived by reading src/OBB.cpp / src/World.cpp) are the first anchor; the second is
oracle/_ref (tests/test_reference_build.py)."""
import json
import math
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def kat():
    with open(os.path.join(HERE, "golden", "kat_small_portal.json")) as f:
        return json.load(f)


def test_kat_stored_aabbs(oracle, config, kat):
    w = oracle.OracleWorld(config)
    w.add_gate(0, kat["gate"])
    obbs = w.obbs()
    names = list(config["component_geometry"]["small_portal"].keys())
    assert len(obbs) == 5
    for o in obbs:
        want = kat["stored_aabbs"][names[o["obb_index"]]]
        np.testing.assert_allclose(o["aabb_min"], want[0], rtol=0, atol=1e-12)
        np.testing.assert_allclose(o["aabb_max"], want[1], rtol=0, atol=1e-12)


def test_kat_points_and_segments(oracle, config, kat):
    w = oracle.OracleWorld(config)
    w.add_gate(0, kat["gate"])
    for p, want in kat["points"]:
        assert w.check_point(p, False) is want, p
    for a, b, want in kat["segments"]:
        assert w.check_ray(a, b, False) is want, (a, b)
    # batched loops agree with the single predicates
    pts = np.array([p for p, _ in kat["points"]], dtype=float)
    mask = w.check_states(pts)
    assert [bool(mask[0] >> i & 1) for i in range(len(pts))] == [v for _, v in kat["points"]]
    a = np.array([s[0] for s in kat["segments"]], dtype=float)
    b = np.array([s[1] for s in kat["segments"]], dtype=float)
    mask = w.check_segments(a, b)
    assert [bool(mask[0] >> i & 1) for i in range(len(a))] == [s[2] for s in kat["segments"]]


def test_kat_non_conservative_broad_phase(oracle, config, kat):
    """stored AABB = un-inflated rotated box +- inflate and the point query is
    strict, so a point inside the inflated OBB but outside the stored AABB is VALID."""
    import ctypes as C
    q = kat["quirk"]
    w = oracle.OracleWorld(config)
    w.add_gate(0, q["gate"])
    bottom = [o for o in w.obbs() if o["obb_index"] == 1][0]
    np.testing.assert_allclose(bottom["aabb_min"], q["bottom_bar_aabb_approx"][0], atol=1e-5)
    np.testing.assert_allclose(bottom["aabb_max"], q["bottom_bar_aabb_approx"][1], atol=1e-5)
    rec = oracle.Obb()
    oracle.lib().orc_world_get(w.h, 1, C.byref(rec))
    p = np.array(q["point"], dtype=float)
    assert bool(oracle.lib().orc_obb_point_hit(C.byref(rec), oracle._dp(p), 0.2)) is q["inside_inflated_obb"]
    assert w.check_point(p, False) is q["valid"]


def test_strict_aabb_inclusive_obb(oracle, config):
    """Point exactly on a stored-AABB face is outside the query (Boost within is strict); point
    exactly on an inflated OBB face (inside the AABB) collides (<=, src/OBB.cpp:75-77)."""
    w = oracle.OracleWorld(config)
    w.add_obstacle(0, [0, 0, 0, 0, 0, 0])  # box 0.1x0.1x1.05 at z=0.525, inflate 0.2 -> H = (0.25,0.25,0.725)
    o = w.obbs()[0]
    assert o["aabb_max"][0] == 0.05 + 0.2
    assert w.check_point([o["aabb_max"][0], 0.0, 0.5]) is True          # on the AABB face: not a candidate
    inside = np.nextafter(o["aabb_max"][0], 0.0)
    assert w.check_point([inside, 0.0, 0.5]) is False                  # strictly inside AABB, |local| <= H
    # yawed obstacle: OBB face reachable strictly inside the (larger) AABB
    w.reset()
    w.add_obstacle(0, [0, 0, 0, 0, 0, 0.3])
    o = w.obbs()[0]
    R = np.array(o["rot"]).reshape(3, 3)
    H = np.array(o["half"]) + 0.2
    on_face = R @ np.array([H[0], 0.0, 0.0]) + np.array(o["center"])
    beyond = R @ np.array([H[0] * (1 + 1e-9), 0.0, 0.0]) + np.array(o["center"])
    assert w.check_point(beyond) is True
    assert w.check_point(R @ np.array([H[0] * (1 - 1e-9), 0.0, 0.0]) + np.array(o["center"])) is False
    assert isinstance(w.check_point(on_face), bool)


def test_filling_semantics(oracle, config):
    """filling boxes: never inflated in the point test, inflated in the ray slab, skipped when
    canPassGate (bool overload) and always in the minDistance overload."""
    from helpers import filling_config
    cfg = filling_config(config)
    w = oracle.OracleWorld(cfg)
    w.add_gate(0, [0, 0, 0, 0, 0, 0, 1])
    centre = [0.0, 0.0, 0.525]
    assert w.check_point(centre, False) is False      # inside the filling box
    assert w.check_point(centre, True) is True        # canPassGate skips it
    assert w.check_point_mindist(centre, 0.1) is True # minDistance overload always skips filling
    # filling AABB is not inflated: y = 0.03 is outside its stored AABB (half 0.025)
    assert w.check_point([0.0, 0.03, 0.525], False) is True
    # ray through the opening: blocked unless canPassGate
    assert w.check_ray([0, -1, 0.525], [0, 1, 0.525], False) is False
    assert w.check_ray([0, -1, 0.525], [0, 1, 0.525], True) is True


def test_mindist_keeps_config_inflated_aabb(oracle, config):
    """minDistance > config inflate is clipped by the stored AABB (src/World.cpp:106-128)."""
    w = oracle.OracleWorld(config)
    w.add_obstacle(0, [0, 0, 0, 0, 0, 0])
    p = [0.3, 0.0, 0.5]                      # outside stored AABB (0.25) but within half+0.5
    assert w.check_point_mindist(p, 0.5) is True
    assert w.check_point_mindist([0.2, 0.0, 0.5], 0.5) is False
    assert w.check_point_mindist([0.2, 0.0, 0.5], 0.1) is True   # H = 0.15 < 0.2
    assert w.check_point([0.2, 0.0, 0.5], False) is False


def test_world_errors(oracle, config):
    w = oracle.OracleWorld(config)
    with pytest.raises(RuntimeError, match="x axis"):
        w.add_gate(0, [0, 0, 0, 1e-3, 0, 0, 1])
    with pytest.raises(RuntimeError, match="y axis"):
        w.add_obstacle(0, [0, 0, 0, 0, 1e-3, 0])
    with pytest.raises(RuntimeError, match="z position"):
        w.add_obstacle(0, [0, 0, 0.5, 0, 0, 0])
    w.add_gate(0, [0, 0, 3.0, 0, 0, 0, 1])   # gate z is forced to 0 (src/World.cpp:16)
    assert w.size() == 5
    with pytest.raises(RuntimeError):
        w.update_gate(7, [0, 0, 0, 0, 0, 0, 1])  # obbs.at() throws for an unknown gate


def test_update_gate_replaces_rtree_values(oracle, config):
    w = oracle.OracleWorld(config)
    w.add_gate(0, [0, 0, 0, 0, 0, 0, 1])
    assert w.check_point([0, 0, 0.4]) is False
    w.update_gate(0, [1.0, 0, 0, 0, 0, 0, 1])
    assert w.size() == 5
    assert w.check_point([0, 0, 0.4]) is True
    assert w.check_point([1.0, 0, 0.4]) is False


def test_duplicate_add_keeps_first_obb(oracle, config):
    """std::map::insert keeps the existing OBB while the R-tree gains the new AABB
    (src/World.cpp:63-65): the second add of the same id tests the OLD box inside the NEW AABB."""
    w = oracle.OracleWorld(config)
    w.add_obstacle(0, [0, 0, 0, 0, 0, 0])
    w.add_obstacle(0, [0.1, 0, 0, 0, 0, 0])
    assert w.size() == 2
    obbs = w.obbs()
    assert obbs[0]["center"] == obbs[1]["center"]          # both resolve to the first OBB
    assert obbs[0]["aabb_min"] != obbs[1]["aabb_min"]


def test_polynomial_horner(oracle):
    c = np.array([1.0, -2.0, 0.5, 0.25, 0, 0, 0, 0, 0, 3.0])
    t = 0.7
    want = c[9]
    for j in range(8, -1, -1):
        want = want * t + c[j]
    assert oracle.poly_eval(c, t) == want
    assert oracle.poly_eval(c, 0.0) == 1.0


def test_evaluate_range_running_sums(oracle):
    """Sample times are running sums acc += dt, not k*dt, and the segment switch subtracts the
    segment time (external/poly_traj/src/trajectory.cpp:121-140)."""
    K, N = 3, 10
    coeffs = np.zeros((K, 3, N))
    coeffs[:, 0, 1] = 1.0                    # x(t) = t within each segment
    coeffs[1, 0, 0] = 10.0
    coeffs[2, 0, 0] = 20.0
    T = np.array([0.35, 0.25, 0.4])
    dt = 0.1
    xyz, times, n = oracle.traj_evaluate_range(coeffs, T, 0.0, oracle.traj_max_time(T), dt)
    acc, tin, seg, exp_t, exp_x = 0.0, 0.0, 0, [], []
    t_end = (T[0] + T[1]) + T[2]
    while acc < t_end:
        if tin > T[seg]:
            tin = tin - T[seg]
            seg += 1
            if seg >= K:
                break
            continue
        exp_t.append(acc)
        exp_x.append(coeffs[seg, 0, 0] + tin)
        tin += dt
        acc += dt
    assert n == len(exp_t)
    assert list(times) == exp_t
    np.testing.assert_allclose(xyz[:, 0], exp_x, rtol=0, atol=1e-15)
    assert times[8] == 0.7999999999999999 and times[8] != 8 * 0.1   # running sum, not k*dt
    assert n == 11 and times[10] == 0.9999999999999999              # 10 steps of 0.1 stay below 1.0


def test_generators_match_numpy_twins(oracle):
    import epp_b200
    lo, hi = [-2.0, -2.0, 0.0], [2.0, 2.0, 2.0]
    a = oracle.gen_states(epp_b200.worlds.SEED_STATES, 1000, 4096, lo, hi)
    b = epp_b200.worlds.gen_states(epp_b200.worlds.SEED_STATES, 1000, 4096, lo, hi)
    assert np.array_equal(a, b)
    assert a.min() >= -2.0 and a.max() < 2.0
    sa, sb = oracle.gen_segments(epp_b200.worlds.SEED_SEGMENTS, 5, 4096, lo, hi, 1.0)
    ta, tb = epp_b200.worlds.gen_segments(epp_b200.worlds.SEED_SEGMENTS, 5, 4096, lo, hi, 1.0)
    assert np.array_equal(sa, ta) and np.array_equal(sb, tb)
    assert np.linalg.norm(sb - sa, axis=1).max() <= 1.0 + 1e-12


def test_check_trajectory_early_exit_equivalence(oracle, config):
    import epp_b200
    w = oracle.OracleWorld(config)
    g, o, lo, hi = epp_b200.worlds.track_world(config)
    w.parse_gates_and_obstacles(g, o)
    xyz = oracle.gen_states(7, 0, 300, lo, hi)
    traj = epp_b200.worlds.traj_matrix_from_positions(xyz, np.arange(300) * 0.1)
    ok, mask, first = w.check_trajectory(traj, 0.1)
    single = [w.check_point_mindist(p, 0.1) for p in xyz]
    assert ok is all(single)
    assert first == (single.index(False) if not all(single) else -1)
    assert [bool(mask[i >> 5] >> (i & 31) & 1) for i in range(300)] == single


def test_multithreaded_batches_equal_single_thread(oracle, config):
    import epp_b200
    w = oracle.OracleWorld(config)
    g, o, lo, hi = epp_b200.worlds.track_world(config)
    w.parse_gates_and_obstacles(g, o)
    xyz = oracle.gen_states(3, 0, 20001, lo, hi)
    assert np.array_equal(w.check_states(xyz, threads=1), w.check_states(xyz, threads=4))
    a, b = oracle.gen_segments(3, 0, 20001, lo, hi, 1.0)
    assert np.array_equal(w.check_segments(a, b, threads=1), w.check_segments(a, b, threads=4))
    m = w.check_states(xyz)
    frac_invalid = 1.0 - np.unpackbits(m.view(np.uint8)).sum() / 20001
    assert 0.02 < frac_invalid < 0.5   # the track blocks a meaningful share of the volume
