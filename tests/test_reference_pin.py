"""Pins the CPU oracle (and, on the GPU box, the CUDA path) against THE REFERENCE'S OWN CODE.

Two sources, both produced from /root/reference/src/{OBB,Object,World,StateValidator,MotionValidator}.cpp
compiled unmodified against oracle/shim (third-party headers only):
  * tests/golden/reference_golden.npz — committed outputs (tests/golden/make_reference_golden.py);
  * oracle/_ref/libepp_ref.so         — the same code, live, when it has been built (it travels to the
                                          GPU box prebuilt; it cannot be rebuilt there).
"""
import os

import numpy as np
import pytest

import epp_b200
from epp_b200 import worlds
from helpers import adversarial_points, adversarial_segments, filling_config

HERE = os.path.dirname(os.path.abspath(__file__))
MDS = (0.05, 0.1, 0.2, 0.35)


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "reference_golden.npz"))


def cases(config):
    g, o, lo, hi = worlds.track_world(config)
    yield "track", config, g, o
    yield "filling", filling_config(config), g, o
    g, o, lo, hi = worlds.world_256()
    yield "w256", config, g, o


def check_world_against_golden(z, name, w, states_fn, mindist_fn, seg_fn):
    states, a, b = z[name + "_states"], z[name + "_seg_a"], z[name + "_seg_b"]
    for cpg in (0, 1):
        assert np.array_equal(states_fn(w, states, bool(cpg)), z["%s_states_cpg%d" % (name, cpg)]), (name, cpg)
        assert np.array_equal(seg_fn(w, a, b, bool(cpg)), z["%s_seg_cpg%d" % (name, cpg)]), (name, cpg)
    for md in MDS:
        assert np.array_equal(mindist_fn(w, states, md), z["%s_states_md%03d" % (name, int(md * 100))]), (name, md)


def test_oracle_matches_reference_golden_masks(oracle, config, golden):
    for name, cfg, g, o in cases(config):
        ow = oracle.OracleWorld(cfg)
        ow.parse_gates_and_obstacles(g, o)
        check_world_against_golden(golden, name, ow,
                                   lambda w, s, c: w.check_states(s, c, threads=4),
                                   lambda w, s, md: w.check_states_mindist(s, md, threads=4),
                                   lambda w, a, b, c: w.check_segments(a, b, c, threads=4))
    # the fixture is not trivially all-valid / all-invalid
    n = golden["track_states"].shape[0]
    pop = int(np.unpackbits(golden["track_states_cpg0"].view(np.uint8)).sum())
    assert 0.3 * n < pop < 0.99 * n


def test_oracle_and_host_world_records_match_reference_golden(oracle, config, golden):
    """Object::createFromDescription / rotateZ / OBB::getAABB bit for bit (center, halfSize, rotation,
    stored AABB), for both portals and the obstacle at eight yaws."""
    want = golden["records"]
    got_o, got_h = [], []
    for yaw in golden["records_yaws"]:
        for is_gate, tid in ((1, 0), (1, 1), (0, 0)):
            ow = oracle.OracleWorld(config)
            hw = epp_b200.World(config)
            if is_gate:
                pose = [0.4, -0.7, 0.0, 0.0, 0.0, float(yaw), tid]
                ow.add_gate(0, pose)
                hw.addGate(0, pose)
            else:
                pose = [0.4, -0.7, 0.0, 0.0, 0.0, float(yaw)]
                ow.add_obstacle(0, pose)
                hw.addObstacle(0, pose)
            for recs, out in ((ow.obbs(), got_o), (hw.obbs(), got_h)):
                for r in sorted(recs, key=lambda r: r["obb_index"]):
                    out.append(list(r["center"]) + list(r["half"]) + list(r["rot"]) + list(r["aabb_min"]) +
                               list(r["aabb_max"]))
    assert np.array_equal(np.array(got_o), want)
    assert np.array_equal(np.array(got_h), want)


# ---- live reference ------------------------------------------------------------------------------------

@pytest.fixture(scope="module")
def ref():
    import ref_py
    if not ref_py.available():
        pytest.skip("oracle/_ref/libepp_ref.so not built (needs /root/reference; run make -C oracle ref)")
    return ref_py


@pytest.mark.parametrize("via_validators", [False, True], ids=["World", "Validators"])
def test_oracle_matches_live_reference(oracle, config, ref, via_validators):
    for name, cfg, g, o in cases(config):
        rw = ref.RefWorld(cfg, via_validators=via_validators)
        rw.parse_gates_and_obstacles(g, o)
        ow = oracle.OracleWorld(cfg)
        ow.parse_gates_and_obstacles(g, o)
        lo = np.array([-2.5, -2.5, -0.2]) if name != "w256" else np.array([-21.0, -21.0, -0.2])
        hi = -lo
        hi[2] = 2.5 if name != "w256" else 4.5
        rng = np.random.default_rng(77)
        states = np.concatenate([worlds.gen_states(0xABC, 0, 60000, lo, hi),
                                 adversarial_points(ow.obbs(), rng, per_face=1 if name != "w256" else 0 or 1)[:60000]])
        a, b = worlds.gen_segments(0xABD, 0, 30000, lo, hi, 1.5)
        aa, ab = adversarial_segments(ow.obbs(), rng, lo, hi, per_obb=2)
        a, b = np.concatenate([a, aa]), np.concatenate([b, ab])
        for cpg in (False, True):
            assert np.array_equal(ow.check_states(states, cpg, threads=4), rw.check_states(states, cpg, threads=4))
            assert np.array_equal(ow.check_segments(a, b, cpg, threads=4), rw.check_segments(a, b, cpg, threads=4))
        for md in (0.0, 0.1, 0.3):
            assert np.array_equal(ow.check_states_mindist(states, md, threads=4),
                                  rw.check_states_mindist(states, md, threads=4))


def test_world_mutation_and_errors_match_live_reference(oracle, config, ref):
    g, o, lo, hi = worlds.track_world(config)
    rw, ow = ref.RefWorld(config), oracle.OracleWorld(config)
    rw.parse_gates_and_obstacles(g, o)
    ow.parse_gates_and_obstacles(g, o)
    xyz = worlds.gen_states(5, 0, 30000, lo, hi)
    new = [0.1, 0.2, 0, 0, 0, 1.0, 1]
    rw.update_gate(2, new)      # World::updateGatePosition: remove old R-tree values, insert new
    ow.update_gate(2, new)
    assert np.array_equal(ow.check_states(xyz), rw.check_states(xyz))
    rw.add_gate(1, [1.5, 1.5, 0, 0, 0, 0.5, 0])   # duplicate id: map keeps the first OBBs, R-tree gets both AABBs
    ow.add_gate(1, [1.5, 1.5, 0, 0, 0, 0.5, 0])
    assert np.array_equal(ow.check_states(xyz), rw.check_states(xyz))
    a, b = worlds.gen_segments(6, 0, 20000, lo, hi, 1.0)
    assert np.array_equal(ow.check_segments(a, b), rw.check_segments(a, b))
    for pose, msg in (([0, 0, 0, 0.1, 0, 0, 1], "Rotation around x axis is not supported"),
                      ([0, 0, 0, 0, 0.1, 0, 1], "Rotation around y axis is not supported")):
        with pytest.raises(RuntimeError, match=msg):
            rw.add_gate(9, pose)
        with pytest.raises(Exception, match=msg):
            ow.add_gate(9, pose)
    with pytest.raises(RuntimeError, match="Center z position must be zero"):
        rw.add_obstacle(9, [0, 0, 0.5, 0, 0, 0])
    with pytest.raises(RuntimeError):
        rw.update_gate(42, new)   # obbs.at() on an unknown key


# ---- GPU path against the reference ----------------------------------------------------------------------

@pytest.mark.gpu
@pytest.mark.parametrize("bp", [epp_b200.BP_BRUTE, epp_b200.BP_BITS, epp_b200.BP_GRID], ids=["brute", "bits", "grid"])
def test_gpu_matches_reference_golden_masks(config, golden, bp):
    for name, cfg, g, o in cases(config):
        gw = epp_b200.World(cfg, broadphase=bp)
        gw.parse_gates_and_obstacles(g, o)
        gw.set_broadphase(bp)
        check_world_against_golden(golden, name, gw,
                                   lambda w, s, c: w.check_states(s, c),
                                   lambda w, s, md: w.check_states_mindist(s, md),
                                   lambda w, a, b, c: w.check_segments(a, b, c))


@pytest.mark.gpu
def test_gpu_matches_live_reference(config, ref):
    """Larger seeded batches than the committed fixture, straight against the reference's own code."""
    for name, cfg, g, o in cases(config):
        rw = ref.RefWorld(cfg)
        rw.parse_gates_and_obstacles(g, o)
        gw = epp_b200.World(cfg)
        gw.parse_gates_and_obstacles(g, o)
        lo, hi = (np.array([-2.0, -2.0, 0.0]), np.array([2.0, 2.0, 2.0])) if name != "w256" else \
                 (np.array([-20.0, -20.0, 0.0]), np.array([20.0, 20.0, 4.0]))
        xyz = worlds.gen_states(0xF00D, 0, 300000, lo, hi)
        a, b = worlds.gen_segments(0xF00E, 0, 100000, lo, hi, 1.0)
        for cpg in (False, True):
            assert np.array_equal(gw.check_states(xyz, cpg), rw.check_states(xyz, cpg, threads=8))
            assert np.array_equal(gw.check_segments(a, b, cpg), rw.check_segments(a, b, cpg, threads=8))
        assert np.array_equal(gw.check_states_mindist(xyz, 0.1), rw.check_states_mindist(xyz, 0.1, threads=8))
