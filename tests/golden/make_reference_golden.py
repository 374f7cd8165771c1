"""Generates tests/golden/reference_golden.npz by RUNNING THE REFERENCE'S OWN CODE in this container:
oracle/_ref/libepp_ref.so = /root/reference/src/{OBB,Object,World,StateValidator,MotionValidator}.cpp
compiled unmodified against the stand-in headers of oracle/shim (make -C oracle ref).  The reference ships
no golden vectors for this path; these are its outputs on seeded inputs, committed so the
oracle and the GPU path can be checked against the reference on boxes where /root/reference is absent.

    python tests/golden/make_reference_golden.py

Contents: three worlds (track M=24, track + filling boxes M=28, 256-OBB lattice), uniform + face-adjacent
adversarial states and segments, every predicate variant; world records of single objects at several yaws.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]

import epp_b200  # noqa: E402
import oracle_py  # noqa: E402
import ref_py  # noqa: E402
from epp_b200 import worlds  # noqa: E402
from helpers import adversarial_points, adversarial_segments, filling_config  # noqa: E402

YAWS = [0.0, 0.3, -0.78, 0.7853981633974483, 1.5707963267948966, 2.35, 3.14, -3.0]


def world_cases(config):
    g, o, lo, hi = worlds.track_world(config)
    yield "track", config, g, o, lo, hi, 12000, 6000, 1, 3
    yield "filling", filling_config(config), g, o, lo, hi, 6000, 3000, 1, 2
    g, o, lo, hi = worlds.world_256()
    yield "w256", config, g, o, lo, hi, 12000, 6000, 0, 1


def main():
    assert ref_py.available(), "build oracle/_ref first: make -C oracle ref"
    oracle_py.build()
    config = epp_b200.load_config()
    out = {}
    for name, cfg, g, o, lo, hi, ns, ne, per_face, per_obb in world_cases(config):
        rw = ref_py.RefWorld(cfg, via_validators=True)   # through StateValidator / MotionValidator
        rw.parse_gates_and_obstacles(g, o)
        ow = oracle_py.OracleWorld(cfg)                   # only used to place adversarial inputs at faces
        ow.parse_gates_and_obstacles(g, o)
        rng = np.random.default_rng(0x601D)
        states = worlds.gen_states(0x601D0001, 0, ns, lo, hi)
        if per_face:
            states = np.concatenate([states, adversarial_points(ow.obbs(), rng, per_face=per_face)])
        a, b = worlds.gen_segments(0x601D0002, 0, ne, lo, hi, 1.0)
        aa, ab = adversarial_segments(ow.obbs(), rng, lo, hi, per_obb=per_obb)
        a, b = np.concatenate([a, aa]), np.concatenate([b, ab])
        out[name + "_states"] = states
        out[name + "_seg_a"], out[name + "_seg_b"] = a, b
        for cpg in (0, 1):
            out["%s_states_cpg%d" % (name, cpg)] = rw.check_states(states, bool(cpg))
            out["%s_seg_cpg%d" % (name, cpg)] = rw.check_segments(a, b, bool(cpg))
        for md in (0.05, 0.1, 0.2, 0.35):
            out["%s_states_md%03d" % (name, int(md * 100))] = rw.check_states_mindist(states, md)
    # world records straight from Object::createFromDescription / OBB::getAABB
    rw = ref_py.RefWorld(config)
    recs = []
    for yaw in YAWS:
        for is_gate, tid in ((1, 0), (1, 1), (0, 0)):
            for r in rw.object_records(is_gate, tid, [0.4, -0.7, 0.0], [0.0, 0.0, yaw], 0.2):
                recs.append(r["center"] + r["half"] + r["rot"] + r["aabb_min"] + r["aabb_max"])
    out["records_yaws"] = np.array(YAWS)
    out["records"] = np.array(recs)
    np.savez_compressed(os.path.join(HERE, "reference_golden.npz"), **out)
    print("wrote reference_golden.npz:", {k: v.shape for k, v in out.items() if "states_cpg0" in k or k == "records"})


if __name__ == "__main__":
    main()
