"""Generates tests/golden/track_golden.npz from the CPU oracle (oracle/libepp_oracle.so).

The reference ships no golden vectors for this path, so these are outputs of the
oracle restatement on the builder-defined track world (config "track").  They freeze the oracle:
a later change that moves the oracle and the GPU path together is caught by the committed bits.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]

import epp_b200  # noqa: E402
import oracle_py  # noqa: E402
from epp_b200 import worlds  # noqa: E402
from helpers import adversarial_points, adversarial_segments  # noqa: E402


def main():
    oracle_py.build()
    config = epp_b200.load_config()
    g, o, lo, hi = worlds.track_world(config)
    ow = oracle_py.OracleWorld(config)
    ow.parse_gates_and_obstacles(g, o)
    rng = np.random.default_rng(0x601D)
    states = np.concatenate([worlds.gen_states(worlds.SEED_STATES, 0, 20000, lo, hi),
                             adversarial_points(ow.obbs(), rng, per_face=1)])
    a, b = worlds.gen_segments(worlds.SEED_SEGMENTS, 0, 10000, lo, hi, 1.0)
    aa, ab = adversarial_segments(ow.obbs(), rng, lo, hi, per_obb=4)
    seg_a, seg_b = np.concatenate([a, aa]), np.concatenate([b, ab])
    out = dict(
        states=states,
        states_mask=ow.check_states(states, False),
        states_mask_cpg=ow.check_states(states, True),
        states_mask_md=ow.check_states_mindist(states, 0.1),
        seg_a=seg_a, seg_b=seg_b,
        seg_mask=ow.check_segments(seg_a, seg_b, False),
    )
    np.savez_compressed(os.path.join(HERE, "track_golden.npz"), **out)
    for k, v in out.items():
        print(k, v.shape, v.dtype)


if __name__ == "__main__":
    main()
