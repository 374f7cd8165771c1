"""CPU tests of the product's host side: world construction (bit-for-bit vs the oracle), error
behaviour, C-ABI symbol coverage, sharding helper.  No compute call is made here (no GPU)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import epp_b200
from epp_b200 import worlds

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REC_KEYS = ("center", "half", "rot", "aabb_min", "aabb_max", "inflate", "filling", "is_gate",
            "object_id", "obb_index")


def records_equal(a, b):
    assert len(a) == len(b)
    for x, y in zip(a, b):
        for k in REC_KEYS:
            xa, ya = np.asarray(x[k], dtype=np.float64), np.asarray(y[k], dtype=np.float64)
            # bit-for-bit (distinguishes -0.0 from 0.0 as well)
            assert np.array_equal(xa.view(np.uint64), ya.view(np.uint64)), (k, x[k], y[k])


def test_header_symbols_are_exported():
    """Every function include/epp_c_api.h declares is exported by libepp_b200.so and bound."""
    text = open(os.path.join(ROOT, "include", "epp_c_api.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    declared = set(re.findall(r"\b(epp_[a-z0-9_]+)\s*\(", text))
    assert len(declared) >= 35
    lib = epp_b200.lib()
    for name in sorted(declared):
        assert hasattr(lib, name), "missing export " + name
        assert name in epp_b200.SIGNATURES, "unbound symbol " + name
    assert set(epp_b200.SIGNATURES) == declared


def test_header_constants_match_the_binding():
    """Limits the header states are the ones the Python layer repeats."""
    text = open(os.path.join(ROOT, "include", "epp_c_api.h")).read()
    m = re.search(r"#define\s+EPP_STATIC_MAX_TRAJECTORIES\s+(\d+)", text)
    assert m and int(m.group(1)) == epp_b200.STATIC_MAX_TRAJECTORIES


def test_epp_tune_environment_sets_knobs():
    """EPP_TUNE="knob=value,..." is applied when the library is loaded; an unknown knob is an error, not ignored."""
    import subprocess
    import sys
    code = "import sys; sys.path.insert(0, %r); import epp_b200; print(epp_b200.get_tuning('poly_ring'), epp_b200.get_tuning('poly_wait'))" % ROOT
    env = dict(os.environ, EPP_TUNE="poly_ring=3, poly_wait=1")
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    assert out.returncode == 0 and out.stdout.split() == ["3", "1"], out.stderr
    env = dict(os.environ, EPP_TUNE="no_such_knob=1")
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    assert out.returncode != 0 and "no_such_knob" in out.stderr


@pytest.mark.parametrize("which", ["track", "w256", "w10k"])
def test_world_records_match_oracle_bit_for_bit(oracle, config, which):
    if which == "track":
        g, o, _, _ = worlds.track_world(config)
    elif which == "w256":
        g, o, _, _ = worlds.world_256()
    else:
        g, o, _, _ = worlds.world_10k()
    ow = oracle.OracleWorld(config)
    ow.parse_gates_and_obstacles(g, o)
    gw = epp_b200.World(config)
    for i, x in enumerate(g):
        gw.addGate(i, x)
    for i, x in enumerate(o):
        gw.addObstacle(i, x)
    expect = {"track": 24, "w256": 256, "w10k": 10000}[which]
    assert gw.size() == ow.size() == expect
    records_equal(gw.obbs(), ow.obbs())


def test_update_and_duplicate_quirks_match_oracle(oracle, config):
    ow = oracle.OracleWorld(config)
    gw = epp_b200.World(config)
    for w, add_g, upd_g, add_o in ((ow, ow.add_gate, ow.update_gate, ow.add_obstacle),
                                   (gw, gw.addGate, gw.updateGatePosition, gw.addObstacle)):
        add_g(0, [0.3, -0.2, 0, 0, 0, 0.4, 1])
        add_g(1, [1.0, 1.0, 0, 0, 0, -2.0, 0])
        add_o(0, [0.5, 0.5, 0, 0, 0, 0.1])
        add_o(0, [0.6, 0.5, 0, 0, 0, 0.2])          # duplicate id: first OBB kept, new AABB added
        upd_g(0, [-0.7, 0.1, 0, 0, 0, 1.1, 1])
    key = lambda r: (r["is_gate"], r["object_id"], r["obb_index"], tuple(r["aabb_min"]))
    records_equal(sorted(gw.obbs(), key=key), sorted(ow.obbs(), key=key))
    assert gw.size() == 12


def test_error_messages_follow_reference(config):
    w = epp_b200.World(config)
    with pytest.raises(epp_b200.EppError, match="Rotation around x axis is not supported"):
        w.addGate(0, [0, 0, 0, 0.1, 0, 0, 1])
    with pytest.raises(epp_b200.EppError, match="Rotation around y axis is not supported"):
        w.addObstacle(0, [0, 0, 0, 0, -0.1, 0])
    with pytest.raises(epp_b200.EppError, match="Center z position must be zero"):
        w.addObstacle(0, [0, 0, 0.2, 0, 0, 0])
    with pytest.raises(epp_b200.EppError) as e:
        w.addGate(0, [0, 0, 0, 0, 0, 0, 9])         # unknown gate type
    assert e.value.status == 4
    with pytest.raises(epp_b200.EppError) as e:
        w.updateGatePosition(3, [0, 0, 0, 0, 0, 0, 1])
    assert e.value.status == 5
    w.addGate(0, [0, 0, 5.0, 1e-7, -1e-7, 0, 1.9])  # z forced to 0, tiny roll/pitch ok, type truncates to 1
    assert w.size() == 5
    w.resetWorld()
    assert w.size() == 0


def test_check_without_gpu_fails_loudly(config):
    """No CPU fallback: without a device every check raises (on the GPU box this test is moot)."""
    try:
        n = epp_b200.device_count()
    except epp_b200.EppError:
        n = 0
    if n > 0:
        pytest.skip("a GPU is present")
    w = epp_b200.World(config)
    w.addObstacle(0, [0, 0, 0, 0, 0, 0])
    with pytest.raises(epp_b200.EppError) as e:
        w.check_states(np.zeros((4, 3)))
    assert e.value.status in (8, 9)
    with pytest.raises(epp_b200.EppError):
        w.checkPointValidity([0, 0, 0], False)


def test_uncommitted_world_is_rejected(config):
    lib = epp_b200.lib()
    h = C.c_void_p()
    assert lib.epp_world_create(0.2, 0.2, C.byref(h)) == 0
    mask = np.zeros(1, dtype=np.uint32)
    xyz = np.zeros((1, 3))
    assert lib.epp_check_states(h, xyz.ctypes.data, 1, 0, mask.ctypes.data) == 7
    assert b"commit" in lib.epp_last_error()
    assert lib.epp_check_states_dev(h, None, 0, 0, None, None) == 7
    lib.epp_world_destroy(h)
    assert lib.epp_check_states(None, xyz.ctypes.data, 1, 0, mask.ctypes.data) == 6


def test_shard_range_partitions_on_word_boundaries():
    for n in (0, 1, 31, 32, 1000, 1 << 20, 10 ** 9 + 7):
        for ws in (1, 2, 3, 4, 8):
            spans = [epp_b200.shard_range(n, r, ws, 1024) for r in range(ws)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (b0, e0), (b1, e1) in zip(spans, spans[1:]):
                assert e0 == b1
            for b, e in spans:
                assert b % 1024 == 0 or b == n
                assert b <= e
            sizes = [e - b for b, e in spans if e > b]
            if sizes:
                assert max(sizes) - min(sizes[:-1] or sizes) <= 0 or len(set(sizes[:-1])) <= 1
    with pytest.raises(epp_b200.EppError):
        epp_b200.shard_range(10, 2, 2, 1024)
    with pytest.raises(epp_b200.EppError):
        epp_b200.shard_range(10, 0, 2, 33)


def test_mask_summary_host_is_position_dependent():
    m = np.array([0xFFFF0000, 0x1, 0x0, 0x80000000], dtype=np.uint32)
    pop, x = epp_b200.mask_summary_host(m)
    assert pop == 18
    pop2, x2 = epp_b200.mask_summary_host(m[::-1].copy())
    assert pop2 == 18 and x2 != x
    # shards XOR / add up to the whole
    pa, xa = epp_b200.mask_summary_host(m[:2], 0)
    pb, xb = epp_b200.mask_summary_host(m[2:], 2)
    assert pa + pb == pop and xa ^ xb == x


def test_unpack_mask_convention():
    m = np.array([0b101, 0x80000000], dtype=np.uint32)
    b = epp_b200.unpack_mask(m, 64)
    assert b[0] and not b[1] and b[2] and b[63] and b.sum() == 3


def test_tuning_knobs():
    epp_b200.set_tuning("states_blocks_per_sm", 6)
    assert epp_b200.get_tuning("states_blocks_per_sm") == 6
    epp_b200.set_tuning("states_blocks_per_sm", 8)
    with pytest.raises(epp_b200.EppError):
        epp_b200.set_tuning("no_such_knob", 1)
