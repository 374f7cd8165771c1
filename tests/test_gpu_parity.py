"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU oracle on the same
inputs.  Bit-exact masks are required everywhere (integer/bit outputs of FP64 predicates with FMA
contraction disabled); the north star's 1e-9 m allowance is not used: the expected mismatch count
is zero and the tests assert zero."""
import os

import numpy as np
import pytest

import epp_b200
from epp_b200 import worlds
from helpers import adversarial_points, adversarial_segments, build_pair, filling_config

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
BPS = [epp_b200.BP_BRUTE, epp_b200.BP_BITS, epp_b200.BP_GRID]
BP_IDS = ["brute", "bits", "grid"]


def mismatches(a, b):
    return int(np.unpackbits((np.asarray(a, np.uint32) ^ np.asarray(b, np.uint32)).view(np.uint8)).sum())


@pytest.fixture(scope="module")
def track(oracle, config):
    g, o, lo, hi = worlds.track_world(config)
    return g, o, lo, hi


@pytest.mark.parametrize("bp", BPS, ids=BP_IDS)
def test_kat_through_c_abi(config, bp):
    import json
    kat = json.load(open(os.path.join(HERE, "golden", "kat_small_portal.json")))
    w = epp_b200.World(config, broadphase=bp)
    w.addGate(0, kat["gate"])
    for p, want in kat["points"]:
        assert w.checkPointValidity(p, False) is want, p
    for a, b, want in kat["segments"]:
        assert w.checkRayValid(a, b, False) is want, (a, b)
    w.resetWorld()
    w.addGate(0, kat["quirk"]["gate"])
    assert w.checkPointValidity(kat["quirk"]["point"], False) is kat["quirk"]["valid"]
    assert w.active_broadphase() == bp


@pytest.mark.parametrize("bp", BPS, ids=BP_IDS)
def test_config2_states_track(oracle, config, track, bp):
    """BASELINE config 2 at a size the oracle finishes in seconds: uniform states + adversarial."""
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o, bp)
    n = 1 << 18
    xyz = worlds.gen_states(worlds.SEED_STATES, 0, n, lo, hi)
    adv = adversarial_points(ow.obbs(), np.random.default_rng(2))
    xyz = np.concatenate([xyz, adv])
    for cpg in (False, True):
        assert mismatches(gw.check_states(xyz, cpg), ow.check_states(xyz, cpg, threads=8)) == 0
    for md in (0.05, 0.1, 0.2, 0.35):
        assert mismatches(gw.check_states_mindist(xyz, md), ow.check_states_mindist(xyz, md, threads=8)) == 0


@pytest.mark.parametrize("bp", BPS, ids=BP_IDS)
def test_filling_boxes(oracle, config, track, bp):
    g, o, lo, hi = track
    cfg = filling_config(config)
    ow, gw = build_pair(oracle, epp_b200, cfg, g, o, bp)
    assert ow.size() == 28
    xyz = np.concatenate([worlds.gen_states(11, 0, 50000, lo, hi),
                          adversarial_points(ow.obbs(), np.random.default_rng(3))])
    for cpg in (False, True):
        assert mismatches(gw.check_states(xyz, cpg), ow.check_states(xyz, cpg, threads=8)) == 0
    assert mismatches(gw.check_states_mindist(xyz, 0.1), ow.check_states_mindist(xyz, 0.1, threads=8)) == 0
    a, b = worlds.gen_segments(12, 0, 50000, lo, hi, 1.0)
    aa, ab = adversarial_segments(ow.obbs(), np.random.default_rng(4), lo, hi)
    a, b = np.concatenate([a, aa]), np.concatenate([b, ab])
    for cpg in (False, True):
        assert mismatches(gw.check_segments(a, b, cpg), ow.check_segments(a, b, cpg, threads=8)) == 0


@pytest.mark.parametrize("bp", BPS, ids=BP_IDS)
def test_config3_segments_256(oracle, config, bp):
    g, o, lo, hi = worlds.world_256()
    ow, gw = build_pair(oracle, epp_b200, config, g, o, bp)
    assert gw.size() == 256
    n = 1 << 17
    a, b = worlds.gen_segments(worlds.SEED_SEGMENTS, 0, n, lo, hi, 1.0)
    aa, ab = adversarial_segments(ow.obbs(), np.random.default_rng(5), lo, hi, per_obb=8)
    a, b = np.concatenate([a, aa]), np.concatenate([b, ab])
    got, want = gw.check_segments(a, b, False), ow.check_segments(a, b, False, threads=8)
    assert mismatches(got, want) == 0
    frac = 1.0 - np.unpackbits(want.view(np.uint8)).sum() / a.shape[0]
    assert frac > 0.001
    # long segments spanning many bins / cells
    la, _ = worlds.gen_segments(99, 0, 4096, lo, hi, 1.0)
    lb = worlds.gen_states(98, 0, 4096, lo, hi)
    assert mismatches(gw.check_segments(la, lb, True), ow.check_segments(la, lb, True, threads=8)) == 0
    xyz = np.concatenate([worlds.gen_states(97, 0, 1 << 16, lo, hi),
                          adversarial_points(ow.obbs(), np.random.default_rng(6), per_face=2)])
    assert mismatches(gw.check_states(xyz, False), ow.check_states(xyz, False, threads=8)) == 0


@pytest.mark.parametrize("n", [0, 1, 31, 32, 33, 255, 256, 257, 1000, 4097])
def test_ragged_sizes(oracle, config, track, n):
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    xyz = worlds.gen_states(5, 0, n, lo, hi)
    got = gw.check_states(xyz)
    assert got.shape[0] == (n + 31) // 32
    assert np.array_equal(got, ow.check_states(xyz))
    a, b = worlds.gen_segments(6, 0, n, lo, hi, 1.0)
    assert np.array_equal(gw.check_segments(a, b), ow.check_segments(a, b))
    if n % 32:
        assert got[-1] >> (n % 32) == 0        # tail bits are zero


def test_outside_bounds_and_empty_world(oracle, config, track):
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    far = worlds.gen_states(8, 0, 10000, [-50, -50, -50], [50, 50, 50])
    assert np.array_equal(gw.check_states(far), ow.check_states(far))
    fa, fb = worlds.gen_segments(9, 0, 10000, [-50, -50, -50], [50, 50, 50], 80.0)
    assert np.array_equal(gw.check_segments(fa, fb), ow.check_segments(fa, fb))
    for bp in BPS:
        e = epp_b200.World(config, broadphase=bp)
        assert e.size() == 0
        m = e.check_states(far[:100])
        assert np.unpackbits(m.view(np.uint8)).sum() == 100       # everything valid in an empty world
        assert np.unpackbits(e.check_segments(fa[:100], fb[:100]).view(np.uint8)).sum() == 100


def test_golden_fixture(config):
    """Committed oracle outputs (tests/golden/make_golden.py): guards the GPU path and the oracle
    against drifting together."""
    z = np.load(os.path.join(HERE, "golden", "track_golden.npz"))
    g, o, lo, hi = worlds.track_world(config)
    w = epp_b200.World(config)
    w.parse_gates_and_obstacles(g, o)
    assert np.array_equal(w.check_states(z["states"], False), z["states_mask"])
    assert np.array_equal(w.check_states(z["states"], True), z["states_mask_cpg"])
    assert np.array_equal(w.check_states_mindist(z["states"], 0.1), z["states_mask_md"])
    assert np.array_equal(w.check_segments(z["seg_a"], z["seg_b"], False), z["seg_mask"])


def test_world_update_epochs(oracle, config, track):
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    xyz = worlds.gen_states(21, 0, 40000, lo, hi)
    assert np.array_equal(gw.check_states(xyz), ow.check_states(xyz))
    e0 = gw.epoch()
    new = [0.1, 0.2, 0, 0, 0, 1.0, 1]
    ow.update_gate(2, new)
    gw.updateGatePosition(2, new)
    assert np.array_equal(gw.check_states(xyz), ow.check_states(xyz))
    assert gw.epoch() == e0 + 1
    g2 = g.copy()
    g2[1, :2] += 0.3
    ow.parse_gates_and_obstacles(g2, o)
    gw.parse_gates_and_obstacles(g2, o)
    assert np.array_equal(gw.check_states(xyz), ow.check_states(xyz))


@pytest.mark.parametrize("bp", BPS, ids=BP_IDS)
def test_presampled_trajectories(oracle, config, bp):
    g, o, lo, hi = worlds.world_256()
    ow, gw = build_pair(oracle, epp_b200, config, g, o, bp)
    n_traj, rows = 37, 333
    trajs = np.zeros((n_traj, rows, 10))
    for t in range(n_traj):
        xyz = worlds.gen_states(100 + t, 0, rows, lo, hi)
        if t % 3 == 0:          # keep some trajectories collision free: fly high
            xyz[:, 2] = 3.5
        trajs[t] = worlds.traj_matrix_from_positions(xyz, np.arange(rows) * 0.1)
    valid, smask, first = gw.check_trajectories(trajs, 0.1)
    n_valid = 0
    for t in range(n_traj):
        ok, m, f = ow.check_trajectory(trajs[t], 0.1)
        assert bool(valid[t >> 5] >> (t & 31) & 1) is ok
        assert np.array_equal(smask[t], m)
        assert first[t] == f
        n_valid += ok
    assert 0 < n_valid < n_traj
    assert gw.checkTrajectoryValidity(trajs[0], 0.1) is ow.check_trajectory(trajs[0], 0.1)[0]
    # N x 11 layout of the time-optimal back-end (position columns unchanged)
    t11 = np.concatenate([trajs[:2], np.zeros((2, rows, 1))], axis=2)
    v11, _, _ = gw.check_trajectories(t11, 0.1)
    assert (v11[0] & 3) == (valid[0] & 3)


@pytest.mark.parametrize("shape", [(9, 1000, 10), (200, 31, 10), (40, 129, 11), (3, 4099, 10), (70, 64, 10)],
                         ids=lambda s: "x".join(map(str, s)))
def test_presampled_trajectories_warp_form(oracle, config, track, shape):
    """K3a's warp-pipelined form (batches of >= 4096 rows: columns 0, 3, 6 by cp.async.bulk into the K1 ring) against
    the oracle and against the direct form (tuning knob traj_variant=1): ragged last chunks, odd row counts, padded
    leading dimensions, a base address that is only 8 B aligned, 10 and 11 columns, with and without sample masks."""
    import torch
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o, epp_b200.BP_AUTO)
    n_traj, rows, cols = shape
    trajs = np.zeros((n_traj, rows, cols))
    for t in range(n_traj):
        xyz = worlds.gen_states(900 + t, 0, rows, lo, hi)
        if t % 4 == 0:
            xyz[:, 2] = 1.9          # above the gates: collision free
        if t % 4 == 1:
            xyz[: rows - 1 - (t % 7), 2] = 1.9   # only the last rows can collide
        trajs[t, :, :10] = worlds.traj_matrix_from_positions(xyz, np.arange(rows) * 0.05)
    want = [ow.check_trajectory(trajs[t, :, :10], 0.1) for t in range(n_traj)]
    try:
        results = {}
        for variant in (0, 1):
            epp_b200.set_tuning("traj_variant", variant)
            results[variant] = gw.check_trajectories(trajs, 0.1)
    finally:
        epp_b200.set_tuning("traj_variant", 0)
    for variant, (valid, smask, first) in results.items():
        for t in range(n_traj):
            ok, m, f = want[t]
            assert bool(valid[t >> 5] >> (t & 31) & 1) is ok, (variant, t)
            assert np.array_equal(smask[t], m), (variant, t)
            assert first[t] == f, (variant, t)
    assert 0 < sum(w[0] for w in want) < n_traj
    # device buffers: leading dimension rows + pad, matrices 8 B (not 16 B) aligned, no sample masks wanted
    for pad, shift in ((0, 0), (3, 0), (2, 1), (0, 1)):
        ld = rows + pad
        stride = ld * cols + (pad & 1)
        host = np.full(n_traj * stride + 2, np.nan)
        for t in range(n_traj):
            for c in range(cols):
                host[shift + t * stride + c * ld: shift + t * stride + c * ld + rows] = trajs[t, :, c]
        dev = torch.from_numpy(host).cuda()
        v, sm, f = gw.check_trajectories_dev(dev[shift:], n_traj, rows, ld, stride, cols, 0.1, want_samples=(pad != 3))
        torch.cuda.synchronize()
        assert np.array_equal(v.cpu().numpy().view(np.uint32), results[1][0]), (pad, shift)
        assert np.array_equal(f.cpu().numpy(), results[1][2]), (pad, shift)
        if sm is not None:
            assert np.array_equal(sm.cpu().numpy().view(np.uint32), results[1][1]), (pad, shift)


def test_presampled_trajectories_warp_form_ring_wrap(oracle, config, track):
    """2 000 x 1 001 rows: every block takes ~110 chunks through a ring of at most 64 slots (several wraps), the last
    chunk of every trajectory is ragged and odd.  Warp form == direct form on all of it, == oracle on a subsample."""
    import torch
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o, epp_b200.BP_AUTO)
    n_traj, rows = 2000, 1001
    xyz = epp_b200.gen_states_dev(worlds.SEED_TRAJ, 0, n_traj * rows, lo, hi).view(n_traj, rows, 3)
    xyz[::3, :, 2] = 1.9                       # a third of the trajectories fly above everything
    mats = torch.zeros((n_traj, 10, rows + 1), dtype=torch.float64, device="cuda")   # ld = rows + 1 (even)
    for c in range(3):
        mats[:, 3 * c, :rows] = xyz[:, :, c]
    got = {}
    try:
        for variant in (0, 1):
            epp_b200.set_tuning("traj_variant", variant)
            v, sm, f = gw.check_trajectories_dev(mats, n_traj, rows, rows + 1, 10 * (rows + 1), 10, 0.1)
            torch.cuda.synchronize()
            got[variant] = (v.cpu().numpy().view(np.uint32), sm.cpu().numpy().view(np.uint32), f.cpu().numpy())
    finally:
        epp_b200.set_tuning("traj_variant", 0)
    for a, b in zip(got[0], got[1]):
        assert np.array_equal(a, b)
    host = xyz.cpu().numpy()
    for t in (0, 1, 2, 777, 1999):
        m = ow.check_states_mindist(host[t], 0.1)
        assert np.array_equal(got[0][1][t], m)
        bad = np.flatnonzero(~np.unpackbits(m.view(np.uint8), bitorder="little")[:rows].astype(bool))
        assert got[0][2][t] == (bad[0] if bad.size else -1)
    nvalid = int(np.unpackbits(got[0][0].view(np.uint8), bitorder="little")[:n_traj].sum())
    assert 0 < nvalid < n_traj


@pytest.mark.parametrize("bp", BPS, ids=BP_IDS)
def test_fused_polynomial_sweep(oracle, config, bp):
    """Config 4 (fused form) at reduced size: coefficients -> sampler -> Horner -> world check."""
    g, o, lo, hi = worlds.world_256()
    ow, gw = build_pair(oracle, epp_b200, config, g, o, bp)
    n_traj = 150
    coeffs, T, dt = worlds.rest_to_rest_trajectories(n_traj, 8, lo, hi, n_samples=1000, max_hop=3.0)
    valid, smask, ns, first = gw.check_poly_trajectories(coeffs, T, dt, 0.1, max_samples=1024)
    n_valid = 0
    for t in range(n_traj):
        ok, m, n, f = ow.check_poly_traj(coeffs[t], T[t], dt[t], 0.1, max_samples=1024)
        assert ns[t] == n, t
        assert 990 <= n <= 1010
        assert np.array_equal(smask[t], m), t
        assert first[t] == f
        assert bool(valid[t >> 5] >> (t & 31) & 1) is ok
        n_valid += ok
    assert 0 < n_valid < n_traj
    # fused == pre-sampled on the oracle's own samples
    xyz, times, n = oracle.traj_evaluate_range(coeffs[0], T[0], 0.0, oracle.traj_max_time(T[0]), dt[0])
    mat = worlds.traj_matrix_from_positions(xyz, times)
    v2, s2, _ = gw.check_trajectories(mat[None], 0.1)
    assert np.array_equal(s2[0], smask[0][: s2.shape[1]])


def test_polynomial_sweep_ragged(oracle, config, track):
    """Different segment counts / sampling steps, dt not dividing the duration, tiny batches."""
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    for n_traj, K, ns in ((1, 1, 40), (33, 3, 100), (65, 5, 257)):
        coeffs, T, dt = worlds.rest_to_rest_trajectories(n_traj, K, lo, hi, seed=77 + K, n_samples=ns)
        dt = dt * (1.0 + 0.37 * np.arange(n_traj) / max(n_traj, 1))
        valid, smask, cnt, first = gw.check_poly_trajectories(coeffs, T, dt, 0.15, max_samples=512)
        for t in range(n_traj):
            ok, m, n, f = ow.check_poly_traj(coeffs[t], T[t], dt[t], 0.15, max_samples=512)
            assert cnt[t] == n and np.array_equal(smask[t], m) and first[t] == f
            assert bool(valid[t >> 5] >> (t & 31) & 1) is ok


@pytest.mark.parametrize("G", [0, 1, 4, 32])
def test_polynomial_sweep_truncation_and_block_shapes(oracle, config, track, G):
    """A trajectory with more rows than max_samples must not pass as valid on its unchecked tail: it is reported
    invalid with first row = max_samples unless an earlier row already is; n_samples = rows emitted; the rows that were
    checked equal the oracle's.  Default max_samples (derived from the durations) checks every row.  All block shapes
    of the producer / consumer kernel (G trajectories per block) give the same answers."""
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    n_traj = 70
    coeffs, T, dt = worlds.rest_to_rest_trajectories(n_traj, 4, lo, hi, seed=4141, n_samples=300)
    dt = dt * (1.0 + 0.9 * (np.arange(n_traj) % 7 == 0))  # every 7th trajectory has ~160 rows, the others ~300
    epp_b200.set_tuning("poly_traj_per_warp", G)
    try:
        valid, smask, cnt, first = gw.check_poly_trajectories(coeffs, T, dt, 0.15, max_samples=192)
        full = gw.check_poly_trajectories(coeffs, T, dt, 0.15)
    finally:
        epp_b200.set_tuning("poly_traj_per_warp", 0)
    truncated = 0
    for t in range(n_traj):
        ok, m, n, f = ow.check_poly_traj(coeffs[t], T[t], dt[t], 0.15, max_samples=1024)
        assert full[2][t] == n and np.array_equal(full[1][t], m[: full[1].shape[1]]) and full[3][t] == f
        assert bool(full[0][t >> 5] >> (t & 31) & 1) is ok
        got_ok = bool(valid[t >> 5] >> (t & 31) & 1)
        if n <= 192:
            assert cnt[t] == n and np.array_equal(smask[t], m[:6]) and first[t] == f and got_ok is ok
        else:
            truncated += 1
            assert cnt[t] == 192 and np.array_equal(smask[t], m[:6]) and not got_ok
            assert first[t] == (f if 0 <= f < 192 else 192)
    assert 0 < truncated < n_traj


@pytest.mark.parametrize("G", [0, 4, 14])
def test_polynomial_sweep_segment_shapes(oracle, config, track, G):
    """K3b over ordinary segments (~125 rows each), segments of a dozen rows, segments shorter than one sampling step
    (several segment changes between two rows), with rings of 4 and 2 rounds, consumers that sleep on the round's
    barrier (knob poly_wait) or poll, and both register shapes of the kernel (knob poly_blocks_per_sm)."""
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    cases = []
    cases.append(worlds.rest_to_rest_trajectories(70, 8, lo, hi, seed=901, n_samples=1000))
    cases.append(worlds.rest_to_rest_trajectories(40, 24, lo, hi, seed=902, n_samples=300))
    coeffs, T, dt = worlds.rest_to_rest_trajectories(50, 12, lo, hi, seed=903, n_samples=600)
    rng = np.random.default_rng(903)
    T = T * np.where(rng.random(T.shape) < 0.4, 0.004, 1.0)  # 40 % of the segments last less than a sampling step
    cases.append((coeffs, np.ascontiguousarray(T), T.sum(axis=1) / 599.0))
    for ring in (4, 2):
        epp_b200.set_tuning("poly_wait", 1 if ring == 2 else 0)
        epp_b200.set_tuning("poly_blocks_per_sm", 4 if ring == 2 else 5)  # 256 x 4 (64 registers) / the other shapes
        epp_b200.set_tuning("poly_threads", {0: 0, 4: 256, 14: 192}[G] if ring == 4 else 0)  # 224 x 5, 256 x 5, 192 x 6
        epp_b200.set_tuning("poly_traj_per_warp", G)
        epp_b200.set_tuning("poly_ring", ring)
        try:
            got = [gw.check_poly_trajectories(c, T, dt, 0.15, max_samples=1024) for c, T, dt in cases]
        finally:
            epp_b200.set_tuning("poly_wait", 0)
            epp_b200.set_tuning("poly_blocks_per_sm", 0)
            epp_b200.set_tuning("poly_threads", 0)
            epp_b200.set_tuning("poly_traj_per_warp", 0)
            epp_b200.set_tuning("poly_ring", 0)
        for (c, T, dt), (valid, smask, cnt, first) in zip(cases, got):
            rows = rows_valid = 0
            for t in range(c.shape[0]):
                ok, m, n, f = ow.check_poly_traj(c[t], T[t], dt[t], 0.15, max_samples=1024)
                assert cnt[t] == n and np.array_equal(smask[t], m) and first[t] == f, (ring, t)
                assert bool(valid[t >> 5] >> (t & 31) & 1) is ok
                rows += n
                rows_valid += int(np.unpackbits(np.asarray(m, np.uint32).view(np.uint8)).sum())
            assert 0 < rows_valid < rows  # both verdicts occur among the rows


def test_polynomial_sweep_launch_shapes_agree(oracle, config):
    """A batch large enough for the launcher's batch-size rules (one wave of 256 x 4 blocks of 64-register threads by
    default; 224 x 5; 256 x 5 at 48 registers; 192 x 6; 512 x 2; the halving rule; several waves of 16): every shape
    gives the same words, and the oracle's on a sample."""
    g, o, lo, hi = worlds.world_256()
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    n_traj = 7400
    coeffs, T, dt = worlds.rest_to_rest_trajectories(n_traj, 4, lo, hi, seed=911, n_samples=96, max_hop=3.0)
    shapes = [{}, {"poly_blocks_per_sm": 5}, {"poly_threads": 256}, {"poly_threads": 192}, {"poly_threads": 512},
              {"poly_balance": 0}, {"poly_traj_per_warp": 16, "poly_ring": 3}]
    results = []
    for knobs in shapes:
        for k, v in knobs.items():
            epp_b200.set_tuning(k, v)
        try:
            results.append(gw.check_poly_trajectories(coeffs, T, dt, 0.1, max_samples=128))
        finally:
            for k in knobs:
                epp_b200.set_tuning(k, {"poly_balance": 1}.get(k, 0))
    for r in results[1:]:
        for got, want in zip(r, results[0]):
            assert np.array_equal(got, want)
    valid, smask, ns, first = results[0]
    n_valid = 0
    for t in range(0, n_traj, 37):
        ok, m, n, f = ow.check_poly_traj(coeffs[t], T[t], dt[t], 0.1, max_samples=128)
        assert ns[t] == n and np.array_equal(smask[t], m) and first[t] == f
        assert bool(valid[t >> 5] >> (t & 31) & 1) is ok
        n_valid += ok
    assert 0 < n_valid < len(range(0, n_traj, 37))


def test_device_buffers_and_generators(oracle, config, track):
    """_dev entry points on torch buffers + on-device generators bit-identical to the host twins."""
    import torch
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    n = 100003
    d_xyz = epp_b200.gen_states_dev(worlds.SEED_STATES, 12345, n, lo, hi)
    h_xyz = oracle.gen_states(worlds.SEED_STATES, 12345, n, lo, hi)
    assert np.array_equal(d_xyz.cpu().numpy(), h_xyz)
    mask = gw.check_states_dev(d_xyz)
    torch.cuda.synchronize()
    want = ow.check_states(h_xyz, threads=8)
    assert np.array_equal(mask.cpu().numpy().view(np.uint32), want)
    d_a, d_b = epp_b200.gen_segments_dev(worlds.SEED_SEGMENTS, 777, n, lo, hi, 1.0)
    h_a, h_b = oracle.gen_segments(worlds.SEED_SEGMENTS, 777, n, lo, hi, 1.0)
    assert np.array_equal(d_a.cpu().numpy(), h_a) and np.array_equal(d_b.cpu().numpy(), h_b)
    smask = gw.check_segments_dev(d_a, d_b)
    wantseg = ow.check_segments(h_a, h_b, threads=8)
    assert np.array_equal(smask.cpu().numpy().view(np.uint32), wantseg)
    md = gw.check_states_dev(d_xyz, min_distance=0.1)
    assert np.array_equal(md.cpu().numpy().view(np.uint32), ow.check_states_mindist(h_xyz, 0.1, threads=8))
    # mask summary: device reduction == host restatement, shards combine
    assert epp_b200.mask_summary_dev(mask) == epp_b200.mask_summary_host(want)
    half = (mask.numel() // 2)
    p0, x0 = epp_b200.mask_summary_dev(mask[:half], 0)
    p1, x1 = epp_b200.mask_summary_dev(mask[half:], half)
    assert (p0 + p1, x0 ^ x1) == epp_b200.mask_summary_host(want)
    assert epp_b200.kernel_launch_count() > 0


def test_pinned_host_buffers(oracle, config, track):
    import torch
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    n = 1 << 16
    h = torch.from_numpy(worlds.gen_states(31, 0, n, lo, hi)).pin_memory()
    m = torch.zeros((n + 31) // 32, dtype=torch.int32).pin_memory()
    epp_b200._chk(epp_b200.lib().epp_check_states(gw.handle, h.data_ptr(), n, 0, m.data_ptr()))
    assert np.array_equal(m.numpy().view(np.uint32), ow.check_states(h.numpy()))


def test_config5_cluttered_subsample(oracle, config):
    """Config 5 at a size the oracle finishes in seconds: 10k-OBB world, grid broad phase, inputs
    generated on device, parity on the same indices + checksum equality across shard counts."""
    import torch
    g, o, lo, hi = worlds.world_10k()
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    assert gw.size() == 10000 and gw.active_broadphase() == epp_b200.BP_GRID
    n = 1 << 20
    first = 123456789
    d_xyz = epp_b200.gen_states_dev(worlds.SEED_CLUTTER, first, n, lo, hi)
    mask = gw.check_states_dev(d_xyz)
    want = ow.check_states(oracle.gen_states(worlds.SEED_CLUTTER, first, n, lo, hi), threads=8)
    assert mismatches(mask.cpu().numpy().view(np.uint32), want) == 0
    whole = epp_b200.mask_summary_dev(mask)
    for shards in (2, 4, 8):
        pop, x = 0, 0
        for r in range(shards):
            b, e = epp_b200.shard_range(n, r, shards, 1024)
            part = gw.check_states_dev(d_xyz[b:e])
            p, c = epp_b200.mask_summary_dev(part, b // 32)
            pop, x = pop + p, x ^ c
        assert (pop, x) == whole
    ne = 1 << 17
    d_a, d_b = epp_b200.gen_segments_dev(worlds.SEED_CLUTTER, first, ne, lo, hi, 1.0)
    got = gw.check_segments_dev(d_a, d_b).cpu().numpy().view(np.uint32)
    h_a, h_b = oracle.gen_segments(worlds.SEED_CLUTTER, first, ne, lo, hi, 1.0)
    assert mismatches(got, ow.check_segments(h_a, h_b, threads=8)) == 0
    # grid vs bits vs brute on the same world (bits allowed up to 1024 OBBs only -> brute here)
    gb = epp_b200.World(config, broadphase=epp_b200.BP_BRUTE)
    gb.parse_gates_and_obstacles(g, o)
    gb.set_broadphase(epp_b200.BP_BRUTE)
    sub = d_xyz[: 1 << 16]
    assert torch.equal(gb.check_states_dev(sub), gw.check_states_dev(sub))


def test_concurrent_host_threads(oracle, config, track):
    """Two planner threads on one World (src/OnlineTrajGenerator.cpp:324-340)."""
    import threading
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    gw.commit()
    out = {}

    def work(tid):
        xyz = worlds.gen_states(1000 + tid, 0, 30000, lo, hi)
        res = [gw.check_states(xyz) for _ in range(5)]
        singles = [gw.checkPointValidity(p, False) for p in xyz[:50]]
        out[tid] = (xyz, res, singles)

    ts = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    for tid, (xyz, res, singles) in out.items():
        want = ow.check_states(xyz)
        assert all(np.array_equal(r, want) for r in res)
        assert singles == [bool(want[i >> 5] >> (i & 31) & 1) for i in range(50)]


@pytest.mark.parametrize("bp", BPS, ids=BP_IDS)
def test_states_warp_form_equals_direct_form(oracle, config, bp):
    """K1 has two forms: warp-pipelined (producer warp + cp.async.bulk ring + per-warp compaction of the
    states that have candidates) and direct (per-lane loads and candidate walk, small batches).  Both must
    give the oracle's mask for whole chunks, ragged tails, batches smaller than one chunk, rings that wrap
    many times, every block shape, and 8 B-aligned (not 16 B) batches (loaded by the lanes instead of TMA)."""
    import torch
    g, o, lo, hi = worlds.world_256()
    ow, gw = build_pair(oracle, epp_b200, config, g, o, bp)
    keys = ("states_variant", "states_per_thread", "states_warp_blocks_per_sm", "states_rec_smem",
            "states_warp_log2_slots", "states_warp_threads", "states_pdl")
    saved = {k: epp_b200.get_tuning(k) for k in keys}
    settings = [(1, 0, 1, 1, 0, 1024, 1), (0, 0, 1, 1, 0, 1024, 1),
                # 4 / 8 words per chunk, ring of 2 slots (wraps many times) .. as many as fit, 1 .. 31 consumer
                # warps, 1 .. 4 blocks per SM, with and without programmatic dependent launch
                (3, 4, 1, 1, 0, 1024, 1), (3, 8, 1, 1, 1, 64, 0), (3, 8, 2, 0, 2, 512, 1), (3, 4, 4, 1, 3, 256, 0)]
    try:
        for n in (1, 127, 128, 129, 1024, 1025, 8192 + 511, 100003, 1 << 19):
            buf = torch.empty(3 * n + 1, dtype=torch.float64, device="cuda")
            d_xyz = epp_b200.gen_states_dev(worlds.SEED_STATES, 7 * n, n, lo, hi, out=buf[: 3 * n].view(n, 3))
            h_xyz = oracle.gen_states(worlds.SEED_STATES, 7 * n, n, lo, hi)
            want = ow.check_states(h_xyz, threads=8)
            want_md = ow.check_states_mindist(h_xyz, 0.1, threads=8)
            for setting in settings:
                for k, v in zip(keys, setting):
                    epp_b200.set_tuning(k, v)
                got = gw.check_states_dev(d_xyz).cpu().numpy().view(np.uint32)
                assert np.array_equal(got, want), (n, setting)
                md = gw.check_states_dev(d_xyz, min_distance=0.1).cpu().numpy().view(np.uint32)
                assert np.array_equal(md, want_md), (n, setting)
            # same states at an address that is only 8 B aligned
            shifted = buf[1:].view(n, 3)
            shifted.copy_(d_xyz.clone())
            for variant in (1, 3):
                epp_b200.set_tuning("states_variant", variant)
                assert np.array_equal(gw.check_states_dev(shifted).cpu().numpy().view(np.uint32), want), (n, variant)
    finally:
        for k, v in saved.items():
            epp_b200.set_tuning(k, v)


@pytest.mark.parametrize("bp", BPS, ids=BP_IDS)
def test_states_warp_form_small_world_records_in_smem(oracle, config, track, bp):
    """The warp-pipelined K1 folds canPassGate / minDistance / filling rules into per-OBB records in
    shared memory for worlds of <= 64 OBBs.  Track + filling boxes (M=28), every call flavour, sizes
    around the word / step boundaries, odd sizes (lane-loaded last chunk) and an 8 B-aligned batch."""
    import torch
    g, o, lo, hi = track
    cfg = filling_config(config)
    ow, gw = build_pair(oracle, epp_b200, cfg, g, o, bp)
    keys = ("states_variant", "states_warp_log2_slots", "states_per_thread", "states_warp_threads",
            "states_warp_blocks_per_sm", "states_rec_smem")
    saved = {k: epp_b200.get_tuning(k) for k in keys}
    adv = adversarial_points(ow.obbs(), np.random.default_rng(8), per_face=2)
    try:
        for n in (1, 2, 31, 33, 127, 128, 129, 4095, 20001, 1 << 17):
            h_xyz = np.array(oracle.gen_states(21, 3 * n, n, lo, hi))
            if n > len(adv):
                h_xyz[-len(adv):] = adv  # face-adjacent points at the end (covers the ragged last chunk)
            buf = torch.empty(3 * n + 1, dtype=torch.float64, device="cuda")
            d_xyz = buf[: 3 * n].view(n, 3)
            d_xyz.copy_(torch.from_numpy(h_xyz))
            shifted = buf[1:].view(n, 3)
            want = {("b", c): ow.check_states(h_xyz, c, threads=8) for c in (False, True)}
            want.update({("m", md): ow.check_states_mindist(h_xyz, md, threads=8) for md in (0.05, 0.3)})
            for setting in [(3, 0, 4, 1024, 1, 1), (3, 1, 8, 64, 1, 1), (3, 2, 8, 256, 4, 0)]:
                for k, v in zip(keys, setting):
                    epp_b200.set_tuning(k, v)
                for src in (d_xyz, None):
                    if src is None:
                        shifted.copy_(d_xyz.clone())
                        src = shifted
                    for (kind, arg), w in want.items():
                        got = gw.check_states_dev(src, can_pass_gate=arg) if kind == "b" else \
                            gw.check_states_dev(src, min_distance=arg)
                        assert np.array_equal(got.cpu().numpy().view(np.uint32), w), (n, setting, kind, arg)
                    d_xyz.copy_(torch.from_numpy(h_xyz))
    finally:
        for k, v in saved.items():
            epp_b200.set_tuning(k, v)


def test_states_warp_form_ring_wrap_stress(oracle, config, track):
    """Many launches of batches large enough for every block's ring to wrap dozens of times, with ring sizes
    from 2 slots to as many as fit: the mask checksum must never change and must equal the direct form's
    (chunk hand-out is dynamic, so a protocol hole shows up as a rare wrong word or a hang, not every run)."""
    import torch
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    keys = ("states_variant", "states_warp_log2_slots", "states_warp_threads", "states_per_thread")
    saved = {k: epp_b200.get_tuning(k) for k in keys}
    n = 1 << 23
    d_xyz = epp_b200.gen_states_dev(worlds.SEED_STATES, 123, n, lo, hi)
    mask = torch.empty(n // 32, dtype=torch.int32, device="cuda")
    try:
        epp_b200.set_tuning("states_variant", 1)
        gw.check_states_dev(d_xyz, mask=mask)
        want = epp_b200.mask_summary_dev(mask)
        sub = 1 << 18
        assert np.array_equal(mask[: sub // 32].cpu().numpy().view(np.uint32),
                              ow.check_states(d_xyz[:sub].cpu().numpy(), threads=8))
        for setting in [(3, 0, 1024, 4), (3, 1, 1024, 4), (3, 3, 1024, 8), (3, 5, 512, 4), (3, 0, 256, 8)]:
            for k, v in zip(keys, setting):
                epp_b200.set_tuning(k, v)
            for rep in range(60):
                mask.zero_()
                gw.check_states_dev(d_xyz, mask=mask)
                assert epp_b200.mask_summary_dev(mask) == want, (setting, rep)
    finally:
        for k, v in saved.items():
            epp_b200.set_tuning(k, v)


# ---- mask exchange fused into the check kernels (epp_gather) -------------------------------------------

def _run_logical_ranks(gathers, fn):
    """Issue fn(rank, gather) for every logical rank on its own stream, then synchronise.  The ranks'
    kernels wait for each other on the device, so on ONE GPU they must fit next to each other: the
    sm_limit knob sizes every grid for 1/world_size of the SMs."""
    import torch
    streams = [torch.cuda.Stream() for _ in gathers]
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    epp_b200.set_tuning("sm_limit", max(1, sms // len(gathers) - 1))
    try:
        torch.cuda.synchronize()
        for r, (g, st) in enumerate(zip(gathers, streams)):
            with torch.cuda.stream(st):
                fn(r, g)
        torch.cuda.synchronize()
    finally:
        epp_b200.set_tuning("sm_limit", 0)


@pytest.mark.parametrize("world_size", [2, 3])
@pytest.mark.parametrize("n", [100_000, 3_000, 1 << 20])
def test_gather_states_logical_ranks(oracle, config, track, world_size, n):
    """world_size logical ranks on ONE GPU (in-process peers): the fused exchange of the warp form
    (n >= 4096 per rank) and the push kernel behind the direct form leave the all-gathered mask, equal to
    the oracle's, in every rank's buffer; sequence numbers survive repeated calls and both buffers."""
    import torch
    from epp_b200 import sharding
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    xyz = worlds.gen_states(worlds.SEED_STATES, 777, n, lo, hi)
    want = ow.check_states(xyz, threads=8)
    d_xyz = torch.from_numpy(xyz).cuda()
    sw = sharding.slot_words(n, world_size)
    gathers = [epp_b200.Gather(r, world_size, sw, n_buffers=2) for r in range(world_size)]
    epp_b200.Gather.connect_local(gathers)
    for call in range(4):
        buf = call & 1
        for gth in gathers:
            gth.tensor(buf).fill_(-1 if call < 2 else 0)  # stale contents must be overwritten
        torch.cuda.synchronize()

        def issue(r, gth):
            b, e = epp_b200.shard_range(n, r, world_size, 1024)
            gth.check_states(gw, d_xyz[b:e], buf)

        _run_logical_ranks(gathers, issue)
        for r, gth in enumerate(gathers):
            t = gth.tensor(buf)
            got = sharding.assemble(t, n)
            assert mismatches(got, want) == 0, "rank %d call %d" % (r, call)
            for p in range(world_size):  # words beyond a rank's units arrive as 0
                b, e = epp_b200.shard_range(n, p, world_size, 1024)
                assert int(t[p][(e - b + 31) // 32:].abs().sum().item()) == 0
    for gth in gathers:
        calls, timeouts = gth.status()
        assert calls == 4 and timeouts == 0


def test_gather_segments_and_mindist_logical_ranks(oracle, config):
    import torch
    from epp_b200 import sharding
    g, o, lo, hi = worlds.world_256()
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    world_size, n = 2, 150_001
    a, b = worlds.gen_segments(worlds.SEED_SEGMENTS, 5, n, lo, hi, 1.0)
    want_seg = ow.check_segments(a, b, threads=8)
    want_md = ow.check_states_mindist(a, 0.3, threads=8)
    d_a, d_b = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    sw = sharding.slot_words(n, world_size)
    gathers = [epp_b200.Gather(r, world_size, sw, n_buffers=2) for r in range(world_size)]
    epp_b200.Gather.connect_local(gathers)

    def seg(r, gth):
        lo_, hi_ = epp_b200.shard_range(n, r, world_size, 1024)
        gth.check_segments(gw, d_a[lo_:hi_], d_b[lo_:hi_], 0)

    def md(r, gth):
        lo_, hi_ = epp_b200.shard_range(n, r, world_size, 1024)
        gth.check_states(gw, d_a[lo_:hi_], 1, min_distance=0.3)

    # The two ranks' ranges differ in size and may take different block sizes of K2.  CUDA loads a kernel lazily at
    # its first launch and that load waits for running kernels — here the other logical rank's push kernel, which is
    # waiting for this one (only an issue with several ranks inside ONE process): launch every form once beforehand.
    for variant, block in ((3, 256), (3, 512), (3, 1024), (2, 0)):
        epp_b200.set_tuning("segments_variant", variant)
        epp_b200.set_tuning("segments_block", block)
        gw.check_segments_dev(d_a[:4096], d_b[:4096])
    epp_b200.set_tuning("segments_block", 0)
    epp_b200.set_tuning("segments_variant", 0)
    torch.cuda.synchronize()
    _run_logical_ranks(gathers, seg)
    _run_logical_ranks(gathers, md)
    for gth in gathers:
        assert mismatches(sharding.assemble(gth.tensor(0), n), want_seg) == 0
        assert mismatches(sharding.assemble(gth.tensor(1), n), want_md) == 0
        assert gth.status() == (2, 0)


def test_gather_missing_peer_times_out_instead_of_hanging(oracle, config, track):
    """A peer that never issues its call must not hang the GPU: the wait is bounded and counted."""
    import torch
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    n = 50_000
    d_xyz = epp_b200.gen_states_dev(worlds.SEED_STATES, 0, n, lo, hi)
    gathers = [epp_b200.Gather(r, 2, 2048, n_buffers=1) for r in range(2)]
    epp_b200.Gather.connect_local(gathers)
    gathers[0].set_timeout_ms(50)
    gathers[0].check_states(gw, d_xyz, 0)
    torch.cuda.synchronize()
    calls, timeouts = gathers[0].status()
    assert calls == 1 and timeouts >= 1
    # the local slot is still correct
    want = gw.check_states_dev(d_xyz)
    assert torch.equal(gathers[0].tensor(0)[0][: (n + 31) // 32], want)


def test_gather_argument_errors(config, track):
    g, o, lo, hi = track
    gw = epp_b200.World(config)
    gw.parse_gates_and_obstacles(g, o)
    with pytest.raises(epp_b200.EppError):
        epp_b200.Gather(0, 9, 1024)          # more than 8 ranks
    with pytest.raises(epp_b200.EppError):
        epp_b200.Gather(2, 2, 1024)          # rank out of range
    gth = epp_b200.Gather(0, 2, 1024)
    d_xyz = epp_b200.gen_states_dev(worlds.SEED_STATES, 0, 1024, lo, hi)
    with pytest.raises(epp_b200.EppError):
        gth.check_states(gw, d_xyz, 0)       # not connected
    one = epp_b200.Gather(0, 1, 28)
    with pytest.raises(epp_b200.EppError):
        one.check_states(gw, d_xyz, 0)       # 1024 states need 32 words, the slot holds 28
    with pytest.raises(epp_b200.EppError):
        one.check_states(gw, d_xyz[:64], 2)  # buffer index out of range (n_buffers = 2)
    one.check_states(gw, d_xyz[:512], 0)     # world_size 1: a plain check into the buffer
    import torch
    assert torch.equal(one.tensor(0)[0][:16], gw.check_states_dev(d_xyz[:512]))


@pytest.mark.parametrize("bp", BPS, ids=BP_IDS)
def test_segments_smem_record_form_equals_global_record_form(oracle, config, bp):
    """K2's forms (warp-pipelined with the end points through the cp.async.bulk ring; direct with records folded into
    shared memory; direct with records in global memory) and the oracle agree, incl. canPassGate with filling boxes,
    ragged tails, zero-length segments, batches smaller than one chunk and 8 B-aligned batches (loaded by the lanes)."""
    import torch
    g, o, lo, hi = worlds.world_256()
    cfgf = filling_config(config)
    ow, gw = build_pair(oracle, epp_b200, cfgf, g, o, broadphase=bp)
    rng = np.random.default_rng(99)
    for n in (1, 31, 33, 4097, 70_001, 160_003):
        a, b = worlds.gen_segments(worlds.SEED_SEGMENTS, 1000 + n, n, lo, hi, 1.0)
        b[::17] = a[::17]  # zero-length segments
        extra_a, extra_b = adversarial_segments(gw.obbs(), rng, lo, hi, per_obb=2)
        m = min(len(extra_a), n // 2)
        a[:m], b[:m] = np.asarray(extra_a)[:m], np.asarray(extra_b)[:m]
        d_a, d_b = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
        for cpg in (False, True):
            want = ow.check_segments(a, b, cpg, threads=8)
            got = {}
            # variant 0 auto, 1 direct + global records, 2 warp-pipelined, 3 direct forms only; block = threads per
            # block of the direct shared-memory form (0 = by batch size)
            for variant, block in ((0, 0), (2, 0), (3, 0), (3, 256), (3, 512), (3, 1024), (1, 0)):
                epp_b200.set_tuning("segments_variant", variant)
                epp_b200.set_tuning("segments_block", block)
                try:
                    got[variant, block] = gw.check_segments_dev(d_a, d_b, can_pass_gate=cpg).cpu().numpy().view(np.uint32)
                    if variant == 2 and n > 64:  # 8 B-aligned end points: the consumers load the chunk themselves
                        sub = gw.check_segments_dev(d_a.view(-1)[3:].view(-1, 3), d_b.view(-1)[3:].view(-1, 3),
                                                    can_pass_gate=cpg).cpu().numpy().view(np.uint32)
                        assert mismatches(sub, ow.check_segments(a[1:], b[1:], cpg, threads=8)) == 0, (n, cpg, "unaligned")
                finally:
                    epp_b200.set_tuning("segments_variant", 0)
                    epp_b200.set_tuning("segments_block", 0)
            for key, words in got.items():
                assert mismatches(words, want) == 0, (n, cpg, key)


@pytest.mark.parametrize("bp", BPS, ids=BP_IDS)
def test_states_warp_form_record_fold_small_batches(oracle, config, bp):
    """Worlds of up to 256 OBBs fold their records into shared memory, one consumer thread per record; a batch
    with few chunks per block has fewer consumer threads than records and must take the global-record form."""
    g, o, lo, hi = worlds.world_256()
    ow, gw = build_pair(oracle, epp_b200, config, g, o, bp)
    for n in (4096, 4097, 9000, 20_000, 70_000, 108_544, 300_000):
        xyz = worlds.gen_states(4242, n, n, lo, hi)
        for cpg in (False, True):
            assert mismatches(gw.check_states(xyz, cpg), ow.check_states(xyz, cpg, threads=8)) == 0, (n, cpg)
        assert mismatches(gw.check_states_mindist(xyz, 0.15), ow.check_states_mindist(xyz, 0.15, threads=8)) == 0, n


def test_bits_rows_of_32_words_and_rejection_above(oracle, config):
    """The bits broad phase summarises a row in ONE 32-bit word, so it holds worlds of up to 1024 OBBs (32-word rows).
    At 1000 OBBs (summary bit 31 in use) points and segments must equal brute force and the oracle; above 1024 OBBs a
    request for bits is refused at commit (it used to alias word w >= 32 onto w mod 32 and miss OBBs >= 1024)."""
    import torch
    g, o, lo, hi = worlds.lattice_world(160, 200, 60.0, 0x5EED0B17)
    ow, gw = build_pair(oracle, epp_b200, config, g, o, epp_b200.BP_BITS)
    assert gw.size() == 1000
    n = 1 << 16
    xyz = np.concatenate([worlds.gen_states(21, 0, n, lo, hi), adversarial_points(ow.obbs()[900:], np.random.default_rng(6), 2)])
    assert mismatches(gw.check_states(xyz, False), ow.check_states(xyz, False, threads=8)) == 0
    assert gw.active_broadphase() == epp_b200.BP_BITS
    assert mismatches(gw.check_states_mindist(xyz, 0.1), ow.check_states_mindist(xyz, 0.1, threads=8)) == 0
    a, b = worlds.gen_segments(22, 0, n, lo, hi, 1.0)
    aa, ab = adversarial_segments(ow.obbs()[960:], np.random.default_rng(7), lo, hi, per_obb=4)
    a, b = np.concatenate([a, aa]), np.concatenate([b, ab])
    assert mismatches(gw.check_segments(a, b, False), ow.check_segments(a, b, False, threads=8)) == 0
    d = torch.from_numpy(xyz[:n]).cuda()
    gb = epp_b200.World(config, broadphase=epp_b200.BP_BRUTE)
    gb.parse_gates_and_obstacles(g, o)
    gb.set_broadphase(epp_b200.BP_BRUTE)
    assert torch.equal(gb.check_states_dev(d), gw.check_states_dev(d))
    # 2000 OBBs: bits refused, auto falls to the grid even when the knob asks for more
    g2, o2, lo2, hi2 = worlds.lattice_world(320, 400, 80.0, 0x5EED0B18)
    big = epp_b200.World(config, broadphase=epp_b200.BP_BITS)
    with pytest.raises(epp_b200.EppError, match="1024"):
        big.parse_gates_and_obstacles(g2, o2)
    epp_b200.set_tuning("bits_max_obbs", 4096)
    try:
        auto = epp_b200.World(config)
        auto.parse_gates_and_obstacles(g2, o2)
        assert auto.size() == 2000 and auto.active_broadphase() == epp_b200.BP_GRID
        ow2 = oracle.OracleWorld(config)
        ow2.parse_gates_and_obstacles(g2, o2)
        xyz2 = worlds.gen_states(23, 0, 1 << 15, lo2, hi2)
        assert mismatches(auto.check_states(xyz2, False), ow2.check_states(xyz2, False, threads=8)) == 0
    finally:
        epp_b200.set_tuning("bits_max_obbs", 1024)


def test_commit_new_epochs_while_checks_are_in_flight(oracle, config, track):
    """SURVEY.md §5 race row: the reference rebuilds its World (src/OnlineTrajGenerator.cpp:146) while planner threads
    may still be querying it (:324-340).  Here a mutator thread moves a gate and commits new epochs as fast as it can
    while three threads keep checking batches and single units; every answer must be the oracle's answer for ONE of
    the two worlds (never a mixture, never a crash), and both worlds must be observed."""
    import threading
    g, o, lo, hi = track
    g2 = np.array(g, copy=True)
    g2[1, 0] += 0.15
    g2[1, 5] += 0.3
    gw = epp_b200.World(config)
    gw.parse_gates_and_obstacles(g, o)
    want = []
    xyz = np.concatenate([worlds.gen_states(777, 0, 60000, lo, hi),
                          np.asarray(g2[1, :3]) + (np.random.default_rng(1).random((4000, 3)) - 0.5) * [1.6, 1.6, 2.0] + [0, 0, 1.0]])
    a, b = worlds.gen_segments(778, 0, 30000, lo, hi, 1.0)
    for gg in (g, g2):
        ow = oracle.OracleWorld(config)
        ow.parse_gates_and_obstacles(gg, o)
        want.append((ow.check_states(xyz, False, threads=8), ow.check_segments(a, b, False, threads=8)))
    assert not np.array_equal(want[0][0], want[1][0])
    stop = threading.Event()
    seen, errors = [set(), set(), set()], []

    def checker(tid):
        try:
            while not stop.is_set():
                ms = gw.check_states(xyz, False)
                me = gw.check_segments(a, b, False)
                for got, idx in ((ms, 0), (me, 1)):
                    which = [k for k in (0, 1) if np.array_equal(got, want[k][idx])]
                    if not which:
                        errors.append("thread %d: a mask that belongs to neither world" % tid)
                        return
                    seen[tid].update(which)
                for i in range(0, 200, 7):  # single units through the mailbox warp, across epoch changes
                    one = gw.checkPointValidity(xyz[60000 + i], False)
                    bits = [bool(want[k][0][(60000 + i) >> 5] >> ((60000 + i) & 31) & 1) for k in (0, 1)]
                    if one not in bits:
                        errors.append("thread %d: single answer of neither world" % tid)
                        return
        except Exception as ex:  # noqa: BLE001
            errors.append(repr(ex))

    def mutator():
        try:
            k = 0
            while not stop.is_set():
                k ^= 1
                gw.updateGatePosition(1, (g2 if k else g)[1])
                gw.commit()
        except Exception as ex:  # noqa: BLE001
            errors.append("mutator: " + repr(ex))

    ts = [threading.Thread(target=checker, args=(i,)) for i in range(3)] + [threading.Thread(target=mutator)]
    [t.start() for t in ts]
    import time
    time.sleep(3.0)
    stop.set()
    [t.join(60) for t in ts]
    assert not errors, errors[:3]
    assert set().union(*seen) == {0, 1}


def test_single_unit_mailbox_life_cycle(oracle, config, track):
    """epp_is_valid / epp_check_motion go through a resident warp polling a mapped request line.  It must answer like
    the batch kernels, leave when idle (so device-wide synchronisations are not held up), come back on the next
    call, follow the world across commits, and the launch + synchronise form behind the knob must agree."""
    import time
    import torch
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    xyz = np.concatenate([worlds.gen_states(901, 0, 300, lo, hi), adversarial_points(ow.obbs()[:4], np.random.default_rng(8), 1)])
    a, b = worlds.gen_segments(902, 0, 300, lo, hi, 1.0)
    for cpg in (False, True):
        wp = np.unpackbits(ow.check_states(xyz, cpg).view(np.uint8), bitorder="little")
        ws = np.unpackbits(ow.check_segments(a, b, cpg).view(np.uint8), bitorder="little")
        assert [gw.checkPointValidity(p, cpg) for p in xyz] == [bool(v) for v in wp[: len(xyz)]]
        assert [gw.checkRayValid(p, q, cpg) for p, q in zip(a, b)] == [bool(v) for v in ws[:300]]
    # idle exit: after 20 ms without a request the warp has left and a device-wide synchronise returns at once
    time.sleep(0.02)
    t0 = time.perf_counter()
    torch.cuda.synchronize()
    assert time.perf_counter() - t0 < 0.01
    assert gw.checkPointValidity([0.0, 0.0, 0.4], False) == ow.check_point([0.0, 0.0, 0.4], False)
    # the world changes between single calls: the next call must see the new epoch
    p = np.asarray(g[0][:3]) + [0.0, 0.0, 0.3]
    before = gw.checkPointValidity(p, False)
    gw.updateGatePosition(0, [g[0][0] + 1.0, g[0][1] + 1.0, 0, 0, 0, g[0][5], g[0][6]])
    assert gw.checkPointValidity(p, False) is True and before is False
    # knob off: one launch + synchronise per call, same answers
    epp_b200.set_tuning("single_mailbox", 0)
    try:
        assert gw.checkPointValidity(p, False) is True
        assert [gw.checkRayValid(p_, q_, False) for p_, q_ in zip(a[:50], b[:50])] == \
               [gw.checkRayValid(p_, q_, False) for p_, q_ in zip(a[:50], b[:50])]
    finally:
        epp_b200.set_tuning("single_mailbox", 1)
    us = gw.single_latency_us(2000)
    print("single isValid through the mailbox: %.2f us per call" % us)
    assert us < 20.0
    # big world (grid broad phase): lane 0 walks the cell list
    g5, o5, lo5, hi5 = worlds.world_10k()
    ow5, gw5 = build_pair(oracle, epp_b200, config, g5, o5)
    pts = np.concatenate([worlds.gen_states(903, 0, 100, lo5, hi5), adversarial_points(ow5.obbs()[:3], np.random.default_rng(9), 1)])
    w5 = np.unpackbits(ow5.check_states(pts, False).view(np.uint8), bitorder="little")
    assert [gw5.checkPointValidity(q, False) for q in pts] == [bool(v) for v in w5[: len(pts)]]
    a5, b5 = worlds.gen_segments(904, 0, 100, lo5, hi5, 1.0)
    s5 = np.unpackbits(ow5.check_segments(a5, b5, False).view(np.uint8), bitorder="little")
    assert [gw5.checkRayValid(p_, q_, False) for p_, q_ in zip(a5, b5)] == [bool(v) for v in s5[:100]]


@pytest.mark.parametrize("which", ["track", "w10k"])
def test_multi_batch_call_equals_single_calls(oracle, config, which):
    """epp_check_states_multi_dev: independent batches in one call (kernels overlap at their boundaries: from the second
    batch on the dependency wait sits at the kernel's end).  Every batch's mask must equal the single-call mask and the
    oracle's, for warp-form and direct-form sizes and ragged tails, run repeatedly; overlapping masks are refused."""
    import torch
    g, o, lo, hi = worlds.track_world(config) if which == "track" else worlds.world_10k()
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    sizes = [1 << 20, 300_001, 4096, 100, 1 << 19, 65_537, 1 << 20]
    batches = [epp_b200.gen_states_dev(4000 + k, 0, n, lo, hi) for k, n in enumerate(sizes)]
    masks = [torch.zeros((n + 31) // 32, dtype=torch.int32, device="cuda") for n in sizes]
    singles = [gw.check_states_dev(b) for b in batches]
    for _ in range(5):
        for m in masks:
            m.zero_()
        gw.check_states_multi_dev(batches, masks)
        torch.cuda.synchronize()
        for k in range(len(sizes)):
            assert torch.equal(masks[k], singles[k]), (which, k)
    for k in (0, 1, 3):
        want = ow.check_states(batches[k].cpu().numpy(), False, threads=8)
        assert mismatches(masks[k].cpu().numpy().view(np.uint32), want) == 0
    with pytest.raises(epp_b200.EppError, match="overlap"):
        gw.check_states_multi_dev([batches[2], batches[3]], [masks[2], masks[2]])


@pytest.mark.parametrize("which", ["track", "w256", "w10k"])
def test_multi_batch_segment_call_equals_single_calls(oracle, config, which):
    """epp_check_segments_multi_dev: independent segment batches in one call (half-SM blocks, dependency wait at the
    kernel's end from the second batch on).  Every batch's mask must equal the single-call mask and the oracle's — brute,
    bits and grid worlds, warp-form and direct-form sizes, ragged tails, both canPassGate values, run repeatedly;
    overlapping masks are refused."""
    import torch
    g, o, lo, hi = {"track": lambda: worlds.track_world(config), "w256": worlds.world_256, "w10k": worlds.world_10k}[which]()
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    sizes = [1 << 20, 300_001, 4096, 100, 0, 1 << 19, 65_537, 1 << 20]
    batches = [epp_b200.gen_segments_dev(5000 + k, 0, max(n, 1), lo, hi, 1.0) for k, n in enumerate(sizes)]
    batches = [(a[:n].contiguous(), b[:n].contiguous()) for (a, b), n in zip(batches, sizes)]  # one of them is empty
    masks = [torch.zeros(max((n + 31) // 32, 1), dtype=torch.int32, device="cuda") for n in sizes]
    for cpg in (False, True):
        singles = [gw.check_segments_dev(a, b, can_pass_gate=cpg) if n else None for (a, b), n in zip(batches, sizes)]
        for _ in range(4):
            for m in masks:
                m.zero_()
            gw.check_segments_multi_dev(batches, masks, can_pass_gate=cpg)
            torch.cuda.synchronize()
            for k, n in enumerate(sizes):
                if n:
                    assert torch.equal(masks[k][: (n + 31) // 32], singles[k]), (which, cpg, k)
        for k in (1, 3, 6):
            a, b = batches[k]
            want = ow.check_segments(a.cpu().numpy().reshape(-1, 3), b.cpu().numpy().reshape(-1, 3), cpg, threads=8)
            assert mismatches(masks[k].cpu().numpy().view(np.uint32)[: want.size], want) == 0
    with pytest.raises(epp_b200.EppError, match="overlap"):
        gw.check_segments_multi_dev([batches[2], batches[3]], [masks[2], masks[2]])


def test_gather_bulk_form_logical_ranks(oracle, config, track):
    """Large slots take the bulk form of the exchange (the slot is copied straight into the peers' gathered buffers,
    fence, header): here forced for a small batch through the knob, 3 logical ranks on one GPU, states and segments,
    repeated calls over two alternating buffers; reusing the previous call's buffer is refused."""
    import torch
    from epp_b200 import sharding
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    world_size, n = 3, 200_003
    xyz = worlds.gen_states(worlds.SEED_STATES, 31, n, lo, hi)
    a, b = worlds.gen_segments(worlds.SEED_SEGMENTS, 32, n, lo, hi, 1.0)
    want_s, want_e = ow.check_states(xyz, threads=8), ow.check_segments(a, b, threads=8)
    d_xyz, d_a, d_b = (torch.from_numpy(x).cuda() for x in (xyz, a, b))
    sw = sharding.slot_words(n, world_size)
    epp_b200.set_tuning("gather_bulk_min_words", 1024)
    try:
        gathers = [epp_b200.Gather(r, world_size, sw, n_buffers=2) for r in range(world_size)]
        epp_b200.Gather.connect_local(gathers)
        for variant in ((3, 256), (2, 0)):   # launch every kernel form once before ranks wait for each other
            epp_b200.set_tuning("segments_variant", variant[0])
            epp_b200.set_tuning("segments_block", variant[1])
            gw.check_segments_dev(d_a[:40000], d_b[:40000])
        epp_b200.set_tuning("segments_variant", 0)
        epp_b200.set_tuning("segments_block", 0)
        gw.check_states_dev(d_xyz[:70000])
        torch.cuda.synchronize()
        for call in range(4):
            buf = call & 1
            for gth in gathers:
                gth.tensor(buf).fill_(-1)
            torch.cuda.synchronize()
            seg = call >= 2

            def issue(r, gth):
                lo_, hi_ = epp_b200.shard_range(n, r, world_size, 1024)
                if seg:
                    gth.check_segments(gw, d_a[lo_:hi_], d_b[lo_:hi_], buf)
                else:
                    gth.check_states(gw, d_xyz[lo_:hi_], buf)

            _run_logical_ranks(gathers, issue)
            for r, gth in enumerate(gathers):
                assert mismatches(sharding.assemble(gth.tensor(buf), n), want_e if seg else want_s) == 0, (r, call)
        with pytest.raises(epp_b200.EppError, match="different buffer"):
            gathers[0].check_states(gw, d_xyz[:1024], 1)   # the previous call used buffer 1
        for gth in gathers:
            assert gth.status()[1] == 0
    finally:
        epp_b200.set_tuning("gather_bulk_min_words", 65536)
        epp_b200.set_tuning("segments_variant", 0)
        epp_b200.set_tuning("segments_block", 0)


@pytest.mark.parametrize("world_size", [2, 3])
def test_gather_deferred_form_logical_ranks(oracle, config, track, world_size):
    """Deferred exchange: call s sends its words and unpacks the peers' words of call s - 1; buffer s - 1 is complete
    after call s, the last one after flush().  Different data per call (so a stale row would show), ranks issued in
    varying order, 7 calls over two alternating buffers (rows are reused from call 3 on: the acknowledgements gate it)."""
    import torch
    from epp_b200 import sharding
    g, o, lo, hi = track
    ow, gw = build_pair(oracle, epp_b200, config, g, o)
    n = 150_001
    sw = sharding.slot_words(n, world_size)
    gathers = [epp_b200.Gather(r, world_size, sw, n_buffers=2).set_deferred(True) for r in range(world_size)]
    epp_b200.Gather.connect_local(gathers)
    batches, wants = [], []
    for call in range(7):
        xyz = worlds.gen_states(5000 + call, 0, n, lo, hi)
        batches.append(torch.from_numpy(xyz).cuda())
        wants.append(ow.check_states(xyz, threads=8))
    gw.check_states_dev(batches[0])
    torch.cuda.synchronize()
    for call in range(7):
        buf = call & 1
        def issue(r, gth):
            b, e = epp_b200.shard_range(n, r, world_size, 1024)
            gth.check_states(gw, batches[call][b:e], buf)
        order = gathers if call % 2 == 0 else gathers[::-1]
        _run_logical_ranks(order, lambda r_, gth: issue(gth.rank, gth))
        if call >= 1:   # the previous call's buffer is complete now
            for gth in gathers:
                assert mismatches(sharding.assemble(gth.tensor((call - 1) & 1), n), wants[call - 1]) == 0, (gth.rank, call)
    _run_logical_ranks(gathers, lambda r_, gth: gth.flush())
    for gth in gathers:
        assert mismatches(sharding.assemble(gth.tensor(6 & 1), n), wants[6]) == 0
        calls, timeouts = gth.status()
        assert calls == 7 and timeouts == 0
