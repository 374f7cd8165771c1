"""Synthetic worlds and inputs of the BASELINE.json configs.

Everything here produces plain numpy arrays (gate / obstacle pose rows, states, segments,
polynomial coefficients) that are fed unchanged to the GPU path and to the CPU oracle.  The
counter-based generator below is the same splitmix64 finaliser the device generators use, so
"generated on device" inputs can be reproduced on the host bit for bit.
"""
import math

import numpy as np

SEED_STATES = 0x5EED0002
SEED_SEGMENTS = 0x5EED0003
SEED_TRAJ = 0x5EED0004
SEED_CLUTTER = 0x5EED0005

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def mix64(seed, counter):
    counter = np.asarray(counter, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = np.uint64(seed) + np.uint64(0x9E3779B97F4A7C15) * (counter + np.uint64(1))
        z = z ^ (z >> np.uint64(30))
        z = z * np.uint64(0xBF58476D1CE4E5B9)
        z = z ^ (z >> np.uint64(27))
        z = z * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def u01(seed, index, lane):
    """Uniform [0,1) double for (index, lane): (mix64(seed, index*8+lane) >> 11) * 2^-53."""
    index = np.asarray(index, dtype=np.uint64)
    with np.errstate(over="ignore"):
        c = index * np.uint64(8) + np.uint64(lane)
    return (mix64(seed, c) >> np.uint64(11)).astype(np.float64) * 2.0 ** -53


def gen_states(seed, first, n, lo, hi):
    """Host twin of epp_gen_states_dev: x = lo + u * (hi - lo), two rounded operations."""
    idx = np.arange(first, first + n, dtype=np.uint64)
    lo = np.asarray(lo, dtype=np.float64)
    hi = np.asarray(hi, dtype=np.float64)
    out = np.empty((n, 3))
    for d in range(3):
        out[:, d] = lo[d] + u01(seed, idx, d) * (hi[d] - lo[d])
    return out


def gen_segments(seed, first, n, lo, hi, max_len):
    """Host twin of epp_gen_segments_dev (same operation order)."""
    idx = np.arange(first, first + n, dtype=np.uint64)
    lo = np.asarray(lo, dtype=np.float64)
    hi = np.asarray(hi, dtype=np.float64)
    a = np.empty((n, 3))
    v = np.empty((n, 3))
    for d in range(3):
        a[:, d] = lo[d] + u01(seed, idx, d) * (hi[d] - lo[d])
        v[:, d] = 2.0 * u01(seed, idx, 3 + d) - 1.0
    len2 = (v[:, 0] * v[:, 0] + v[:, 1] * v[:, 1]) + v[:, 2] * v[:, 2]
    norm = np.sqrt(len2)
    L = max_len * (1.0 - u01(seed, idx, 6))
    b = np.empty((n, 3))
    safe = np.where(norm > 0.0, norm, 1.0)
    for d in range(3):
        u = np.where(norm > 0.0, v[:, d] / safe, 0.0)
        e = a[:, d] + L * u
        b[:, d] = np.minimum(np.maximum(e, lo[d]), hi[d])
    return a, b


# ---------------------------------------------------------------------------------- worlds

def track_world(config):
    """Config 1/2: the builder-defined 4-gate + 4-obstacle track (M = 24)."""
    t = config["track"]
    gates = np.asarray(t["gates"], dtype=np.float64)
    obstacles = np.asarray(t["obstacles"], dtype=np.float64)
    wp = config["world_properties"]
    return gates, obstacles, np.asarray(wp["lower_bound"], float), np.asarray(wp["upper_bound"], float)


def lattice_world(n_gates, n_obstacles, half_extent, seed, z_top=4.0):
    """Jittered square lattice of gates (alternating type, yaw in (-pi, pi]) and obstacles."""
    total = n_gates + n_obstacles
    side = int(math.ceil(math.sqrt(total)))
    cell = 2.0 * half_extent / side
    k = np.arange(total, dtype=np.uint64)
    # deterministic permutation of lattice sites: sort by a hash
    order = np.argsort(mix64(seed ^ 0xA5A5, k), kind="stable")
    site = np.empty(total, dtype=np.int64)
    site[order] = np.arange(total)
    ix, iy = site % side, site // side
    jx = (u01(seed, k, 0) - 0.5) * 0.6 * cell
    jy = (u01(seed, k, 1) - 0.5) * 0.6 * cell
    x = -half_extent + (ix + 0.5) * cell + jx
    y = -half_extent + (iy + 0.5) * cell + jy
    yaw = (1.0 - 2.0 * u01(seed, k, 2)) * math.pi  # (-pi, pi]
    gates = np.zeros((n_gates, 7))
    gates[:, 0], gates[:, 1], gates[:, 5] = x[:n_gates], y[:n_gates], yaw[:n_gates]
    gates[:, 6] = np.arange(n_gates) % 2
    obstacles = np.zeros((n_obstacles, 6))
    obstacles[:, 0], obstacles[:, 1], obstacles[:, 5] = x[n_gates:], y[n_gates:], yaw[n_gates:]
    lo = np.array([-half_extent, -half_extent, 0.0])
    hi = np.array([half_extent, half_extent, z_top])
    return gates, obstacles, lo, hi


def world_256():
    """Config 3: 40 gates + 56 obstacles = 256 OBBs inside [-20,20]^2 x [0,4]."""
    return lattice_world(40, 56, 20.0, SEED_SEGMENTS)


def world_10k():
    """Config 5: 1600 gates + 2000 obstacles = 10 000 OBBs inside [-200,200]^2 x [0,4]."""
    return lattice_world(1600, 2000, 200.0, SEED_CLUTTER)


# ---------------------------------------------------------------------- trajectories (config 4)

# s(tau) = 126 tau^5 - 420 tau^6 + 540 tau^7 - 315 tau^8 + 70 tau^9: the unique degree-9 rest-to-rest
# profile with zero velocity .. snap at both ends (Vertex::makeStartOrEnd(.., SNAP), vertex.cpp:146-170).
_REST_TO_REST = np.array([0.0, 0.0, 0.0, 0.0, 0.0, 126.0, -420.0, 540.0, -315.0, 70.0])


def nfabian_time(distance, v_max, a_max, magic=6.5):
    """estimateSegmentTimesNfabian, external/poly_traj/src/vertex.cpp:272-289."""
    return distance / v_max * 2 * (1.0 + magic * v_max / a_max * np.exp(-distance / v_max * 2))


def rest_to_rest_trajectories(n_traj, n_segments, lo, hi, seed=SEED_TRAJ, v_max=1.0, a_max=2.0,
                              n_samples=1000, max_hop=None):
    """Synthetic min-snap-shaped trajectories: per segment p(t) = p0 + D * s(t / T).

    Returns coeffs (n_traj, K, 3, 10), durations (n_traj, K), dt (n_traj,) with
    dt = sum(T) / (n_samples - 1) so evaluateRange yields ~n_samples rows.
    """
    lo = np.asarray(lo, float)
    hi = np.asarray(hi, float)
    K = n_segments
    idx = np.arange(n_traj * (K + 1), dtype=np.uint64)
    wp = np.empty((n_traj * (K + 1), 3))
    for d in range(3):
        wp[:, d] = lo[d] + u01(seed, idx, d) * (hi[d] - lo[d])
    wp = wp.reshape(n_traj, K + 1, 3)
    if max_hop is not None:  # keep hops short in large worlds: random walk clipped to the bounds
        step = (wp[:, 1:] - 0.5 * (lo + hi)) / (0.5 * (hi - lo)) * max_hop
        for k in range(1, K + 1):
            wp[:, k] = np.clip(wp[:, k - 1] + step[:, k - 1], lo, hi)
    delta = wp[:, 1:] - wp[:, :-1]                      # (n_traj, K, 3)
    dist = np.sqrt((delta ** 2).sum(axis=2))
    T = np.maximum(nfabian_time(dist, v_max, a_max), 1e-3)  # (n_traj, K)
    coeffs = np.zeros((n_traj, K, 3, 10))
    coeffs[..., 0] = wp[:, :-1]
    for j in range(5, 10):
        coeffs[..., j] = delta * _REST_TO_REST[j] / (T[..., None] ** j)
    dt = T.sum(axis=1) / (n_samples - 1)
    return np.ascontiguousarray(coeffs), np.ascontiguousarray(T), np.ascontiguousarray(dt)


def traj_matrix_from_positions(xyz, times):
    """N x 10 trajectory matrix of poly_traj::generateTrajectory
    (external/poly_traj/src/trajectory_generator.cpp:82-96): position in columns 0, 3, 6, time in 9
    (velocity / acceleration columns left zero, as TrajInterpolation does)."""
    n = xyz.shape[0]
    m = np.zeros((n, 10))
    m[:, 0], m[:, 3], m[:, 6], m[:, 9] = xyz[:, 0], xyz[:, 1], xyz[:, 2], times
    return m


def adversarial_states(obbs, n, seed=0x5EED0A2):
    """n states on, and 1-2 ulp either side of, the faces of the stored AABBs and of the (inflated) OBBs of a world
    (where a rounding difference between two implementations of the predicates would show).

    obbs: records as returned by World.obbs() (center, half, rot row-major, aabb_min, aabb_max, inflate, filling).
    State i cycles over (OBB, AABB face | OBB face, axis, side, ulp offset in -2..2); its position on the face comes
    from the counter-based generator, so the batch is reproducible bit for bit."""
    M = len(obbs)
    mn = np.array([o["aabb_min"] for o in obbs], dtype=np.float64)
    mx = np.array([o["aabb_max"] for o in obbs], dtype=np.float64)
    c = np.array([o["center"] for o in obbs], dtype=np.float64)
    R = np.array([o["rot"] for o in obbs], dtype=np.float64).reshape(M, 3, 3)
    H = np.array([np.asarray(o["half"], float) + (0.0 if o["filling"] else o["inflate"]) for o in obbs])
    i = np.arange(n, dtype=np.uint64)
    k = (i % np.uint64(M)).astype(np.int64)
    j = (i // np.uint64(M)).astype(np.int64)
    kind, axis, side, ulp = j % 2, (j // 2) % 3, (j // 6) % 2, (j // 12) % 5 - 2
    u = np.stack([u01(seed, i, d) for d in range(3)], axis=1)
    rows = np.arange(n)
    # stored-AABB faces
    p_box = mn[k] + u * (mx[k] - mn[k])
    p_box[rows, axis] = np.where(side == 0, mn[k, axis], mx[k, axis])
    # OBB faces: local point on the face, rotated into the world
    loc = (2.0 * u - 1.0) * H[k]
    loc[rows, axis] = np.where(side == 0, -1.0, 1.0) * H[k, axis]
    p_obb = np.einsum("nij,nj->ni", R[k], loc) + c[k]
    p = np.where((kind == 0)[:, None], p_box, p_obb)
    v = p[rows, axis]
    for step in (1, 2):
        v = np.where(ulp >= step, np.nextafter(v, np.inf), v)
        v = np.where(ulp <= -step, np.nextafter(v, -np.inf), v)
    p[rows, axis] = v
    return np.ascontiguousarray(p)
