#!/usr/bin/env python
"""Run ONE workload of the hot path a few times (for ncu captures and quick timings):

    python profiles/one_kernel.py <workload> [n=...] [reps=...] [key=value tuning knobs ...]

workloads: c2_states (track, 24 OBBs) | c3_segments (256 OBBs) | w256_states | c4_fused | c4_pre |
           c5_states | c5_edges (10 000 OBBs, grid) | minsnap
Prints microseconds per launch (CUDA events, direct launches) and the mask popcount / XOR checksum."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import epp_b200  # noqa: E402
from epp_b200 import worlds  # noqa: E402

wl = sys.argv[1]
kv = dict(a.split("=") for a in sys.argv[2:])
reps = int(kv.pop("reps", 5))
n = int(kv.pop("n", 0))
config = epp_b200.load_config()
dev = torch.device("cuda", 0)


def world(kind):
    g, o, lo, hi = {"track": lambda: worlds.track_world(config), "w256": worlds.world_256, "w10k": worlds.world_10k}[kind]()
    w = epp_b200.World(config)
    w.parse_gates_and_obstacles(g, o)
    return w, lo, hi


for k, v in kv.items():
    epp_b200.set_tuning(k, int(v))

summary = None
if wl in ("c2_states", "w256_states", "c5_states"):
    kind, seed = {"c2_states": ("track", worlds.SEED_STATES), "w256_states": ("w256", worlds.SEED_SEGMENTS),
                  "c5_states": ("w10k", worlds.SEED_CLUTTER)}[wl]
    w, lo, hi = world(kind)
    n = n or (1 << 24 if wl == "c5_states" else 1 << 20)
    ring = [epp_b200.gen_states_dev(seed, s * n, n, lo, hi) for s in range(max(2, min(16, (400 << 20) // (24 * n))))]
    mask = torch.empty((n + 31) // 32, dtype=torch.int32, device=dev)
    fn = lambda r: w.check_states_dev(ring[r % len(ring)], mask=mask)  # noqa: E731
    summary = lambda: epp_b200.mask_summary_dev(mask)  # noqa: E731
    units = n
elif wl in ("c3_segments", "c5_edges"):
    kind, seed = {"c3_segments": ("w256", worlds.SEED_SEGMENTS), "c5_edges": ("w10k", worlds.SEED_CLUTTER)}[wl]
    w, lo, hi = world(kind)
    n = n or (1 << 22 if wl == "c5_edges" else 1 << 20)
    ring = [epp_b200.gen_segments_dev(seed, s * n, n, lo, hi, 1.0) for s in range(max(2, min(8, (400 << 20) // (48 * n))))]
    mask = torch.empty((n + 31) // 32, dtype=torch.int32, device=dev)
    fn = lambda r: w.check_segments_dev(ring[r % len(ring)][0], ring[r % len(ring)][1], mask=mask)  # noqa: E731
    summary = lambda: epp_b200.mask_summary_dev(mask)  # noqa: E731
    units = n
elif wl in ("c4_fused", "c4_pre"):
    w, lo, hi = world("w256")
    n_traj = n or 10000
    if wl == "c4_fused":
        coeffs, T, dt = worlds.rest_to_rest_trajectories(n_traj, 8, lo, hi, n_samples=1000, max_hop=3.0)
        d_c, d_T, d_dt = (torch.from_numpy(x).to(dev) for x in (coeffs, T, dt))
        res = w.check_poly_trajectories_dev(d_c, d_T, d_dt, 0.1, max_samples=1024)
        torch.cuda.synchronize()
        units = int(res[2].sum().item())
        fn = lambda r: w.check_poly_trajectories_dev(d_c, d_T, d_dt, 0.1, max_samples=1024)  # noqa: E731
        summary = lambda: epp_b200.mask_summary_dev(res[1].view(-1))  # noqa: E731
    else:
        rows = 1000
        mats = torch.zeros((n_traj, 10, rows), dtype=torch.float64, device=dev)
        xyz = epp_b200.gen_states_dev(worlds.SEED_TRAJ, 0, n_traj * rows, lo, hi, device=dev).view(n_traj, rows, 3)
        for c in range(3):
            mats[:, 3 * c, :] = xyz[:, :, c]
        units = n_traj * rows
        fn = lambda r: w.check_trajectories_dev(mats, n_traj, rows, rows, 10 * rows, 10, 0.1)  # noqa: E731
elif wl == "minsnap":
    g, o, lo, hi = worlds.track_world(config)
    n_traj = n or 10000
    rng = np.random.default_rng(4)
    wp = rng.uniform(lo, hi, size=(n_traj, 9, 3))
    T = epp_b200.segment_times(wp, 1.0, 2.0)
    d_wp, d_T8 = torch.from_numpy(wp).to(dev), torch.from_numpy(T).to(dev)
    d_c8 = epp_b200.minsnap_dev(d_wp, d_T8)
    fn = lambda r: epp_b200.minsnap_dev(d_wp, d_T8, out=d_c8)  # noqa: E731
    units = n_traj
else:
    raise SystemExit("unknown workload " + wl)

for r in range(3):
    fn(r)
torch.cuda.synchronize()
best = 1e30
for _ in range(3):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for r in range(reps):
        fn(r)
    e.record()
    torch.cuda.synchronize()
    best = min(best, s.elapsed_time(e) * 1e3 / reps)
line = "%s units %d %s: %.2f us/launch, %.4g units/s" % (wl, units, kv, best, units / (best * 1e-6))
if summary is not None:
    fn(0)
    torch.cuda.synchronize()
    pop, xor = summary()
    line += "  popcount %d xor %016x" % (pop, xor)
print(line)
