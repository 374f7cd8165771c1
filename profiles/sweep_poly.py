"""K3b sweep: trajectories per warp (tuning knob poly_traj_per_warp) on config 4 (10k trajectories x ~1k samples)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import torch
import epp_b200
from epp_b200 import worlds

dev = torch.device("cuda", 0)
config = epp_b200.load_config()
g, o, lo, hi = worlds.world_256()
w = epp_b200.World(config)
w.parse_gates_and_obstacles(g, o)
n_traj = 10000
coeffs, T, dt = worlds.rest_to_rest_trajectories(n_traj, 8, lo, hi, n_samples=1000, max_hop=3.0)
d_c, d_T, d_dt = (torch.from_numpy(x).to(dev) for x in (coeffs, T, dt))
ref = None
for G in (0, 1, 2, 4, 8, 16, 32):
    epp_b200.set_tuning("poly_traj_per_warp", G)
    out = w.check_poly_trajectories_dev(d_c, d_T, d_dt, 0.1, max_samples=1024)
    torch.cuda.synchronize()
    if ref is None:
        ref = [x.clone() for x in out]
    else:
        assert all(torch.equal(a, b) for a, b in zip(ref, out)), G
    samples = int(out[2].sum().item())
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(5):
        w.check_poly_trajectories_dev(d_c, d_T, d_dt, 0.1, max_samples=1024)
    e.record()
    torch.cuda.synchronize()
    ms = s.elapsed_time(e) / 5
    print(json.dumps({"G": G, "ms": ms, "samples_per_s": samples / (ms * 1e-3)}))
