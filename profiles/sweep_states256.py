"""K1 on the 256-OBB world: records in shared memory (states_rec_smem_max=256) vs global records (=64)."""
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
B = 1 << 20
pts = [epp_b200.gen_states_dev(worlds.SEED_SEGMENTS, s * B, B, lo, hi, device=dev) for s in range(8)]
m = torch.empty(B // 32, dtype=torch.int32, device=dev)
ref = None
for mx in (64, 256):
    epp_b200.set_tuning("states_rec_smem_max", mx)
    out = w.check_states_dev(pts[0]).clone()
    ref = out if ref is None else ref
    assert torch.equal(ref, out)
    for _ in range(3):
        [w.check_states_dev(p, mask=m) for p in pts]
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); s.record()
    for _ in range(10):
        [w.check_states_dev(p, mask=m) for p in pts]
    e.record(); torch.cuda.synchronize()
    us = s.elapsed_time(e) / 80 * 1e3
    print(json.dumps({"states_rec_smem_max": mx, "us_per_1M": us, "states_per_s": B / us * 1e6}))
