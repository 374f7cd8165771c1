#!/usr/bin/env python
"""Run K1 a few times with one tuning setting (for ncu): python profiles/one_states.py key=value ... [n=1048576] [world=track|w256|w10k]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import epp_b200  # noqa: E402
from epp_b200 import worlds  # noqa: E402

kv = dict(a.split("=") for a in sys.argv[1:])
n = int(kv.pop("n", 1 << 20))
world = kv.pop("world", "track")
reps = int(kv.pop("reps", 8))
config = epp_b200.load_config()
g, o, lo, hi = {"track": lambda: worlds.track_world(config), "w256": worlds.world_256, "w10k": worlds.world_10k}[world]()
w = epp_b200.World(config)
w.parse_gates_and_obstacles(g, o)
for k, v in kv.items():
    epp_b200.set_tuning(k, int(v))
ring = [epp_b200.gen_states_dev(worlds.SEED_STATES, s * n, n, lo, hi) for s in range(max(2, min(16, (400 << 20) // (24 * n))))]
mask = torch.empty((n + 31) // 32, dtype=torch.int32, device="cuda")
for r in range(reps):
    w.check_states_dev(ring[r % len(ring)], mask=mask)
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for r in range(reps):
    w.check_states_dev(ring[r % len(ring)], mask=mask)
e.record()
torch.cuda.synchronize()
print("n %d world %s %s: %.2f us/launch (direct launches)" % (n, world, kv, s.elapsed_time(e) * 1e3 / reps))
if os.environ.get("GRAPH", "1") == "1":
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for x in ring:
            w.check_states_dev(x, mask=mask)
    graph.replay()
    torch.cuda.synchronize()
    s.record()
    for _ in range(50):
        graph.replay()
    e.record()
    torch.cuda.synchronize()
    print("   graph of %d launches: %.2f us/launch" % (len(ring), s.elapsed_time(e) * 1e3 / (50 * len(ring))))
