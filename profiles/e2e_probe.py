#!/usr/bin/env python
"""Where the host-buffer call spends its time: python profiles/e2e_probe.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, epp_b200
from epp_b200 import worlds
config = epp_b200.load_config()
g, o, lo, hi = worlds.track_world(config)
w = epp_b200.World(config); w.parse_gates_and_obstacles(g, o)
L = epp_b200.lib()
n = 1 << 20
ring = [epp_b200.gen_states_dev(worlds.SEED_STATES, s * n, n, lo, hi).cpu().pin_memory() for s in range(8)]
mask = torch.zeros(n // 32, dtype=torch.int32).pin_memory()
d = torch.empty((n, 3), dtype=torch.float64, device="cuda")
def wall(fn, reps=40):
    fn(0); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(reps): fn(i)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e6
print("H2D 24 MiB copy + sync: %.1f us" % wall(lambda i: (d.copy_(ring[i % 8], non_blocking=True), torch.cuda.synchronize())))
print("epp_check_states (host buffers): %.1f us" % wall(lambda i: L.epp_check_states(w.handle, ring[i % 8].data_ptr(), n, 0, mask.data_ptr())))
dm = torch.empty(n // 32, dtype=torch.int32, device="cuda")
def manual(i):
    d.copy_(ring[i % 8], non_blocking=True)
    w.check_states_dev(d, mask=dm)
    mask.copy_(dm, non_blocking=True)
    torch.cuda.synchronize()
print("torch copy + epp_check_states_dev + D2H + sync: %.1f us" % wall(manual))
for k in (2, 4, 8):
    m = n // k
    print("host call on %d states x %d: %.1f us per 1M" % (m, k, wall(lambda i: [L.epp_check_states(w.handle, ring[i % 8].data_ptr() + j * m * 24, m, 0, mask.data_ptr() + j * (m // 8)) for j in range(k)])))
