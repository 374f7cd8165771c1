#!/usr/bin/env python
"""Per-warp timeline of the warp form of K1 (needs `make -C efficient-path-planner_b200/csrc trace`):
EPP_B200_LIB=efficient-path-planner_b200/libepp_b200_trace.so python profiles/trace_states.py key=value ...
Prints, relative to the earliest warp entry of one launch, percentiles (min / p50 / p90 / max, in us) of every stamp."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import epp_b200  # noqa: E402
from epp_b200 import worlds  # noqa: E402

kv = dict(a.split("=") for a in sys.argv[1:])
n = int(kv.pop("n", 1 << 20))
config = epp_b200.load_config()
g, o, lo, hi = worlds.track_world(config)
w = epp_b200.World(config)
w.parse_gates_and_obstacles(g, o)
for k, v in kv.items():
    epp_b200.set_tuning(k, int(v))
ring = [epp_b200.gen_states_dev(worlds.SEED_STATES, s * n, n, lo, hi) for s in range(16)]
mask = torch.empty((n + 31) // 32, dtype=torch.int32, device="cuda")
L = epp_b200.lib()
L.epp_debug_trace_copy.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
buf = np.zeros(1 << 20, dtype=np.uint64)
for r in range(6):
    w.check_states_dev(ring[r], mask=mask)
torch.cuda.synchronize()
L.epp_debug_trace_copy(buf.ctypes.data, buf.size, 1)
names = ["entry", "records stored", "consumer sync", "table here", "data0", "step0 done", "data1", "step1 done", "data2", "step2 done", "data3", "step3 done", "d4", "s4", "d5", "s5"]
names_old = ["entry", "tma issued", "records built", "block sync", "table here", "data0", "step0 done", "data1",
         "step1 done", "data2", "step2 done", "data3", "step3 done", "data4", "step4 done", "data5"]
names[11], names[12], names[13], names[14], names[15] = "prod. bar init", "bar init done", "producer start", "kernel start", "exit"
L.epp_debug_trace_base.argtypes = [ctypes.c_uint, ctypes.c_void_p]
st = torch.cuda.current_stream().cuda_stream
for rep in range(2):
    # two back-to-back launches, each into its own half of the trace buffer
    L.epp_debug_trace_base(0, st)
    w.check_states_dev(ring[8 + 2 * rep], mask=mask)
    L.epp_debug_trace_base(1 << 19, st)
    w.check_states_dev(ring[9 + 2 * rep], mask=mask)
    torch.cuda.synchronize()
    L.epp_debug_trace_copy(buf.ctypes.data, buf.size, 1)
    halves = []
    for h in range(2):
        t = buf[h << 19:(h + 1) << 19].reshape(-1, 16).astype(np.int64)
        halves.append(t[t[:, 14] > 0])
    t0 = halves[0][:, 14].min()
    for h, t in enumerate(halves):
        print("pair %d launch %d: %d warps traced %s" % (rep, h, len(t), kv))
        for e in [13, 14, 11, 12] + list(range(11)) + [15]:
            col = t[:, e]
            col = col[col > 0]
            if len(col) == 0:
                continue
            rel = (col - t0) / 1e3
            print("  %-14s n=%5d  min %6.2f  p50 %6.2f  p90 %6.2f  max %6.2f us" %
                  (names[e], len(col), rel.min(), np.percentile(rel, 50), np.percentile(rel, 90), rel.max()))
