#!/usr/bin/env python
"""K2 (segments vs the 256-OBB world, config 3) for 256 / 512 / 1024 threads per block, CUDA-event timed over a ring of
4 batches of 1 M segments (192 MiB > L2); masks compared between settings."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import epp_b200  # noqa: E402
from epp_b200 import worlds  # noqa: E402

BATCH = 1 << 20
config = epp_b200.load_config()
g, o, lo, hi = worlds.world_256()
w = epp_b200.World(config)
w.parse_gates_and_obstacles(g, o)
segs = [epp_b200.gen_segments_dev(worlds.SEED_SEGMENTS, s * BATCH, BATCH, lo, hi, 1.0) for s in range(4)]
masks = [torch.empty(BATCH // 32, dtype=torch.int32, device="cuda") for _ in range(4)]
ref = None
for tb in (256, 512, 1024, 256):
    epp_b200.set_tuning("segments_block", tb)
    for _ in range(3):
        for (a, b), m in zip(segs, masks):
            w.check_segments_dev(a, b, mask=m)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        for (a, b), m in zip(segs, masks):
            w.check_segments_dev(a, b, mask=m)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 40
    got = torch.stack(masks).cpu()
    if ref is None:
        ref = got
    print("segments_block %4d  %.2f us per 1M segments  %.3e segments/s  masks equal: %s" % (tb, us, BATCH / us * 1e6, bool(torch.equal(ref, got))))
