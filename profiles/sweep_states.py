#!/usr/bin/env python
"""Tuning sweep of K1 (states) on a B200: form x chunk size x block shape, CUDA-event timed from
a CUDA graph of one pass over a ring of batches larger than L2.  Prints one line per setting:
    variant words/chunk blocks/SM threads pdl  us/launch  GB/s(algorithmic)  frac of measured HBM peak
Usage: python profiles/sweep_states.py [workload=states|clutter] [batch_log2=20]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import epp_b200  # noqa: E402
from epp_b200 import worlds  # noqa: E402


def main():
    workload = sys.argv[1] if len(sys.argv) > 1 else "states"
    batch = 1 << (int(sys.argv[2]) if len(sys.argv) > 2 else 20)
    ring_n = max(4, (400 << 20) // (batch * 24))
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    config = epp_b200.load_config()
    if workload == "states":
        g, o, lo, hi = worlds.track_world(config)
    elif workload == "w256":
        g, o, lo, hi = worlds.world_256()
    else:
        g, o, lo, hi = worlds.world_10k()
    w = epp_b200.World(config)
    w.parse_gates_and_obstacles(g, o)
    ring = [epp_b200.gen_states_dev(worlds.SEED_STATES, s * batch, batch, lo, hi) for s in range(ring_n)]
    mask = torch.empty(batch // 32, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    # (variant, words per chunk, blocks/SM, threads per block, programmatic dependent launch)
    settings = [(1, 0, b, 0, 0) for b in (4, 8)]
    settings += [(3, spt, b, thr, pdl) for pdl in (0, 1) for spt in (4, 8)
                 for (b, thr) in ((1, 1024), (1, 768), (1, 512), (2, 512), (2, 1024), (4, 256))]
    print("workload %s batch %d ring %d world_obbs %d bp %s" % (workload, batch, ring_n, w.size(),
                                                              epp_b200.BP_NAMES[w.active_broadphase()]))
    for variant, spt, bps, thr, pdl in settings:
        epp_b200.set_tuning("states_variant", variant)
        if variant == 1:
            epp_b200.set_tuning("states_blocks_per_sm", bps)
        else:
            epp_b200.set_tuning("states_per_thread", spt)
            epp_b200.set_tuning("states_warp_blocks_per_sm", bps)
            epp_b200.set_tuning("states_warp_threads", thr)
            epp_b200.set_tuning("states_pdl", pdl)
        try:
            for x in ring[:3]:
                w.check_states_dev(x, mask=mask)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                for x in ring:
                    w.check_states_dev(x, mask=mask)
            graph.replay()
            torch.cuda.synchronize()
            reps = max(4, 2000 // ring_n)
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(reps):
                graph.replay()
            e.record()
            torch.cuda.synchronize()
            us = s.elapsed_time(e) * 1e3 / (reps * ring_n)
            gbs = batch * 24.125 / (us * 1e-6) / 1e9
            print("variant %d words/chunk %d blocks/SM %d threads %4d pdl %d  %8.2f us  %8.1f GB/s  frac %.3f" %
                  (variant, spt, bps, thr, pdl, us, gbs, gbs / peak), flush=True)
        except Exception as ex:
            print("variant %d spt %d blocks/SM %d threads %d pdl %d failed: %r" % (variant, spt, bps, thr, pdl, ex), flush=True)


if __name__ == "__main__":
    main()
