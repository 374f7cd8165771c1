#!/bin/bash
OUT=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "multi_batch or segment or motion or edges" > $OUT/r02_segmulti_pytest.log 2>&1; echo "pytest exit $?" >> $OUT/r02_segmulti_pytest.log
tail -4 $OUT/r02_segmulti_pytest.log
timeout 600 python - > $OUT/r02_segmulti_ab.log 2>&1 <<'PY'
import sys; sys.path.insert(0, '.')
import torch, epp_b200
from epp_b200 import worlds
dev = torch.device('cuda', 0)
config = epp_b200.load_config()
def timeit(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(4):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(reps): fn()
        e.record(); torch.cuda.synchronize()
        best = min(best, s.elapsed_time(e) * 1e3 / reps)
    return best
for name, mk, n in (("w256", worlds.world_256, 1 << 20), ("track", lambda: worlds.track_world(config), 1 << 20), ("w10k", worlds.world_10k, 1 << 22)):
    g, o, lo, hi = mk()
    w = epp_b200.World(config); w.parse_gates_and_obstacles(g, o)
    segs = [epp_b200.gen_segments_dev(77, s * n, n, lo, hi, 1.0, device=dev) for s in range(4)]
    m8 = [torch.empty(n // 32, dtype=torch.int32, device=dev) for _ in range(8)]
    single = timeit(lambda: [w.check_segments_dev(a, b, mask=m8[k]) for k, (a, b) in enumerate(segs + segs)]) / 8
    line = "%s n=%d: single calls %.2f us/batch" % (name, n, single)
    for bps in (2, 1, 3):
        epp_b200.set_tuning("segments_multi_blocks_per_sm", bps)
        t = timeit(lambda: w.check_segments_multi_dev(segs + segs, m8)) / 8
        line += ", multi bps=%d %.2f" % (bps, t)
    epp_b200.set_tuning("segments_multi_blocks_per_sm", 2)
    print(line, "(%.3g segments/s at bps=2)" % (n / (timeit(lambda: w.check_segments_multi_dev(segs + segs, m8)) / 8 * 1e-6)))
PY
cat $OUT/r02_segmulti_ab.log
