import sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/oracle'); sys.path.insert(0, '/root/repo/tests')
import numpy as np, torch, epp_b200, oracle_py
from epp_b200 import worlds, sharding
from test_gpu_parity import _run_logical_ranks
config = epp_b200.load_config()
g, o, lo, hi = worlds.track_world(config)
gw = epp_b200.World(config); gw.parse_gates_and_obstacles(g, o)
ow = oracle_py.OracleWorld(config); ow.parse_gates_and_obstacles(g, o)
import itertools
for rep, (world_size, n) in itertools.product(range(12), ((3, 100000), (2, 100000), (3, 1 << 20))):
    xyz = worlds.gen_states(worlds.SEED_STATES, 777, n, lo, hi)
    want = ow.check_states(xyz, threads=8)
    d_xyz = torch.from_numpy(xyz).cuda()
    sw = sharding.slot_words(n, world_size)
    gathers = [epp_b200.Gather(r, world_size, sw, n_buffers=2) for r in range(world_size)]
    epp_b200.Gather.connect_local(gathers)
    for call in range(4):
        buf = call & 1
        for gth in gathers: gth.tensor(buf).fill_(-1)
        torch.cuda.synchronize()
        def issue(r, gth):
            b, e = epp_b200.shard_range(n, r, world_size, 1024)
            gth.check_states(gw, d_xyz[b:e], buf)
        _run_logical_ranks(gathers, issue)
        for r, gth in enumerate(gathers):
            got = sharding.assemble(gth.tensor(buf), n)
            bad = np.nonzero(got ^ want)[0]
            if len(bad): print("world", world_size, "n", n, "call", call, "rank", r, "bad words", len(bad), bad[:12], bad[-6:] if len(bad) else "", "status", gth.status())
print("done")
