"""Multi-rank path on CPU: world_size 2 (and 3) over gloo.  The check callable is the CPU oracle here (there
is no GPU in this container); what is under test is the sharding, the in-place all-gather layout and the
reassembly, which are the same code the NCCL run uses (efficient-path-planner_b200/sharding.py)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world_size, port, n, q):
    for p in (ROOT, os.path.join(ROOT, "oracle")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch
    import torch.distributed as dist
    import epp_b200
    import oracle_py
    from epp_b200 import sharding, worlds
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    config = epp_b200.load_config()
    g, o, lo, hi = worlds.track_world(config)
    ow = oracle_py.OracleWorld(config)
    ow.parse_gates_and_obstacles(g, o)

    def check_local(begin, end, out_words):
        # inputs generated from the rank's GLOBAL unit indices, like the on-device generators of config 5
        xyz = worlds.gen_states(worlds.SEED_CLUTTER, begin, end - begin, lo, hi)
        out_words.copy_(torch.from_numpy(ow.check_states(xyz).view(np.int32)))

    gathered = sharding.sharded_check(n, check_local, rank, world_size)
    full = sharding.assemble(gathered, n)
    pop, x = epp_b200.mask_summary_host(full)
    # per-shard summaries combine to the whole-batch summary (what the GPU run compares across G)
    b, e = sharding.shard_range(n, rank, world_size)
    mine = full[b // 32: (e + 31) // 32] if e > b else full[:0]
    p_loc, x_loc = epp_b200.mask_summary_host(mine, b // 32)
    t = torch.tensor([p_loc, x_loc & 0x7FFFFFFFFFFFFFFF, x_loc >> 63], dtype=torch.int64)
    parts = [torch.zeros_like(t) for _ in range(world_size)]
    dist.all_gather(parts, t)
    p_sum = sum(int(p[0]) for p in parts)
    x_all = 0
    for p in parts:
        x_all ^= int(p[1]) | (int(p[2]) << 63)
    q.put((rank, full.tobytes(), pop, x, p_sum, x_all))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world_size,n", [(2, 50000), (2, 1024), (2, 1000), (3, 70001), (2, 0)])
def test_sharded_masks_equal_single_process(oracle, config, world_size, n):
    import torch.multiprocessing as mp
    import epp_b200
    from epp_b200 import worlds
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world_size, port, n, q)) for r in range(world_size)]
    [p.start() for p in procs]
    results = [q.get(timeout=240) for _ in procs]
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    g, o, lo, hi = worlds.track_world(config)
    ow = oracle.OracleWorld(config)
    ow.parse_gates_and_obstacles(g, o)
    want = ow.check_states(worlds.gen_states(worlds.SEED_CLUTTER, 0, n, lo, hi)) if n else np.zeros(0, np.uint32)
    wpop, wx = epp_b200.mask_summary_host(want)
    for rank, full, pop, x, p_sum, x_all in results:
        assert np.array_equal(np.frombuffer(full, dtype=np.uint32), want), rank
        assert (pop, x) == (wpop, wx) and (p_sum, x_all) == (wpop, wx), rank


def test_shard_range_twin_matches_c_abi():
    import epp_b200
    from epp_b200 import sharding
    for n in (0, 1, 31, 1024, 1025, 50000, (1 << 20) + 7, 10**9):
        for ws in (1, 2, 3, 4, 8):
            covered = 0
            for r in range(ws):
                assert sharding.shard_range(n, r, ws) == epp_b200.shard_range(n, r, ws, 1024)
                b, e = sharding.shard_range(n, r, ws)
                assert b == covered and (b % 32 == 0 or b == e == n)
                covered = e
            assert covered == n
            assert sharding.slot_words(n, ws) * 32 * ws >= n
