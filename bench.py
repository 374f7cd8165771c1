#!/usr/bin/env python
"""bench.py — validity checks per second of the B200 path (and of the CPU reference arm).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload ...]

Default workload = BASELINE.json configs[1]: 1 M uniform random states isValid
(World::checkPointValidity(point, canPassGate=false), src/World.cpp:80-104) against the
config.json track world (4 gates + 4 obstacles = 24 OBBs).  A step = one pass of the hot path
over one 1 M-state batch.  Batches rotate over a ring of distinct buffers whose total size
exceeds the 126 MB L2, so every step streams its states from HBM.

  value     device-resident inputs, K steps timed with CUDA events on the launching stream
            (steps are replayed from a CUDA graph of one ring pass so the Python launch cost
            does not hide the kernel), max over ranks.
  e2e       the same batches through the host-buffer C-ABI call epp_check_states(): pinned host
            states -> H2D -> kernel -> D2H mask -> host, every step, copies inside the region.
  roofline  dominant kernel (k_states_warp) algorithmic bytes / its mean launch duration vs the
            measured HBM copy bandwidth in MEASURED_PEAKS.json.
  cpu_baseline  the CPU oracle (restatement of the reference validators), all host cores, on a
            bounded sample of the same workload.  Rank 0, N=1 only.

--impl reference times the reference's CPU implementation of the same path on the host cores
(oracle/_ref = the reference's own sources when built, else the oracle port).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

BATCH = 1 << 20           # "1M random states" (BASELINE.json configs[1])
RING = 16                 # 16 x 24 MiB = 384 MiB of states > 126 MB L2
BYTES_PER_STATE = 24.125  # 3 doubles read + 1 mask bit written
BYTES_PER_SEGMENT = 48.125

WORKLOADS = {
    "states": "config2: 1M uniform states isValid (canPassGate=false) vs config.json track world, M=24 OBBs",
    "segments": "config3: 1M segments checkMotion (exact slab test) vs synthetic 256-OBB world",
    "config5": "config5: 1B states isValid + 100M edges checkMotion vs synthetic 10k-OBB world, "
               "index ranges sharded over the ranks, NCCL all-gather of the mask words",
}
C5_STATES = 1_000_000_000
C5_EDGES = 100_000_000


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None, help="default 4000 (config5: 20)")
    ap.add_argument("--warmup", type=int, default=None, help="default 64 (config5: 3)")
    ap.add_argument("--c5-states", type=int, default=C5_STATES)
    ap.add_argument("--c5-edges", type=int, default=C5_EDGES)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="states", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl", "none", "none-ipc"],
                    help="N > 1: mask exchange fused into the check kernel (default) or NCCL all-gather")
    ap.add_argument("--tune", action="append", default=[], metavar="KEY=VALUE",
                    help="epp_set_tuning knob (profiling sweeps; the default run sets none)")
    ap.add_argument("--no-verify", action="store_true", help="profiling sweeps with gather_debug only")
    args = ap.parse_args()
    if args.steps is None:
        args.steps = 20 if args.workload == "config5" else 4000
    if args.warmup is None:
        args.warmup = 3 if args.workload == "config5" else 64
    return args


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------ workloads

def make_world_inputs(workload):
    import epp_b200
    from epp_b200 import worlds
    config = epp_b200.load_config()
    if workload == "states":
        g, o, lo, hi = worlds.track_world(config)
        seed = worlds.SEED_STATES
    elif workload == "config5":
        g, o, lo, hi = worlds.world_10k()
        seed = worlds.SEED_CLUTTER
    else:
        g, o, lo, hi = worlds.world_256()
        seed = worlds.SEED_SEGMENTS
    return config, g, o, lo, hi, seed


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed regions run."""

    def __init__(self, index, period=0.01):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.active = [], set(), False
        self._stop_evt = threading.Event()
        self.sm_max = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
            nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake",
        }
        while not self._stop_evt.is_set():
            if self.active:
                try:
                    self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    for bit, name in names.items():
                        if r & bit:
                            self.reasons.add(name)
                except Exception:
                    pass
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons), "samples": len(s)}


# ------------------------------------------------------------------------------ B200 arm

def dist_env():
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)),
            int(os.environ.get("LOCAL_RANK", 0)))


def timed_events(torch, fn):
    """Run fn() between two CUDA events on the current stream; returns milliseconds."""
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    s.record()
    fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e)


def run_b200(args):
    import torch
    import torch.distributed as dist
    import epp_b200

    rank, world, local_rank = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = epp_b200.lib()
    for kv in args.tune:
        k, v = kv.split("=")
        epp_b200.set_tuning(k, int(v))

    config, g, o, lo, hi, seed = make_world_inputs(args.workload)
    gw = epp_b200.World(config)
    gw.parse_gates_and_obstacles(g, o)
    segs = args.workload == "segments"
    unit_bytes = BYTES_PER_SEGMENT if segs else BYTES_PER_STATE
    words = BATCH // 32

    # Ring of distinct batches.  Global unit index of (slot s, rank r, unit i) = (s*world + r)*BATCH + i:
    # every rank owns a contiguous index range of each global batch.
    ring = []
    for s in range(RING):
        first = (s * world + rank) * BATCH
        if segs:
            ring.append(epp_b200.gen_segments_dev(seed, first, BATCH, lo, hi, 1.0, device=dev))
        else:
            ring.append((epp_b200.gen_states_dev(seed, first, BATCH, lo, hi, device=dev),))
    # Gathered masks, double buffered.  N > 1: the check kernel itself stores its mask words into row r of
    # the gathered buffer of EVERY rank over NVLink and ends when the peers' rows have arrived
    # (epp_gather) — no collective launch per step.  --exchange nccl keeps the older form
    # (kernel writes row r locally, in-place NCCL all-gather behind it) for comparison.
    gather = None
    exchange = "none (single GPU)"
    ipc_probe = None
    if world > 1 and args.exchange == "none-ipc":  # profiling only: plain kernels writing into IPC-exported memory
        ipc_probe = epp_b200.Gather(rank, world, words, n_buffers=2).connect()
    if world > 1 and args.exchange == "fused":
        try:
            gather = epp_b200.Gather(rank, world, words, n_buffers=2).connect()
            gathered = [gather.tensor(0, dev), gather.tensor(1, dev)]
            exchange = ("fused: every mask word goes to every peer as a (word, call number) pair straight from "
                        "the check kernel; its warps unpack the peers' pairs; no collective launch")
        except epp_b200.EppError as ex:  # CUDA IPC not permitted on this box: say so, use NCCL
            gather = None
            exchange = "nccl all-gather (fused exchange unavailable: %s)" % ex
    if gather is None:
        gathered = [torch.zeros((world, words), dtype=torch.int32, device=dev) for _ in range(2)]
        if ipc_probe is not None:
            gathered = [ipc_probe.tensor(0, dev), ipc_probe.tensor(1, dev)]
        if world > 1 and not exchange.startswith("nccl"):
            exchange = "nccl all-gather, in place"
    torch.cuda.synchronize()

    def step_dev(s):
        buf = gathered[s & 1]
        mine = buf[rank]
        if gather is not None:
            if segs:
                gather.check_segments(gw, ring[s % RING][0], ring[s % RING][1], s & 1)
            else:
                gather.check_states(gw, ring[s % RING][0], s & 1)
            return
        if segs:
            gw.check_segments_dev(ring[s % RING][0], ring[s % RING][1], mask=mine)
        else:
            gw.check_states_dev(ring[s % RING][0], mask=mine)
        if world > 1 and not args.exchange.startswith("none"):  # "none*": profiling only, the ranks never exchange
            dist.all_gather_into_tensor(buf.view(-1), mine)

    K, W = args.steps, args.warmup
    sampler = ClockSampler(local_rank)
    sampler.start()

    # ---- value: device-resident inputs ------------------------------------------------------
    use_graph = (world == 1 or gather is not None or args.exchange.startswith("none")) and not args.no_graph and K >= RING
    for s in range(max(W, 3)):
        step_dev(s)
    torch.cuda.synchronize()
    launches0 = epp_b200.kernel_launch_count()
    graph = None
    if use_graph:
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            for s in range(RING):
                step_dev(s)
        graph.replay()  # warm the instantiated graph
        torch.cuda.synchronize()
    captured = epp_b200.kernel_launch_count() - launches0  # launches recorded into the graph
    launches1 = epp_b200.kernel_launch_count()

    def timed_region():
        done = 0
        if graph is not None:
            for _ in range(K // RING):
                graph.replay()
            done = (K // RING) * RING
        for s in range(done, K):
            step_dev(s)

    if world > 1:
        dist.barrier()
    sampler.active = True
    ms = timed_events(torch, timed_region)
    sampler.active = False
    direct = epp_b200.kernel_launch_count() - launches1
    gpu_launches = direct + (captured // RING) * ((K // RING) * RING if graph is not None else 0)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = world * BATCH * K / (ms_total * 1e-3)
    pop, xor = epp_b200.mask_summary_dev(gathered[0][rank])
    exchange_ok = None
    if world > 1:
        # what the last two steps left in the gathered buffers vs an NCCL all-gather of plain local checks
        exchange_ok = True
        for s_ in (K - 2, K - 1):
            slot = ring[s_ % RING]
            ref_local = (gw.check_segments_dev(slot[0], slot[1]) if segs else gw.check_states_dev(slot[0]))
            ref_all = torch.empty((world, words), dtype=torch.int32, device=dev)
            dist.all_gather_into_tensor(ref_all.view(-1), ref_local)
            exchange_ok = exchange_ok and bool(torch.equal(ref_all, gathered[s_ & 1]))
        if gather is not None:
            calls, timeouts = gather.status()
            exchange_ok = exchange_ok and timeouts == 0
        flag = torch.tensor([1 if exchange_ok else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        exchange_ok = bool(flag.item())
        if not exchange_ok and not args.no_verify:
            raise SystemExit("bench: gathered masks differ from the NCCL all-gather of the local masks")

    # ---- e2e: host buffers through the C ABI, copies inside the timed region ------------------
    e2e_steps = max(8, min(K, 64))
    h_ring = [tuple(t_.cpu().pin_memory() for t_ in slot) for slot in ring]
    h_mask = torch.zeros(words, dtype=torch.int32).pin_memory()

    def step_host(s):
        slot = h_ring[s % RING]
        if segs:
            rc = L.epp_check_segments(gw.handle, slot[0].data_ptr(), slot[1].data_ptr(), BATCH, 0, h_mask.data_ptr())
        else:
            rc = L.epp_check_states(gw.handle, slot[0].data_ptr(), BATCH, 0, h_mask.data_ptr())
        if rc:
            raise RuntimeError(L.epp_last_error().decode())

    for s in range(3):
        step_host(s)
    e2e_pop = int(np.unpackbits(h_mask.numpy().view(np.uint8)).sum())
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler.active = True
    t0 = time.perf_counter()
    for s in range(e2e_steps):
        step_host(s)   # synchronous: returns after the mask is back in host memory
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    sampler.active = False
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * BATCH * e2e_steps / float(t.item())
    in_bytes = BATCH * (48 if segs else 24)

    extras = None
    if rank == 0 and world == 1 and not args.no_extras:
        try:
            extras = run_extras(torch, epp_b200, dev)
        except Exception as ex:  # extras never take the headline down
            extras = {"error": repr(ex)}
    sampler.stop()

    peak, peak_src = measured_peaks()
    kernel_ms = ms_total / K
    achieved = unit_bytes * BATCH / (kernel_ms * 1e-3) / 1e9
    line = {
        "metric": "OBB state+motion validity checks/sec",
        "value": value, "unit": "checks/s", "n_gpus": world, "steps": K, "warmup": max(W, 3),
        "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload], "units_per_step_per_gpu": BATCH,
                   "world_obbs": gw.size(), "broadphase": epp_b200.BP_NAMES[gw.active_broadphase()],
                   "l2": "inputs rotate over a ring of %d distinct batches (%d MiB) > 126 MB L2" %
                         (RING, RING * in_bytes >> 20),
                   "launch": "cuda graph of one ring pass" if graph is not None else "direct launches",
                   "sharding": "contiguous index range per rank; every rank ends each step with all ranks' mask words"
                               if world > 1 else "single GPU",
                   "exchange": exchange, "exchange_verified": exchange_ok, "tuning": args.tune,
                   "mask_popcount_slot0": pop, "mask_xor_slot0": "%016x" % xor},
        "e2e": {"value": e2e_value, "unit": "checks/s", "h2d_bytes_per_step": in_bytes,
                "d2h_bytes_per_step": words * 4, "steps": e2e_steps,
                "api": "epp_check_segments" if segs else "epp_check_states",
                "mask_popcount": e2e_pop},
        "gpu_launches": gpu_launches,
        "clocks": sampler.summary(),
        "roofline": {"bound": "hbm", "kernel": "k_segments" if segs else "k_states_warp",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": load_traffic("k_segments" if segs else "k_states_warp"),
                     "peak_source": peak_src, "bytes_per_launch": unit_bytes * BATCH,
                     "launch_ms": kernel_ms},
    }
    if extras is not None:
        line["extras"] = extras
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args.workload, budget_s=8.0)
    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_b200_config5(args):
    """BASELINE.json configs[4]: one step = the whole 1 B-state + 100 M-edge job against the 10 000-OBB
    world, STRONG scaling: rank r owns the contiguous index range epp_shard_range(n, r, G, 1024) of the
    states and of the edges (inputs generated on its device from the counter-based generator, resident
    before the timed region), writes its mask words into row r of the gathered buffers and the ranks
    exchange them with one in-place NCCL all-gather each."""
    import torch
    import torch.distributed as dist
    import epp_b200
    from epp_b200 import sharding

    rank, world, local_rank = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = epp_b200.lib()
    config, g, o, lo, hi, seed = make_world_inputs("config5")
    gw = epp_b200.World(config)
    gw.parse_gates_and_obstacles(g, o)
    gw.commit()

    NS, NE = args.c5_states, args.c5_edges
    sb, se = epp_b200.shard_range(NS, rank, world, 1024)
    eb, ee = epp_b200.shard_range(NE, rank, world, 1024)
    d_xyz = epp_b200.gen_states_dev(seed, sb, se - sb, lo, hi, device=dev)
    d_a, d_b = epp_b200.gen_segments_dev(seed, eb, ee - eb, lo, hi, 1.0, device=dev)
    sw_s, sw_e = sharding.slot_words(NS, world), sharding.slot_words(NE, world)
    ws, we = (se - sb + 31) // 32, (ee - eb + 31) // 32
    gat_s = gat_e = None
    exchange = "none (single GPU)"
    if world > 1 and args.exchange == "fused":
        try:
            gat_s = epp_b200.Gather(rank, world, sw_s, n_buffers=1).connect()
            gat_e = epp_b200.Gather(rank, world, sw_e, n_buffers=1).connect()
            gs, ge = gat_s.tensor(0, dev), gat_e.tensor(0, dev)
            exchange = ("fused: states - peer stores from the check kernel; edges - push kernel behind "
                        "k_segments; signal/wait on the device, no collective launch")
        except epp_b200.EppError as ex:
            gat_s = gat_e = None
            exchange = "nccl all-gather (fused exchange unavailable: %s)" % ex
    if gat_s is None:
        gs = torch.zeros((world, sw_s), dtype=torch.int32, device=dev)
        ge = torch.zeros((world, sw_e), dtype=torch.int32, device=dev)
        if world > 1 and not exchange.startswith("nccl"):
            exchange = "nccl all-gather, in place"
    torch.cuda.synchronize()

    def step():
        if gat_s is not None:
            gat_s.check_states(gw, d_xyz, 0)
            gat_e.check_segments(gw, d_a, d_b, 0)
            return
        if se > sb:
            gw.check_states_dev(d_xyz, mask=gs[rank][:ws])
        if world > 1:
            dist.all_gather_into_tensor(gs.view(-1), gs[rank])
        if ee > eb:
            gw.check_segments_dev(d_a, d_b, mask=ge[rank][:we])
        if world > 1:
            dist.all_gather_into_tensor(ge.view(-1), ge[rank])

    K, W = args.steps, max(args.warmup, 3)
    sampler = ClockSampler(local_rank)
    sampler.start()
    for _ in range(W):
        step()
    torch.cuda.synchronize()
    l0 = epp_b200.kernel_launch_count()
    if world > 1:
        dist.barrier()
    sampler.active = True
    ms = timed_events(torch, lambda: [step() for _ in range(K)])
    sampler.active = False
    gpu_launches = epp_b200.kernel_launch_count() - l0
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = (NS + NE) * K / (ms_total * 1e-3)

    # states-only kernel time on this rank (dominant kernel of the step) for the roofline
    timeouts = 0
    if gat_s is not None:
        timeouts = gat_s.status()[1] + gat_e.status()[1]
        if timeouts:
            raise SystemExit("bench: %d gather waits timed out" % timeouts)
    tmp_s = torch.empty(max(ws, 1), dtype=torch.int32, device=dev)
    tmp_e = torch.empty(max(we, 1), dtype=torch.int32, device=dev)
    ms_states = timed_events(torch, lambda: [gw.check_states_dev(d_xyz, mask=tmp_s) for _ in range(5)]) / 5
    ms_edges = timed_events(torch, lambda: [gw.check_segments_dev(d_a, d_b, mask=tmp_e) for _ in range(5)]) / 5
    if not (torch.equal(tmp_s[:ws], gs[rank][:ws]) and torch.equal(tmp_e[:we], ge[rank][:we])):
        raise SystemExit("bench: gathered slot of this rank differs from a plain local check")

    # Checksums of the WHOLE gathered masks: must not depend on the number of ranks.
    full_s = torch.from_numpy(sharding.assemble(gs, NS).view(np.int32)).to(dev)
    full_e = torch.from_numpy(sharding.assemble(ge, NE).view(np.int32)).to(dev)
    sum_s, sum_e = epp_b200.mask_summary_dev(full_s), epp_b200.mask_summary_dev(full_e)

    # e2e: host buffers through the C ABI on a bounded slice of the rank's range (the full 24 GB of
    # states would not fit a pinned host allocation on every box); copies inside the timed region.
    n_s, n_e = min(se - sb, 1 << 24), min(ee - eb, 1 << 22) // 32 * 32
    n_s = n_s // 32 * 32
    h_xyz = d_xyz[:n_s].cpu().pin_memory()
    h_a, h_b = d_a[:n_e].cpu().pin_memory(), d_b[:n_e].cpu().pin_memory()
    h_ms = torch.zeros(max(n_s // 32, 1), dtype=torch.int32).pin_memory()
    h_me = torch.zeros(max(n_e // 32, 1), dtype=torch.int32).pin_memory()

    def step_host():
        rc = L.epp_check_states(gw.handle, h_xyz.data_ptr(), n_s, 0, h_ms.data_ptr())
        rc = rc or L.epp_check_segments(gw.handle, h_a.data_ptr(), h_b.data_ptr(), n_e, 0, h_me.data_ptr())
        if rc:
            raise RuntimeError(L.epp_last_error().decode())

    step_host()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler.active = True
    t0 = time.perf_counter()
    for _ in range(3):
        step_host()
    e2e_s = time.perf_counter() - t0
    sampler.active = False
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * (n_s + n_e) * 3 / float(t.item())
    assert np.array_equal(h_ms.numpy()[: n_s // 32], gs[rank][: n_s // 32].cpu().numpy())
    assert np.array_equal(h_me.numpy()[: n_e // 32], ge[rank][: n_e // 32].cpu().numpy())
    sampler.stop()

    peak, peak_src = measured_peaks()
    achieved = BYTES_PER_STATE * (se - sb) / (ms_states * 1e-3) / 1e9
    line = {
        "metric": "OBB state+motion validity checks/sec",
        "value": value, "unit": "checks/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOADS["config5"], "states": NS, "edges": NE,
                   "world_obbs": gw.size(), "broadphase": epp_b200.BP_NAMES[gw.active_broadphase()],
                   "l2": "per-rank inputs %.1f GB >> 126 MB L2" % ((24 * (se - sb) + 48 * (ee - eb)) / 1e9),
                   "sharding": "contiguous index range per rank (align 1024); every rank ends each step with "
                               "all ranks' mask words" if world > 1 else "single GPU",
                   "exchange": exchange,
                   "states_mask_popcount": sum_s[0], "states_mask_xor": "%016x" % sum_s[1],
                   "edges_mask_popcount": sum_e[0], "edges_mask_xor": "%016x" % sum_e[1],
                   "rank0_ms_states_kernel": ms_states, "rank0_ms_edges_kernel": ms_edges},
        "e2e": {"value": e2e_value, "unit": "checks/s", "h2d_bytes_per_step": 24 * n_s + 96 * n_e,
                "d2h_bytes_per_step": (n_s + n_e) // 8, "steps": 3,
                "api": "epp_check_states + epp_check_segments",
                "sample": "first %d states + %d edges of each rank's range per step" % (n_s, n_e)},
        "gpu_launches": gpu_launches,
        "clocks": sampler.summary(),
        "roofline": {"bound": "hbm", "kernel": "k_states_warp<grid>", "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                     "bytes_per_launch": BYTES_PER_STATE * (se - sb), "launch_ms": ms_states},
    }
    if rank == 0 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_parity_config5(gw, gs, ge, sb, eb, lo, hi, seed)
    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def cpu_parity_config5(gw, gs, ge, sb, eb, lo, hi, seed, n_s=1 << 20, n_e=1 << 17):
    """CPU arm on a subsample with the SAME global indices as the head of rank 0's range, timed, and
    compared bit for bit with the GPU masks (mismatch count reported, target 0)."""
    import oracle_py
    threads = os.cpu_count() or 1
    kind, w, _ = cpu_world("config5")
    n_s = min(n_s, gs.shape[1] * 32) // 32 * 32
    n_e = min(n_e, ge.shape[1] * 32) // 32 * 32
    xyz = oracle_py.gen_states(seed, sb, n_s, lo, hi)
    a, b = oracle_py.gen_segments(seed, eb, n_e, lo, hi, 1.0)
    t0 = time.perf_counter()
    ms = w.check_states(xyz, False, threads=threads)
    me = w.check_segments(a, b, False, threads=threads)
    el = time.perf_counter() - t0
    got_s = gs[0][: n_s // 32].cpu().numpy().view(np.uint32)
    got_e = ge[0][: n_e // 32].cpu().numpy().view(np.uint32)
    bad = int(np.unpackbits((got_s ^ ms).view(np.uint8)).sum() + np.unpackbits((got_e ^ me).view(np.uint8)).sum())
    return {"value": (n_s + n_e) / el, "unit": "checks/s", "cores": threads, "kind": kind,
            "sample": "%d states + %d edges with the global indices of rank 0's range head, one pass (%.1f s); "
                      "CPU broad phase = linear scan of the 10 000 stored AABBs (the R-tree stand-in is a list)"
                      % (n_s, n_e, el),
            "mask_mismatches_vs_gpu": bad}


def load_traffic(kernel):
    """dram bytes per launch from the committed ncu capture (profiles/traffic.json), else null."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(path) as f:
            return json.load(f).get(kernel)
    except Exception:
        return None


def run_extras(torch, epp_b200, dev):
    """Kernel-only throughput of the other BASELINE configs (device-resident inputs, CUDA events).
    Informational: the headline stays the workload named in `config`."""
    from epp_b200 import worlds
    config = epp_b200.load_config()
    out = {}

    def rate(fn, units, reps=20):
        for _ in range(3):
            fn()
        # best of three timed blocks: these are short (milliseconds) and one host hiccup halves a single block
        ms = min(timed_events(torch, lambda: [fn() for _ in range(reps)]) for _ in range(3))
        return units * reps / (ms * 1e-3)

    # config 3: 1M segments vs the 256-OBB world (ring of 4 x 48 MiB > L2)
    g, o, lo, hi = worlds.world_256()
    w3 = epp_b200.World(config)
    w3.parse_gates_and_obstacles(g, o)
    segs = [epp_b200.gen_segments_dev(worlds.SEED_SEGMENTS, s * BATCH, BATCH, lo, hi, 1.0, device=dev) for s in range(4)]
    m = torch.empty(BATCH // 32, dtype=torch.int32, device=dev)
    out["config3_segments_256obb_checks_per_s"] = rate(
        lambda: [w3.check_segments_dev(a, b, mask=m) for a, b in segs], 4 * BATCH, reps=5)
    pts = [epp_b200.gen_states_dev(worlds.SEED_SEGMENTS, s * BATCH, BATCH, lo, hi, device=dev) for s in range(8)]
    out["states_256obb_checks_per_s"] = rate(lambda: [w3.check_states_dev(p, mask=m) for p in pts], 8 * BATCH, reps=5)
    del segs, pts

    # config 4: 10k trajectories x ~1k samples, fused and pre-sampled forms
    n_traj = 10000
    coeffs, T, dt = worlds.rest_to_rest_trajectories(n_traj, 8, lo, hi, n_samples=1000, max_hop=3.0)
    d_c, d_T, d_dt = (torch.from_numpy(x).to(dev) for x in (coeffs, T, dt))
    valid, smask, ns, first = w3.check_poly_trajectories_dev(d_c, d_T, d_dt, 0.1, max_samples=1024)
    torch.cuda.synchronize()
    samples = int(ns.sum().item())
    out["config4_samples"] = samples
    out["config4_fused_samples_per_s"] = rate(
        lambda: w3.check_poly_trajectories_dev(d_c, d_T, d_dt, 0.1, max_samples=1024), samples, reps=5)
    rows = 1000
    mats = torch.zeros((n_traj, 10, rows), dtype=torch.float64, device=dev)  # column-major rows x 10 each
    xyz = epp_b200.gen_states_dev(worlds.SEED_TRAJ, 0, n_traj * rows, lo, hi, device=dev).view(n_traj, rows, 3)
    for c in range(3):
        mats[:, 3 * c, :] = xyz[:, :, c]
    out["config4_presampled_samples_per_s"] = rate(
        lambda: w3.check_trajectories_dev(mats, n_traj, rows, rows, 10 * rows, 10, 0.1), n_traj * rows, reps=5)
    del mats, xyz

    # config 5 (single-GPU share): 10k-OBB world, grid broad phase, inputs generated on device
    g, o, lo, hi = worlds.world_10k()
    w5 = epp_b200.World(config)
    w5.parse_gates_and_obstacles(g, o)
    n5 = 1 << 24
    pts = epp_b200.gen_states_dev(worlds.SEED_CLUTTER, 0, n5, lo, hi, device=dev)
    m5 = torch.empty(n5 // 32, dtype=torch.int32, device=dev)
    out["config5_states_10k_obb_checks_per_s"] = rate(lambda: w5.check_states_dev(pts, mask=m5), n5, reps=5)
    a, b = epp_b200.gen_segments_dev(worlds.SEED_CLUTTER, 0, n5 // 4, lo, hi, 1.0, device=dev)
    out["config5_segments_10k_obb_checks_per_s"] = rate(
        lambda: w5.check_segments_dev(a, b, mask=m5), n5 // 4, reps=5)
    out["config5_broadphase"] = epp_b200.BP_NAMES[w5.active_broadphase()]

    # batched min-snap solve, then generate -> sample + check entirely on the device
    g, o, lo, hi = worlds.track_world(config)
    rng = np.random.default_rng(4)
    wp = rng.uniform(lo, hi, size=(n_traj, 9, 3))
    T = epp_b200.segment_times(wp, 1.0, 2.0)
    d_wp, d_T8 = torch.from_numpy(wp).to(dev), torch.from_numpy(T).to(dev)
    d_c8 = epp_b200.minsnap_dev(d_wp, d_T8)
    out["minsnap_8seg_trajectories_per_s"] = rate(lambda: epp_b200.minsnap_dev(d_wp, d_T8, out=d_c8), n_traj, reps=5)
    d_dt8 = torch.from_numpy(T.sum(axis=1) / 999.0).to(dev)
    w1 = epp_b200.World(config)
    w1.parse_gates_and_obstacles(g, o)

    def gen_and_check():
        epp_b200.minsnap_dev(d_wp, d_T8, out=d_c8)
        return w1.check_poly_trajectories_dev(d_c8, d_T8, d_dt8, 0.1, max_samples=1024)
    ns8 = int(gen_and_check()[2].sum().item())
    out["config4_minsnap_generate_and_sweep_samples_per_s"] = rate(gen_and_check, ns8, reps=5)

    # config 1: the reference's Python surface end to end on the track (plan through 4 gates, min-snap, sample)
    import online_traj_planner as otp
    tr = config["track"]
    import tempfile
    otp.OnlineTrajGenerator.path_segments_folder = os.path.join(tempfile.mkdtemp(prefix="epp_bench_"), "path_segments")
    gen = otp.OnlineTrajGenerator(tr["start"], tr["goal"], tr["gates"], tr["obstacles"], config)
    gen.pre_compute_traj(0.0)  # warm-up (allocations, first launches)
    t0 = time.perf_counter()
    gen.pre_compute_traj(0.0)
    out["config1_pre_compute_traj_ms"] = (time.perf_counter() - t0) * 1e3
    out["config1_traj_rows"] = int(gen.get_planned_traj().shape[0])

    # single-state latency through the OMPL-shaped entry point (epp_is_valid)
    t0 = time.perf_counter()
    for i in range(200):
        w3.checkPointValidity([0.1 * i, 0.0, 1.0], False)
    out["single_is_valid_latency_us"] = (time.perf_counter() - t0) / 200 * 1e6
    return out


# ------------------------------------------------------------------------------ CPU arms

def cpu_world(workload):
    """(kind, world object with check_states/check_segments(threads=...)) for the CPU arms:
    oracle/_ref (the reference's own sources) when it was built, else the oracle port."""
    import oracle_py
    config, g, o, lo, hi, seed = make_world_inputs(workload)
    kind = "port"
    w = None
    try:
        import ref_py
        if ref_py.available():
            w = ref_py.RefWorld(config)
            kind = "reference"
    except ImportError:
        pass
    if w is None:
        w = oracle_py.OracleWorld(config)
    w.parse_gates_and_obstacles(g, o)
    return kind, w, (lo, hi, seed)


def cpu_inputs(workload, n, lo, hi, seed, first=0):
    import oracle_py
    if workload == "segments":
        return oracle_py.gen_segments(seed, first, n, lo, hi, 1.0)
    return (oracle_py.gen_states(seed, first, n, lo, hi),)


def cpu_step(w, workload, inputs, threads):
    if workload == "segments":
        return w.check_segments(inputs[0], inputs[1], False, threads=threads)
    return w.check_states(inputs[0], False, threads=threads)


def cpu_baseline(workload, budget_s=8.0):
    threads = os.cpu_count() or 1
    kind, w, (lo, hi, seed) = cpu_world(workload)
    inputs = cpu_inputs(workload, BATCH, lo, hi, seed)
    cpu_step(w, workload, inputs, threads)  # warm-up
    reps, t0 = 0, time.perf_counter()
    while True:
        cpu_step(w, workload, inputs, threads)
        reps += 1
        el = time.perf_counter() - t0
        if el > budget_s or reps >= 200:
            break
    t1 = time.perf_counter()
    cpu_step(w, workload, tuple(x[: BATCH // 8] for x in inputs), 1)
    single = (BATCH // 8) / (time.perf_counter() - t1)
    return {"value": BATCH * reps / el, "unit": "checks/s", "cores": threads, "kind": kind,
            "sample": "%d passes over the step's own 1M-unit batch (%.1f s), %d threads over contiguous ranges"
                      % (reps, el, threads),
            "single_thread_value": single}


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    K, W = args.steps, max(args.warmup, 1)
    threads = os.cpu_count() or 1
    kind, w, (lo, hi, seed) = cpu_world(args.workload)
    full = cpu_inputs(args.workload, BATCH, lo, hi, seed)
    t0 = time.perf_counter()
    cpu_step(w, args.workload, full, threads)
    est = time.perf_counter() - t0
    # Bound the run to ~100 s: a step is the first `n` units of the batch.
    n = BATCH
    while n > (1 << 15) and est * (n / BATCH) * (K + W) > 100.0:
        n //= 2
    if est * (n / BATCH) * (K + W) > 100.0:
        K = max(1, int(100.0 / (est * n / BATCH)) - W)
    inputs = tuple(x[:n] for x in full)
    for _ in range(W):
        cpu_step(w, args.workload, inputs, threads)
    t0 = time.perf_counter()
    for _ in range(K):
        cpu_step(w, args.workload, inputs, threads)
    el = time.perf_counter() - t0
    value = n * K / el
    segs = args.workload == "segments"
    line = {
        "impl": "reference", "metric": "OBB state+motion validity checks/sec", "value": value,
        "unit": "checks/s", "n_gpus": args.gpus, "steps": K, "warmup": W, "ms_per_step": el / K * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload], "units_per_step": n,
                   "note": "host CPU only; runs on rank 0 regardless of --gpus"},
        "cpu_baseline": {"value": value, "unit": "checks/s", "cores": threads, "kind": kind,
                         "sample": "first %d units of the 1M-unit batch per step, %d steps" % (n, K)},
        "e2e": {"value": value, "unit": "checks/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


_REAL_STDOUT = None


def emit(line):
    """The ONE JSON line goes to the real stdout; everything else any library prints (NCCL's version
    banner goes to stdout) was redirected to stderr by main()."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is not None:
        os.write(_REAL_STDOUT, data)
    else:
        sys.stdout.write(data.decode())
        sys.stdout.flush()


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "config5":
        run_b200_config5(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
