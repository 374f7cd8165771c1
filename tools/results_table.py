#!/usr/bin/env python
"""Markdown rows of BASELINE.md §5 / DESIGN.md §7 from bench lines:

    python tools/results_table.py profiles/r02_x_bench.json [n2.json n4.json n8.json]

The first file is a single-GPU `bench.py` line (with extras); the optional others are `bench.py --gpus N` lines whose
headline (config 2, weak scaling) and extras.config5 (strong scaling) fill the multi-GPU columns."""
import json
import sys


def load(path):
    lines = [x for x in open(path).read().splitlines() if x.startswith("{")]
    return json.loads(lines[-1])


def e(x):
    return "—" if x is None else "%.3g" % x


def main():
    one = load(sys.argv[1])
    multi = {}
    for p in sys.argv[2:]:
        d = load(p)
        multi[d["n_gpus"]] = d
    ex = one["extras"]
    r = ex["rooflines"]
    cb = one.get("cpu_baseline", {})
    c5 = ex.get("config5", {})
    c5cpu = c5.get("cpu_baseline", {})
    fp64 = ex["fp64_nofma_instr_per_s"]

    def m2(n):
        return e(multi[n]["value"]) if n in multi else "—"

    def m5(n, key):
        if n not in multi or "extras" not in multi[n] or "config5" not in multi[n]["extras"]:
            return "—"
        c = multi[n]["extras"]["config5"]
        units = c["config"]["states"] if key == "states" else c["config"]["edges"]
        ms = c["config"]["rank0_ms_states_kernel" if key == "states" else "rank0_ms_edges_kernel"]
        return "%s (rank 0 kernel %.2f ms)" % (e(units / n / (ms * 1e-3) * n), ms)

    print("FP64 non-fused roof measured: %.3e instr/s; HBM roof %.1f GB/s" % (fp64, one["roofline"]["peak"]))
    print("| Config | CPU 1 thr (checks/s) | CPU all cores (N = %s) | 1×B200 | 2× | 4× | 8× | %% HBM roof | %% FP64 roof | mask mismatches |" % cb.get("cores"))
    print("|---|---|---|---|---|---|---|---|---|---|")
    mm = one.get("mask_mismatches", {})
    print("| 2 (1 M states / step / GPU, weak) | %s | %s | %s | %s | %s | %s | %.1f | %.1f | %s |" % (
        e(cb.get("single_thread_value")), e(cb.get("value")), e(one["value"]), m2(2), m2(4), m2(8),
        100 * one["roofline"]["frac"], 100 * one["value"] * r["config2_states_16M_per_launch"]["fp64_ops_per_unit"] / fp64,
        mm.get("total")))
    for name, key in (("3", "config3_segments_256obb"), ("4 pre-sampled", "config4_presampled"), ("4 fused", "config4_fused")):
        x = r[key]
        cpu = x.get("cpu", {})
        print("| %s | %s | %s | %s | — | — | — | %.1f | %.1f | %s |" % (
            name, e(cpu.get("single_thread")), e(cpu.get("all_cores")), e(x["units_per_s"]), 100 * x["frac"], 100 * x["fp64_frac"],
            x.get("mask_mismatches_vs_cpu")))
    if c5:
        rs, re_ = c5["roofline"], c5["roofline_edges"]
        print("| 5 states (1 B, strong) | %s | %s | %s | %s | %s | %s | %.1f | %.1f | %s |" % (
            e(c5cpu.get("states_single_thread")), e(c5cpu.get("states_all_cores")), e(rs["units_per_s"]),
            m5(2, "states"), m5(4, "states"), m5(8, "states"), 100 * rs["frac"], 100 * rs["fp64_frac"], c5cpu.get("mask_mismatches_vs_gpu")))
        print("| 5 edges (100 M, strong) | %s | %s | %s | %s | %s | %s | %.1f | %.1f | %s |" % (
            e(c5cpu.get("edges_single_thread")), e(c5cpu.get("edges_all_cores")), e(re_["units_per_s"]),
            m5(2, "edges"), m5(4, "edges"), m5(8, "edges"), 100 * re_["frac"], 100 * re_["fp64_frac"], c5cpu.get("mask_mismatches_vs_gpu")))
        print("config 5 whole step (states + edges + exchange), ms: 1 GPU %.3f%s" % (
            c5["ms_per_step"], "".join(", %d GPUs %.3f" % (n, multi[n]["extras"]["config5"]["ms_per_step"])
                                        for n in sorted(multi) if "extras" in multi[n] and "config5" in multi[n]["extras"])))
    print("e2e %.3e checks/s = %.1f GB/s of %.1f GB/s pinned H2D (%.2f)" % (
        one["e2e"]["value"], one["e2e"]["achieved_h2d_gbs_per_gpu"], one["e2e"]["h2d_pinned_gbs"], one["e2e"]["frac_of_h2d"]))
    print("single isValid %.2f us (launch + sync form %.2f us)" % (ex["single_is_valid_latency_us"], ex["single_is_valid_launch_sync_latency_us"]))


if __name__ == "__main__":
    main()
