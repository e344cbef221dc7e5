#!/usr/bin/env python
"""profiles/ncu_traffic*.json (tools/ncu_traffic.py) -> the per-kernel Markdown table of profiles/ncu_summary_*.md.
usage: ncu_table.py <ncu_traffic.json> [hbm peak GB/s]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORDER = ["k_transform", "k_morton_plan", "k_morton", "k_histogram", "k_scan_histograms", "k_onesweep", "k_tree", "k_tree_top",
         "k_origin_chain", "k_tile_entries", "k_trace"]


def main():
    d = json.load(open(sys.argv[1]))
    peak = float(sys.argv[2]) if len(sys.argv) > 2 else json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    ks = d["kernels"]
    total = sum(k["duration_ns"] * k["launches_averaged"] for k in ks.values())
    print(f"source hash {d['source_hash']}; HBM peak {peak:.1f} GB/s; sum of kernel durations under ncu {total / 1e6:.3f} ms\n")
    print("| kernel | launches | ms (all launches) | share | DRAM MB | DRAM GB/s | of HBM peak | issue slots busy | live lanes / 32 | "
          "warps active | L1 data pipe | L1 / L2 hit | regs |")
    print("|---|---|---|---|---|---|---|---|---|---|---|---|---|")
    for name in ORDER + sorted(set(ks) - set(ORDER)):
        if name not in ks:
            continue
        k = ks[name]
        n = k["launches_averaged"]
        ms = k["duration_ns"] * n / 1e6
        mb = k.get("dram_bytes", 0.0) * n / 1e6
        gbs = k.get("dram_bytes", 0.0) / k["duration_ns"]
        print(f"| `{name}` | {n} | {ms:.3f} | {100 * ms * 1e6 / total:.1f} % | {mb:.0f} | {gbs:.0f} | {gbs / peak:.2f} | "
              f"{k.get('issue_slots_busy_pct', 0):.0f} % | {k.get('warp_exec_eff_threads', 0):.1f} | {k.get('warps_active_pct', 0):.0f} % | "
              f"{k.get('l1_data_pipe_pct', 0):.0f} % | {k.get('l1_hit_pct', 0):.0f} / {k.get('l2_hit_pct', 0):.0f} % | {k.get('registers', 0):.0f} |")


if __name__ == "__main__":
    main()
